#!/bin/bash
# round 2, GPU call C (1 GPU): parity suite, frame-kernel occupancy variants, points-kernel variants (location-grid size, points per thread)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
echo "== pytest -m gpu"
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/c_pytest.log
run() { tag=$1; shift; env "$@" timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/c_pq_$tag.json 2> gpurun_out/c_pq_$tag.err; echo "$tag rc=$?"; python - "$tag" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/c_pq_%s.json" % sys.argv[1]))
    print({k: d[k] for k in ("opaque", "translucent", "points", "points_locate", "points_sorted_us", "points_local128_us") if k in d})
except Exception as e:
    print("no result", e)
PY
}
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
run base PQ_CLEAN_FLUSH=1 PQ_DIAG=1
run mb6 EXA_LIB=owlexabrick_b200/libexabrick_b200.mb6.so PQ_CLEAN_FLUSH=1
run mb8 EXA_LIB=owlexabrick_b200/libexabrick_b200.mb8.so PQ_CLEAN_FLUSH=1
run grid16m EXA_LOCGRID_CELLS=16777216 PQ_CLEAN_FLUSH=1 PQ_DIAG=1
run grid128m EXA_LOCGRID_CELLS=134217728 PQ_CLEAN_FLUSH=1
run ppt2 EXA_LIB=owlexabrick_b200/libexabrick_b200.ppt2.so PQ_CLEAN_FLUSH=1
run ppt4 EXA_LIB=owlexabrick_b200/libexabrick_b200.ppt4.so PQ_CLEAN_FLUSH=1
run ppt2_grid16m EXA_LIB=owlexabrick_b200/libexabrick_b200.ppt2.so EXA_LOCGRID_CELLS=16777216 PQ_CLEAN_FLUSH=1
run ppt4_grid16m EXA_LIB=owlexabrick_b200/libexabrick_b200.ppt4.so EXA_LOCGRID_CELLS=16777216 PQ_CLEAN_FLUSH=1
