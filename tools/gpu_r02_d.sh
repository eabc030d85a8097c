#!/bin/bash
# round 2, GPU call D (1 GPU): location-grid size sweep for the points kernel, frame kernel with the scene-rectangle tile skip, bench line
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/d_pytest.log
run() { tag=$1; shift; env "$@" timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/d_pq_$tag.json 2> gpurun_out/d_pq_$tag.err; echo "$tag rc=$?"; python - "$tag" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/d_pq_%s.json" % sys.argv[1]))
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
run base PQ_CLEAN_FLUSH=1
run grid8m EXA_LOCGRID_CELLS=8388608 PQ_CLEAN_FLUSH=1
run grid32m EXA_LOCGRID_CELLS=33554432 PQ_CLEAN_FLUSH=1 PQ_DIAG=1
run grid128m EXA_LOCGRID_CELLS=134217728 PQ_CLEAN_FLUSH=1
run grid512m EXA_LOCGRID_CELLS=536870912 PQ_CLEAN_FLUSH=1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/d_bench_c3.json 2> gpurun_out/d_bench_c3.err; echo "bench rc=$?"; tail -c 600 gpurun_out/d_bench_c3.json; tail -4 gpurun_out/d_bench_c3.err
