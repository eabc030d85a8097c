#!/bin/bash
# round 2, GPU call R (1 GPU): two-level point-location grid: parity suite, point query with / without the sub-grids (C3, C2)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r_pytest.log
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
pq() { tag=$1; shift; env "$@" PQ_CLEAN_FLUSH=1 PQ_DIAG=1 timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/r_pq_$tag.json 2> gpurun_out/r_pq_$tag.err; python - "$tag" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/r_pq_%s.json" % sys.argv[1])); print(sys.argv[1], {k: d[k] for k in ("opaque", "translucent", "points", "points_locate", "points_sorted_us", "points_local128_us", "points_host") if k in d})
except Exception as e:
    print(sys.argv[1], "no result", e)
PY
}
pq sub
pq nosub EXA_NO_LOCSUB=1
pq sub_grid512k EXA_LOCGRID_CELLS=524288
pq sub_grid128k EXA_LOCGRID_CELLS=131072
pq sub_c2 PQ_WORKLOAD=c2
pq nosub_c2 PQ_WORKLOAD=c2 EXA_NO_LOCSUB=1
