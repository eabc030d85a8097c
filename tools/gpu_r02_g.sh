#!/bin/bash
# round 2, GPU call G (1 GPU): the no-alloc point query alone in a fresh process, then the whole GPU suite; persistent-lane
# points kernel with 1 / 2 / 4 points per lane against the thread-per-point kernel.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/g_pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -15 gpurun_out/g_pytest_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/g_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/g_pytest.log
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
run() { tag=$1; shift; env "$@" timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/g_pq_$tag.json 2> gpurun_out/g_pq_$tag.err; echo "$tag rc=$?"; python - "$tag" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/g_pq_%s.json" % sys.argv[1]))
    print({k: d[k] for k in ("points", "points_locate", "points_sorted_us", "points_local128_us", "points_host") if k in d})
except Exception as e:
    print("no result", e)
PY
}
for v in q1 q2 q4; do run $v EXA_LIB=owlexabrick_b200/libexabrick_b200.$v.so EXA_POINTS_QUEUE=1 PQ_CLEAN_FLUSH=1 PQ_DIAG=1 PQ_POINTS_ONLY=1; done
run thread EXA_POINTS_QUEUE=0 PQ_CLEAN_FLUSH=1 PQ_DIAG=1 PQ_POINTS_ONLY=1
