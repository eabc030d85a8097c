#!/bin/bash
# round 2, GPU call F (1 GPU): persistent-lane points kernel A/B, host-buffer point query, GPU builder with pinned staging
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/f_pytest.log
run() { tag=$1; shift; env "$@" timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/f_pq_$tag.json 2> gpurun_out/f_pq_$tag.err; echo "$tag rc=$?"; python - "$tag" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/f_pq_%s.json" % sys.argv[1]))
    print({k: d[k] for k in ("opaque", "translucent", "points", "points_locate", "points_sorted_us", "points_local128_us", "points_host") if k in d})
except Exception as e:
    print("no result", e)
PY
}
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
print(bench.BUILDER_INFO)
PY
run queue EXA_POINTS_QUEUE=1 PQ_CLEAN_FLUSH=1 PQ_DIAG=1
run thread EXA_POINTS_QUEUE=0 PQ_CLEAN_FLUSH=1
run queue_c2 EXA_POINTS_QUEUE=1 PQ_CLEAN_FLUSH=1 PQ_WORKLOAD=c2
