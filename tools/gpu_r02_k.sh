#!/bin/bash
# round 2, GPU call K (1 GPU): GPU suite after the sort-last rectangle change + quick frame timing (no regression of the plain frame)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/k_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/k_pytest.log
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
env PQ_CLEAN_FLUSH=1 timeout 300 python tools/perf_quick.py --workload c3 --tag k > gpurun_out/k_pq.json 2> gpurun_out/k_pq.err; python - <<'PY'
import json
d = json.load(open("gpurun_out/k_pq.json")); print({k: d[k] for k in ("opaque", "translucent", "points") if k in d})
PY
