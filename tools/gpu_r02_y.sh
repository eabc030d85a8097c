#!/bin/bash
# round 2, GPU call Y (1 GPU): host-buffer point query after the pipeline reorder, chunk sizes; point-query tests
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py tests/test_gpu_parity.py -x -q -m gpu -k "noalloc or reconstruction" > gpurun_out/y_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/y_pytest.log
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
for ch in default; do
  if [ $ch = default ]; then unset EXA_POINTS_CHUNK; else export EXA_POINTS_CHUNK=$ch; fi
  PQ_CLEAN_FLUSH=1 timeout 200 python tools/perf_quick.py --workload c3 --tag y_$ch > gpurun_out/y_pq_$ch.json 2> gpurun_out/y_pq_$ch.err
  python - $ch <<'PY'
import json, sys
d = json.load(open("gpurun_out/y_pq_%s.json" % sys.argv[1])); print("chunk", sys.argv[1], d.get("points_host"), d.get("points_host_fresh_arrays_ms"), d.get("points"))
PY
done
