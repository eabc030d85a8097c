#!/bin/bash
# round 2, GPU call X (1 GPU): the point query with .L2::64B on its 256-bit loads: GPU suite, timings, DRAM sectors (ncu metric pass)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/x_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/x_pytest.log
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
env PQ_CLEAN_FLUSH=1 PQ_DIAG=1 timeout 300 python tools/perf_quick.py --workload c3 --tag x > gpurun_out/x_pq.json 2> gpurun_out/x_pq.err; python - <<'PY'
import json
d = json.load(open("gpurun_out/x_pq.json")); print({k: d[k] for k in ("opaque", "translucent", "points", "points_sorted_us", "points_local128_us", "points_host") if k in d})
PY
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__sectors_read.sum,dram__sectors_write.sum
timeout 300 ncu --metrics $M --clock-control none --profile-from-start off --csv --log-file gpurun_out/x_ncu_points.csv python tools/ncu_points.py --workload c3 > gpurun_out/x_ncu_points.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv
rows = [r for r in csv.reader(open("gpurun_out/x_ncu_points.csv")) if len(r) > 10]
h = rows[0]; mi, vi, ii, ui = h.index("Metric Name"), h.index("Metric Value"), h.index("ID"), h.index("Metric Unit")
out = {}
for r in rows[1:]:
    out.setdefault(r[ii], {})[r[mi]] = r[vi] + " " + r[ui]
for k, d in out.items(): print(["random", "no-out", "sorted", "local128", "grad"][int(k)], d)
PY
