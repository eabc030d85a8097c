#!/bin/bash
# round 2, GPU call W (1 GPU): profiler evidence on the final tree - launch list of the bench's timed region, ncu --set full of the
# frame and points kernels (summaries exported on the box), whole GPU suite once more
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/w_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/w_pytest.log
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-builder-timing > gpurun_out/w_bench_plain.json 2> gpurun_out/w_bench_plain.err; echo "plain bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/w_launches_timed_region.csv python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-builder-timing --profile-range > gpurun_out/w_ncu_launches.log 2>&1; echo "launch list rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/w_launches_timed_region.csv")) if len(r) > 10]
h = rows[0]; ki, vi = h.index("Kernel Name"), h.index("Metric Value")
t = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    t[r[ki]][0] += 1; t[r[ki]][1] += float(r[vi])
tot = sum(v[1] for v in t.values())
for k, v in sorted(t.items(), key=lambda kv: -kv[1][1]): print("%6d launches %10.1f us %5.1f %%  %s" % (v[0], v[1] / 1e3, 100 * v[1] / tot, k[:90]))
PY
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/w_ncu_c3 python tools/ncu_target.py --workload c3 > gpurun_out/w_ncu.log 2>&1; echo "ncu rc=$?"
if [ -f gpurun_out/w_ncu_c3.ncu-rep ]; then
  ncu -i gpurun_out/w_ncu_c3.ncu-rep --page details > gpurun_out/w_ncu_c3_details.txt 2>&1
  ncu -i gpurun_out/w_ncu_c3.ncu-rep --page raw --csv > gpurun_out/w_ncu_c3_raw.csv 2>&1
  for i in 0 1 2; do python tools/ncu_lines.py gpurun_out/w_ncu_c3.ncu-rep 40 $i > gpurun_out/w_ncu_c3_lines_$i.txt 2>&1; done
  ls -la gpurun_out/w_ncu_c3*
fi
