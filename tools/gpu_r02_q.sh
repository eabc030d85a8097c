#!/bin/bash
# round 2, GPU call Q (1 GPU): L2 prefetch-size qualifier of the 256-bit record loads (.L2::64B / 128B / 256B) and the device's
# L2 fetch-granularity limit: DRAM sectors (ncu metric pass) and time (perf_quick) of the point query and the frames
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
M=gpu__time_duration.sum,dram__sectors_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum,lts__t_requests_srcunit_tex_op_read.sum
ncupass() { tag=$1; shift; env "$@" timeout 600 ncu --metrics $M --clock-control none --profile-from-start off --csv --log-file gpurun_out/q_ncu_$tag.csv python tools/ncu_points.py --workload c3 > gpurun_out/q_ncu_$tag.log 2>&1
  python - gpurun_out/q_ncu_$tag.csv $tag <<'PY'
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]; mi, vi, ii = h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
out = {}
for r in rows[1:]:
    out.setdefault(r[ii], {})[r[mi]] = r[vi]
names = ["random", "no-out", "sorted", "local128", "grad"]
for k, d in out.items():
    print(sys.argv[2], names[int(k)], "dram_rd", d.get("dram__sectors_read.sum"), "ns", d.get("gpu__time_duration.sum"), "l2rd", d.get("lts__t_sectors_srcunit_tex_op_read.sum"), "req", d.get("lts__t_requests_srcunit_tex_op_read.sum"))
PY
}
pq() { tag=$1; shift; env "$@" PQ_CLEAN_FLUSH=1 timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/q_pq_$tag.json 2> gpurun_out/q_pq_$tag.err; python - "$tag" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/q_pq_%s.json" % sys.argv[1])); print(sys.argv[1], {k: d[k] for k in ("opaque", "translucent", "points") if k in d})
except Exception as e:
    print(sys.argv[1], "no result", e)
PY
}
ncupass base
ncupass l2_64 EXA_LIB=owlexabrick_b200/libexabrick_b200.l2_64.so
ncupass fetch32 EXA_L2_FETCH=32
pq base
pq l2_64 EXA_LIB=owlexabrick_b200/libexabrick_b200.l2_64.so
pq l2_128 EXA_LIB=owlexabrick_b200/libexabrick_b200.l2_128.so
pq l2_256 EXA_LIB=owlexabrick_b200/libexabrick_b200.l2_256.so
pq fetch32 EXA_L2_FETCH=32
pq l2_64_fetch32 EXA_LIB=owlexabrick_b200/libexabrick_b200.l2_64.so EXA_L2_FETCH=32
