#!/bin/bash
# round 2, GPU call P (1 GPU): DRAM-sector accounting of the point query (ncu metric pass over five point sets, two grid sizes)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
M=gpu__time_duration.sum,dram__sectors_read.sum,dram__sectors_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum,lts__t_requests_srcunit_tex_op_read.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld_lookup_miss.sum,lts__t_sectors_srcunit_ltcfabric.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,lts__d_sectors_fill_device.sum
for g in default 4096; do
  if [ $g = default ]; then unset EXA_LOCGRID_CELLS; else export EXA_LOCGRID_CELLS=$g; fi
  timeout 600 ncu --metrics $M --clock-control none --profile-from-start off --csv --log-file gpurun_out/p_points_grid_$g.csv python tools/ncu_points.py --workload c3 > gpurun_out/p_points_grid_$g.log 2>&1; echo "grid $g rc=$?"
  python - gpurun_out/p_points_grid_$g.csv <<'PY'
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]; ki, mi, vi = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"); ii = h.index("ID")
out = {}
for r in rows[1:]:
    out.setdefault(r[ii], {})[r[mi]] = r[vi]
for k, d in out.items():
    print(k, {m.replace("lts__t_sectors_srcunit_tex_op_read", "l2tex_rd").replace("l1tex__t_sectors_pipe_lsu_mem_global_op_ld", "l1_ld").replace(".sum", ""): v for m, v in d.items()})
PY
done
