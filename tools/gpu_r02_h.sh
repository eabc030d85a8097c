#!/bin/bash
# round 2, GPU call H (1 GPU): the no-alloc point query alone in a fresh process, the whole GPU suite, smaller location grids
# for the point query, ncu --set full of the final frame / points kernels on C3, bench line (C3, N=1).
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/h_pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -5 gpurun_out/h_pytest_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/h_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/h_pytest.log
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
run() { tag=$1; shift; env "$@" timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/h_pq_$tag.json 2> gpurun_out/h_pq_$tag.err; echo "$tag rc=$?"; python - "$tag" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/h_pq_%s.json" % sys.argv[1]))
    print({k: d[k] for k in ("opaque", "translucent", "points", "points_locate", "points_sorted_us", "points_local128_us", "points_host") if k in d})
except Exception as e:
    print("no result", e)
PY
}
run base PQ_CLEAN_FLUSH=1 PQ_DIAG=1
run grid1m EXA_LOCGRID_CELLS=1048576 PQ_CLEAN_FLUSH=1 PQ_DIAG=1
run grid512k EXA_LOCGRID_CELLS=524288 PQ_CLEAN_FLUSH=1 PQ_DIAG=1
run grid128k EXA_LOCGRID_CELLS=131072 PQ_CLEAN_FLUSH=1 PQ_DIAG=1
run grid4m EXA_LOCGRID_CELLS=4194304 PQ_CLEAN_FLUSH=1 PQ_DIAG=1
echo "== ncu --set full (c3: opaque frame, translucent frame, 1 M points)"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/h_ncu_c3 python tools/ncu_target.py --workload c3 > gpurun_out/h_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/h_ncu.log
if [ -f gpurun_out/h_ncu_c3.ncu-rep ]; then
  ncu -i gpurun_out/h_ncu_c3.ncu-rep --page details > gpurun_out/h_ncu_c3_details.txt 2>&1
  ncu -i gpurun_out/h_ncu_c3.ncu-rep --page raw --csv > gpurun_out/h_ncu_c3_raw.csv 2>&1
  ncu -i gpurun_out/h_ncu_c3.ncu-rep --page source --csv > gpurun_out/h_ncu_c3_source.csv 2>&1
  ls -la gpurun_out/h_ncu_c3*
  sz=$(stat -c %s gpurun_out/h_ncu_c3.ncu-rep); if [ "$sz" -gt 30000000 ]; then rm gpurun_out/h_ncu_c3.ncu-rep; fi
  gzip -f gpurun_out/h_ncu_c3_source.csv
fi
echo "== bench c3"
timeout 1200 python bench.py --steps 20 --warmup 5 > gpurun_out/h_bench_c3.json 2> gpurun_out/h_bench_c3.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/h_bench_c3.json; tail -5 gpurun_out/h_bench_c3.err
