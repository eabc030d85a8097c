#!/bin/bash
# round 2, GPU call B (1 GPU): full parity suite (incl. the GPU builder), kernel-variant A/B on C3, GPU builder on C3,
# ncu --set full of the frame and points kernels (summaries exported on the box; the .ncu-rep only if small).
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
echo "== pytest -m gpu"
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/b_pytest.log
echo "== GPU builder on c3 (+ scene cache for the steps below)"
timeout 600 python - > gpurun_out/b_builder_c3.json 2> gpurun_out/b_builder_c3.err <<'EOF'
import json, sys, time
sys.path.insert(0, ".")
import numpy as np
import bench
from owlexabrick_b200 import renderer as R
bricks, field = bench.load_or_build_scene("c3", R, gpu_builder=True)
out = dict(bench.BUILDER_INFO)
cells, _, _, _ = bench.make_workload("c3", with_field=False)
t = time.time(); hb = R.ExaBricks.build(cells); out["host_builder_s"] = round(time.time() - t, 2)
out["gpu_equals_host"] = bool(hb.headers.tobytes() == bricks.headers.tobytes() and np.array_equal(hb.cell_ids, bricks.cell_ids) and hb.kd_nodes.tobytes() == bricks.kd_nodes.tobytes())
t = time.time(); b2 = R.ExaBricks.build_gpu(cells); out["second_gpu_build"] = b2.timings
print(json.dumps(out))
EOF
echo "builder rc=$?"; cat gpurun_out/b_builder_c3.json; tail -3 gpurun_out/b_builder_c3.err
echo "== perf_quick variants on c3"
run() { tag=$1; shift; env "$@" timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/b_pq_$tag.json 2> gpurun_out/b_pq_$tag.err; echo "$tag rc=$?"; python - "$tag" <<'EOF'
import json, sys
try:
    d = json.load(open("gpurun_out/b_pq_%s.json" % sys.argv[1]))
    print({k: d[k] for k in ("opaque", "translucent", "points") if k in d})
except Exception as e:
    print("no result", e)
EOF
}
run base PQ_CLEAN_FLUSH=1
run mb6 EXA_LIB=owlexabrick_b200/libexabrick_b200.mb6.so PQ_CLEAN_FLUSH=1
run mb8 EXA_LIB=owlexabrick_b200/libexabrick_b200.mb8.so PQ_CLEAN_FLUSH=1
run ppt2 EXA_LIB=owlexabrick_b200/libexabrick_b200.ppt2.so PQ_CLEAN_FLUSH=1
run ppt4 EXA_LIB=owlexabrick_b200/libexabrick_b200.ppt4.so PQ_CLEAN_FLUSH=1
echo "== ncu --set full (c3: opaque frame, translucent frame, 1 M points)"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/b_ncu_c3 python tools/ncu_target.py --workload c3 > gpurun_out/b_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/b_ncu.log
if [ -f gpurun_out/b_ncu_c3.ncu-rep ]; then
  ncu -i gpurun_out/b_ncu_c3.ncu-rep --page details > gpurun_out/b_ncu_c3_details.txt 2>&1
  ncu -i gpurun_out/b_ncu_c3.ncu-rep --page raw --csv > gpurun_out/b_ncu_c3_raw.csv 2>&1
  ncu -i gpurun_out/b_ncu_c3.ncu-rep --page source --csv > gpurun_out/b_ncu_c3_source.csv 2>&1
  ls -la gpurun_out/b_ncu_c3*
  sz=$(stat -c %s gpurun_out/b_ncu_c3.ncu-rep); if [ "$sz" -gt 30000000 ]; then rm gpurun_out/b_ncu_c3.ncu-rep; fi
  gzip -f gpurun_out/b_ncu_c3_source.csv
fi
