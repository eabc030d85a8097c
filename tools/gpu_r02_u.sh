#!/bin/bash
# round 2, GPU call U (N GPUs): the driver's scaling line on the final tree (default bench: image tiles + the sort-last sub-object)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
( time timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/u_bench_n${N}.json 2> gpurun_out/u_bench_n${N}.err ) 2>&1 | grep real; echo "rc=$?"
python - gpurun_out/u_bench_n${N}.json <<'PY'
import json, sys
d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
print({k: d[k] for k in ("n_gpus", "value", "ms_per_step", "frames_per_s", "frame_kernel_only_ms_per_step", "barrier_alone_ms", "multi_gpu_check", "gpu_launches")})
print("e2e", {k: d["e2e"][k] for k in ("value", "frames_per_s", "d2h_bytes_per_step_per_rank", "strip_rows", "d2h_gbs_per_rank_all_copying", "diagnostics")})
print("sortlast", {k: d["sortlast"][k] for k in ("resident_cells_per_rank", "opaque", "translucent")})
print("translucent", d.get("dvr_translucent"))
PY
[ -n "$SKIPREF" ] || timeout 300 $TR bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/u_ref_n${N}.json 2> gpurun_out/u_ref_n${N}.err; echo "reference arm under torchrun rc=$?"; tail -c 300 gpurun_out/u_ref_n${N}.json
