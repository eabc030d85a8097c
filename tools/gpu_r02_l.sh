#!/bin/bash
# round 2, GPU call L (N GPUs): sort-last as the main mode on C3 (opaque + translucent), after the per-part rectangles
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --partition sortlast --no-sortlast --no-builder-timing > gpurun_out/l_bench_n${N}_sortlast.json 2> gpurun_out/l_bench_n${N}_sortlast.err; echo "rc=$?"
python - gpurun_out/l_bench_n${N}_sortlast.json <<'PY'
import json, sys
d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
print({k: d[k] for k in ("n_gpus", "ms_per_step", "frames_per_s", "frame_kernel_only_ms_per_step", "multi_gpu_check")}, "e2e fps", d["e2e"]["frames_per_s"])
print("translucent", d.get("dvr_translucent"))
PY
grep -o "own frame kernel[^;]*; per rank \[[^]]*\]" gpurun_out/l_bench_n${N}_sortlast.err | head -1
grep -iE "Traceback|Error" gpurun_out/l_bench_n${N}_sortlast.err | head -5
