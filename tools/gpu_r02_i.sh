#!/bin/bash
# round 2, GPU call I (N GPUs): e2e diagnostics of the tile pipeline (own copy alone per rank, loop without copies)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi topo -m > gpurun_out/i_topo.txt 2>&1; numactl -H >> gpurun_out/i_topo.txt 2>&1; lscpu | grep -i -E "numa|socket|model name|^CPU\(s\)" >> gpurun_out/i_topo.txt
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/i_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -3 gpurun_out/i_pytest_multi.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus $N --steps 40 --warmup 5 --no-sortlast --no-builder-timing > gpurun_out/i_bench_n${N}_tiles.json 2> gpurun_out/i_bench_n${N}_tiles.err; echo "rc=$?"
grep -E "e2e diagnostics|own frame kernel" gpurun_out/i_bench_n${N}_tiles.err | head -4
python - $N <<'PY'
import json, sys
d = json.loads([l for l in open("gpurun_out/i_bench_n%s_tiles.json" % sys.argv[1]) if l.startswith("{")][-1])
print("ms", d["ms_per_step"], "fps", d["frames_per_s"], "e2e", d["e2e"]["frames_per_s"], d["e2e"].get("diagnostics"))
PY
cat gpurun_out/i_topo.txt | head -30
