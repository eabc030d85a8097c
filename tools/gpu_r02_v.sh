#!/bin/bash
# round 2, GPU call V (1 GPU): group rebalance test, config C4 at N = 1 with the parity object (iso + DVR against the oracle at full size)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/v_pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/v_pytest_multi.log
( time timeout 1200 python bench.py --workload c4 --gpus 1 --steps 20 --warmup 5 > gpurun_out/v_bench_c4.json 2> gpurun_out/v_bench_c4.err ) 2>&1 | grep real; echo "bench c4 rc=$?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/v_bench_c4.json") if l.startswith("{")][-1])
print({k: d[k] for k in ("value", "ms_per_step", "frames_per_s")}, "e2e", d["e2e"]["frames_per_s"], "roofline", {k: d["roofline"][k] for k in ("frac", "dram_frac")})
print("parity", json.dumps(d["parity"])[:1500])
print("build", d["build"])
PY
tail -5 gpurun_out/v_bench_c4.err
