#!/bin/bash
# round 2, GPU call T (1 GPU): what the driver runs at round end - smoke(), pytest -m gpu, bench.py (both arms) - plus the flat-layout C3 line
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/t_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/t_smoke.log
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/t_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/t_pytest.log
( time timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/t_bench_reference.json 2> gpurun_out/t_bench_reference.err ) 2>&1 | grep real; echo "reference rc=$?"; tail -c 600 gpurun_out/t_bench_reference.json
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/t_bench_c3.json 2> gpurun_out/t_bench_c3.err ) 2>&1 | grep real; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/t_bench_c3.json") if l.startswith("{")][-1])
print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches", "clocks")}, "e2e", d["e2e"]["value"], d["e2e"]["frames_per_s"], "roofline", {k: d["roofline"][k] for k in ("frac", "dram_frac", "limiter")})
print("recon", {k: d["recon"][k] for k in ("ms", "frac", "dram_frac", "e2e_samples_per_s")}, "parity", d["parity"]["frame"]["rgba_max_abs"], d["parity"]["abr_bit_exact"])
PY
( time timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --flat-layout --no-cpu-baseline --no-builder-timing > gpurun_out/t_bench_c3_flat.json 2> gpurun_out/t_bench_c3_flat.err ) 2>&1 | grep real; echo "flat rc=$?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/t_bench_c3_flat.json") if l.startswith("{")][-1])
print("flat", {k: d[k] for k in ("ms_per_step",)}, d["config"]["scalar_layout"], "translucent", d["dvr_translucent"]["ms_per_frame"], "recon ms", d["recon"]["ms"])
PY
