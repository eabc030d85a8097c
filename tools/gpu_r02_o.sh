#!/bin/bash
# round 2, GPU call O (1 GPU): does a lower CTA occupancy shorten the N=8 rank's frame (longest-ray floor)?  One GPU plays each rank.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
for k in 0 5 4 3 2; do
  EXA_FRAME_MAX_CTAS=$k timeout 300 python tools/scale_emulate.py --workload c3 --ranks 1,8 --steps 10 > gpurun_out/o_emulate_ctas$k.json 2> gpurun_out/o_emulate_ctas$k.err
  python - $k <<'PY'
import json, sys
for l in open("gpurun_out/o_emulate_ctas%s.json" % sys.argv[1]):
    if l.startswith("{"):
        d = json.loads(l); print("max_ctas", sys.argv[1], d["xf"], "N1", d["N1_ms"], "N8", d["N8_ms"], d["N8_ranks"], "hist8", d["N8_hist"])
PY
done
