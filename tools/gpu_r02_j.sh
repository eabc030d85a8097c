#!/bin/bash
# round 2, GPU call J (N GPUs, meant for 8): group API, the driver's default bench line (tiles + sort-last sub-object),
# sort-last as the main mode, config C4 (iso + DVR)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
N=${1:-8}
nvidia-smi topo -m > gpurun_out/j_topo_n$N.txt 2>&1
echo "== group API on $N devices"
timeout 300 python - $N > gpurun_out/j_group_n$N.log 2>&1 <<'PY'
import sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import scenes
from owlexabrick_b200 import capi, renderer as R
n = int(sys.argv[1])
cells, fields, dom = scenes.small_scene("nb64_off", 1)
b = R.ExaBricks.build(cells, max_leaf_width=8)
W, H = 640, 360
fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H)); fd.opacity_scale = 0.05; fd.ert = 2.0
r1 = R.Renderer(b, fields); fd.apply(r1); r1.render(); ref = r1.read_rgba8()
for mode, name in ((capi.PARTITION_IMAGE_TILES, "tiles"), (capi.PARTITION_SORT_LAST, "sortlast")):
    g = R.RendererGroup(b, fields, list(range(n)), mode)
    g.apply(fd)
    img = g.render(); img = g.render()
    d = np.abs(scenes.unpack_rgba8(img) - scenes.unpack_rgba8(ref)).max()
    t = time.perf_counter()
    for _ in range(50): g.render(img)
    print("group %s on %d devices: max RGBA8 diff vs 1 GPU %d, %.3f ms/frame incl. host read-back" % (name, n, d, (time.perf_counter() - t) / 50 * 1e3), flush=True)
    assert d <= (0 if name == "tiles" else 1)
    g.close()
print("group ok")
PY
echo "group rc=$?"; tail -3 gpurun_out/j_group_n$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
summ() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print({k: d[k] for k in ("n_gpus", "ms_per_step", "frames_per_s", "frame_kernel_only_ms_per_step", "barrier_alone_ms", "multi_gpu_check")}, "e2e fps", d["e2e"]["frames_per_s"])
    if d.get("sortlast"): print("sortlast", {k: d["sortlast"][k] for k in ("resident_cells_per_rank", "opaque", "translucent")})
    if d.get("dvr_translucent"): print("translucent", d["dvr_translucent"])
except Exception as e:
    print("no result", e)
PY
}
echo "== bench --gpus $N (driver's line)"
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/j_bench_n${N}_tiles.json 2> gpurun_out/j_bench_n${N}_tiles.err; echo "rc=$?"; summ gpurun_out/j_bench_n${N}_tiles.json
echo "== bench --gpus $N --partition sortlast"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --partition sortlast --no-sortlast --no-builder-timing > gpurun_out/j_bench_n${N}_sortlast.json 2> gpurun_out/j_bench_n${N}_sortlast.err; echo "rc=$?"; summ gpurun_out/j_bench_n${N}_sortlast.json
echo "== bench --gpus $N --workload c4"
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 --workload c4 --no-builder-timing > gpurun_out/j_bench_n${N}_c4.json 2> gpurun_out/j_bench_n${N}_c4.err; echo "rc=$?"; summ gpurun_out/j_bench_n${N}_c4.json; grep -iE "error|Traceback" gpurun_out/j_bench_n${N}_c4.err | head -5
