#!/bin/bash
# round 2, GPU call E (2 GPUs): the multi-GPU pipelines through bench.py under torchrun, and the in-process group API on two real devices
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
N=${1:-2}
echo "== group API on $N devices"
timeout 600 python - $N > gpurun_out/e_group.log 2>&1 <<'PY'
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
echo "group rc=$?"; tail -5 gpurun_out/e_group.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
echo "== bench --gpus $N (tiles, peer barrier; sort-last sub-object)"
timeout 1200 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/e_bench_n${N}_tiles.json 2> gpurun_out/e_bench_n${N}_tiles.err; echo "rc=$?"; tail -c 2500 gpurun_out/e_bench_n${N}_tiles.json; grep -E "rank 0|multi-GPU|sort-last|Error|error" gpurun_out/e_bench_n${N}_tiles.err | tail -12
echo "== bench --gpus $N --nccl-barrier --static-tiles --no-sortlast"
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 --nccl-barrier --static-tiles --no-sortlast --no-builder-timing > gpurun_out/e_bench_n${N}_tiles_nccl.json 2> gpurun_out/e_bench_n${N}_tiles_nccl.err; echo "rc=$?"; tail -c 700 gpurun_out/e_bench_n${N}_tiles_nccl.json
echo "== bench --gpus $N --partition sortlast"
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 --partition sortlast --no-sortlast --no-builder-timing > gpurun_out/e_bench_n${N}_sortlast.json 2> gpurun_out/e_bench_n${N}_sortlast.err; echo "rc=$?"; tail -c 900 gpurun_out/e_bench_n${N}_sortlast.json; tail -3 gpurun_out/e_bench_n${N}_sortlast.err
