#!/usr/bin/env python
"""What one rank of an N-GPU image-tile run does, timed on ONE GPU (GPU box only; not the bench).

    python tools/scale_emulate.py [--workload c2] [--image 1024x1024,3840x2160]

For N in 1,2,4,8 the renderer is told it is rank 0 of N (exa_renderer_set_partition, image tiles)
and its frame kernel is timed alone: the time of the N-GPU frame is max over ranks of that plus
the frame-completion all-reduce, so t(1)/t(N) is the upper bound of the scaling the kernel allows.
The per-tile cost histogram (log2 buckets of CTA cycles / 1024) shows the longest tile.
"""
import argparse, ctypes as C, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--image", default=None, help="WxH[,WxH...]; default: the workload's image")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--ranks", default="1,2,4,8")
    a = ap.parse_args()
    import torch
    import bench
    from owlexabrick_b200 import capi, renderer as R
    domain, boxes, _ = bench.WORKLOADS[a.workload]
    bricks, field = bench.load_or_build_scene(a.workload, R)
    r = R.Renderer(bricks, [field]); torch.cuda.synchronize()
    stream = torch.cuda.current_stream(); r.set_stream(stream.cuda_stream)
    frange = (float(field.min()), float(field.max()))

    def time_frames(n):
        for _ in range(4): r.render(sync=False)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(n): r.render(sync=False)
        e1.record(stream); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    for img in (a.image or "%dx%d" % tuple(bench.WORKLOADS[a.workload][2])).split(","):
        W, H = (int(x) for x in img.split("x"))
        fd = bench.frame_desc(R, domain, (W, H), frange); fd.apply(r)
        alpha, rgb = fd.xf[0]
        for name, scale in (("opaque", 1.0), ("translucent", 0.02)):
            r.updateXF(0, alpha, rgb, fd.xf_domain[0], scale)
            row = {"image": img, "xf": name}
            for n in (int(x) for x in a.ranks.split(",")):
                worst, per_rank = 0.0, []
                for rank in range(n):
                    r.set_partition(capi.PARTITION_IMAGE_TILES if n > 1 else capi.PARTITION_NONE, rank, n)
                    per_rank.append(round(time_frames(a.steps), 4))
                worst = max(per_rank)
                row["N%d_ranks" % n] = per_rank
                hist = (C.c_int * 16)(); capi.check(capi.lib.exa_renderer_get_tile_cost_histogram(r.h, hist))
                row["N%d_ms" % n] = round(worst, 4)
                row["N%d_hist" % n] = list(hist)
            row["speedup_bound_N8"] = round(row["N1_ms"] / row["N%d_ms" % int(a.ranks.split(",")[-1])], 2)
            print(json.dumps(row), flush=True)
        r.set_partition(capi.PARTITION_NONE, 0, 1)


if __name__ == "__main__":
    main()
