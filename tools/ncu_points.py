#!/usr/bin/env python
"""Target for an `ncu --metrics ...` pass over k_samplePoints with different point sets (GPU box only): where do the DRAM
sectors of the point query come from?  Launches, in this order, between cudaProfilerStart/Stop, each after an L2 flush:
random points; the same without outputs (values / flags not written); the same sorted by coarse cell; points confined to a
128^3 box; random points again with the gradient."""
import argparse, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    a = ap.parse_args()
    import torch
    import bench
    from owlexabrick_b200 import renderer as R, synth
    domain, _, image = bench.WORKLOADS[a.workload]
    bricks, field = bench.load_or_build_scene(a.workload, R)
    r = R.Renderer(bricks, [field]); torch.cuda.synchronize()
    stream = torch.cuda.current_stream(); r.set_stream(stream.cuda_stream)
    npts = 1 << 20
    dom = bench.dom3(domain)
    pts = synth.random_points(npts, [0] * 3, dom, seed=1)
    key = (np.floor(pts[:, 2] / 16).astype(np.int64) * 4096 + np.floor(pts[:, 1] / 16).astype(np.int64)) * 4096 + np.floor(pts[:, 0] / 16).astype(np.int64)
    srt = pts[np.argsort(key, kind="stable")]
    c = [d / 2 for d in dom]
    loc = synth.random_points(npts, [x - 64 for x in c], [x + 64 for x in c], seed=2)
    d_val = torch.empty(npts, dtype=torch.float32, device="cuda"); d_ok = torch.empty(npts, dtype=torch.uint8, device="cuda")
    d_g = torch.empty((npts, 3), dtype=torch.float32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def go(p, values=True, grad=False):
        d = torch.from_numpy(np.ascontiguousarray(p)).cuda()
        r.sample_points_device(d.data_ptr(), npts, 0, d_val.data_ptr(), d_ok.data_ptr()); torch.cuda.synchronize()     # warm (untimed, unprofiled)
        flush.fill_(1); flush.sum(); torch.cuda.synchronize()
        torch.cuda.profiler.start()
        r.sample_points_device(d.data_ptr(), npts, 0, d_val.data_ptr() if values else 0, d_ok.data_ptr() if values else 0, d_g.data_ptr() if grad else None)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    go(pts); go(pts, values=False); go(srt); go(loc); go(pts, grad=True)
    print("ncu points target done", flush=True)


if __name__ == "__main__":
    main()
