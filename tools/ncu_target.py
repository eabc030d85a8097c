#!/usr/bin/env python
"""Target for the `ncu --set full` capture (GPU box only): warm up, then ONE opaque frame, ONE
translucent frame and ONE 1 M-point sampling launch between cudaProfilerStart/Stop.

    ncu --set full --clock-control none --import-source on --profile-from-start off \
        -o gpurun_out/ncu_c3 python tools/ncu_target.py --workload c3
"""
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
    fd = bench.frame_desc(R, domain, image, (float(field.min()), float(field.max()))); fd.apply(r)
    alpha, rgb = fd.xf[0]
    npts = 1 << 20
    pts = synth.random_points(npts, [0] * 3, bench.dom3(domain), seed=1)
    d_pts = torch.from_numpy(pts).cuda(); d_val = torch.empty(npts, dtype=torch.float32, device="cuda"); d_ok = torch.empty(npts, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def frames(scale, n):
        r.updateXF(0, alpha, rgb, fd.xf_domain[0], scale)
        for _ in range(n): r.render(sync=False)
        torch.cuda.synchronize()
    frames(0.02, 4); frames(1.0, 4)
    r.sample_points_device(d_pts.data_ptr(), npts, 0, d_val.data_ptr(), d_ok.data_ptr()); flush.fill_(1); torch.cuda.synchronize()
    torch.cuda.profiler.start()
    r.render(sync=False); torch.cuda.synchronize()                          # opaque frame (tile order from the previous opaque frame)
    torch.cuda.profiler.stop()
    frames(0.02, 3)
    torch.cuda.profiler.start()
    r.render(sync=False); torch.cuda.synchronize()                          # translucent frame
    r.sample_points_device(d_pts.data_ptr(), npts, 0, d_val.data_ptr(), d_ok.data_ptr()); torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("ncu target done", flush=True)


if __name__ == "__main__":
    main()
