#!/usr/bin/env python
"""Measure the C5-shaped workload (BASELINE.json configs[4]: molecular-cloud-shaped 3-channel AMR,
multichannel DVR at 8K + streamline tracing) at a reduced scale on ONE GPU (GPU box only; not the
bench — bench.py measures C3).

    python tools/c5_measure.py [--scale 2] [--steps 10]

scale k: synth.cloud_shaped(k), domain 2048/k, ~2.07 B / k^3 cells (k = 1 is the frozen C5).
Prints one JSON line: 3-channel DVR frame at 7680x4320 (ms, frames/s, reconstruction samples/s,
algorithmic GB/s and fraction of the measured HBM peak), the RK4 tracer (1000 seeds x 100 steps,
12 reconstructions x 3 channels per step) and the frame with the traces drawn as cones.
"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=int, default=2)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--image", default="7680x4320")
    a = ap.parse_args()
    import torch
    import bench, scenes
    from concurrent.futures import ThreadPoolExecutor
    from owlexabrick_b200 import capi, renderer as R, synth
    W, H = (int(x) for x in a.image.split("x"))
    t = time.time()
    cells, dom = synth.cloud_shaped(a.scale)
    n = len(cells)
    # 3 velocity channels, chunks on a thread pool (numpy releases the GIL in the ufuncs)
    fields = [np.empty(n, np.float32) for _ in range(3)]
    chunk = 1 << 22

    def work(b):
        v = synth.cloud_velocity(cells[b:b + chunk], dom)
        for c in range(3):
            fields[c][b:b + chunk] = v[c]
    with ThreadPoolExecutor(max_workers=min(32, os.cpu_count() or 1)) as ex:
        list(ex.map(work, range(0, n, chunk)))
    bench.log("c5 scale %d: %d cells, domain %d^3, generated in %.1fs" % (a.scale, n, dom, time.time() - t))
    t = time.time()
    bricks = R.ExaBricks.build(cells)
    del cells
    bench.log("exa_bricks_build: %d bricks in %.1fs" % (bricks.num_bricks, time.time() - t))
    t = time.time()
    r = R.Renderer(bricks, fields)
    torch.cuda.synchronize()
    regions, leaf = r.get_abr()
    bench.log("renderer created in %.1fs (%d regions, L=%d, kd %s)" % (time.time() - t, len(regions), len(leaf), r.kd_info()))
    stream = torch.cuda.current_stream(); r.set_stream(stream.cuda_stream)
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H), 3)
    fd.opacity_scale, fd.dt = bench.OPACITY_SCALE, bench.DT
    fd.xf_domain = [(float(f.min()), float(f.max())) for f in fields]
    fd.apply(r)
    peak = bench.peaks()[0]

    def time_frames(k):
        for _ in range(3): r.render(sync=False)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(k): r.render(sync=False)
        e1.record(stream); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / k

    def frame_line(tag):
        st = r.render_stats()
        ms = time_frames(a.steps)
        ab = bench.algorithmic_bytes(st, W, H)
        return {"ms_per_frame": round(ms, 4), "frames_per_s": round(1e3 / ms, 2), "samples_per_frame": st["samples"],
                "samples_per_s": st["samples"] / ms * 1e3, "algorithmic_gbs": round(ab / ms / 1e6, 1), "frac_of_hbm_peak": round(ab / ms / 1e6 / peak, 4)}
    lay, nbytes = r.scalar_layout()
    out = {"workload": "c5 shape (synth.cloud_shaped(%d)): %d cells, domain %d^3, 3 channels, multichannel DVR %dx%d + streamlines" % (a.scale, n, dom, W, H),
           "cells": n, "bricks": bricks.num_bricks, "regions": len(regions), "n_gpus": 1,
           "scalar_layout": "dual-cell records" if lay == capi.LAYOUT_DUAL else "flat", "scalar_bytes": nbytes, "hbm_peak_gbs": peak,
           "dvr_3ch": frame_line("dvr")}
    r.updateXF(0, fd.xf[0][0], fd.xf[0][1], fd.xf_domain[0], 0.02)
    r.updateXF(1, fd.xf[1][0], fd.xf[1][1], fd.xf_domain[1], 0.02)
    r.updateXF(2, fd.xf[2][0], fd.xf[2][1], fd.xf_domain[2], 0.02)
    out["dvr_3ch_translucent"] = dict(frame_line("translucent"), xf_opacity_scale=0.02)
    # tracer: 1000 seeds x 100 time steps (viewer defaults, OptixRenderer.h:155-165), seed region = viewer default
    T, NT = 100, 1000
    r.setTracerParams(NT, T, dom / 100.0, (0, 1, 2))
    r.setTracerEnabled(True)
    ts = []
    for i in range(4):
        r.resetTracer()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); r.traceAll(); e1.record(stream); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    tr = r.readTraces()
    alive = int((tr[..., 0] < 2e10).sum())
    ms = float(np.mean(ts[1:]))
    out["streamlines"] = {"traces": NT, "timesteps": T, "ms_trace_all": round(ms, 4), "alive_vertices": alive,
                          "reconstructions_per_s": 12.0 * (alive - NT) / ms * 1e3}
    out["dvr_3ch_translucent_with_cones"] = frame_line("cones")
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
