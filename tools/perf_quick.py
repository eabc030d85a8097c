#!/usr/bin/env python
"""Quick A/B timing of the hot kernels on the C2 scene (GPU box only; not the bench).

    python tools/perf_quick.py [--workload c2] [--tag name]
    EXA_LIB=owlexabrick_b200/libexabrick_b200.<variant>.so python tools/perf_quick.py --tag <variant>

The scene (bricks + field) is cached under /dev/shm so that several variants can be timed in one
gpurun call.  Prints one JSON line: opaque / translucent frame ms, 1 M point sampling us.
"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--tag", default="base")
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    a.workload = os.environ.get("PQ_WORKLOAD", a.workload)
    import torch
    import bench
    from owlexabrick_b200 import capi, renderer as R, synth
    domain, boxes, image = bench.WORKLOADS[a.workload]
    bricks, field = bench.load_or_build_scene(a.workload, R)
    r = R.Renderer(bricks, [field]); torch.cuda.synchronize()
    stream = torch.cuda.current_stream(); r.set_stream(stream.cuda_stream)
    W, H = image
    fd = bench.frame_desc(R, domain, image, (float(field.min()), float(field.max()))); fd.apply(r)
    alpha, rgb = fd.xf[0]
    out = {"tag": a.tag, "lib": os.path.basename(capi.LIB_PATH), "layout": r.scalar_layout()}

    def time_frames(n):
        for _ in range(3): r.render(sync=False)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(n): r.render(sync=False)
        e1.record(stream); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    for name, scale in (("opaque", 1.0), ("translucent", 0.02)):
        r.updateXF(0, alpha, rgb, fd.xf_domain[0], scale)
        st = r.render_stats()
        import ctypes as C
        dbg = (C.c_uint64 * 4)(); capi.check(capi.lib.exa_renderer_debug_counters(r.h, dbg))
        out[name + "_dbg"] = {"max_kd_nodes_per_ray": int(dbg[0]), "max_samples_per_ray": int(dbg[1]), "kd_nodes_total": int(dbg[2]), "segments": st["segments"]}
        ms = time_frames(a.steps)
        import ctypes as C
        hist = (C.c_int * 16)(); capi.check(capi.lib.exa_renderer_get_tile_cost_histogram(r.h, hist))
        out[name + "_tile_cost_hist"] = list(hist)
        out[name] = {"ms": round(ms, 4), "Gsamples_s": round(st["samples"] / ms / 1e6, 3), "alg_GBs": round(bench.algorithmic_bytes(st, W, H) / ms / 1e6, 1)}
    for gs in (False, True):
        r.setGradientShadingDVR(gs)
    r.setGradientShadingDVR(False); out["translucent_nograd_ms"] = round(time_frames(a.steps), 4); r.setGradientShadingDVR(True)
    npts = 1 << 20
    pts = synth.random_points(npts, [0] * 3, bench.dom3(domain), seed=1)
    d_pts = torch.from_numpy(pts).cuda(); d_val = torch.empty(npts, dtype=torch.float32, device="cuda"); d_ok = torch.empty(npts, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    pst = r.sample_points_stats(d_pts.data_ptr(), npts, 0)
    import ctypes as C
    dbg = (C.c_uint64 * 4)(); capi.check(capi.lib.exa_renderer_debug_counters(r.h, dbg))
    out["points_locate"] = {"kd_steps_per_point": round(int(dbg[2]) / npts, 3), "max_kd_steps": int(dbg[1]), "kd_steps_paid_per_warp": round(int(dbg[0]) / (npts / 32), 2),
                            "locgrid_cells": os.environ.get("EXA_LOCGRID_CELLS", "default 2M")}
    ts = []
    flush2 = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for i in range(8):
        flush.fill_(i)
        if os.environ.get("PQ_CLEAN_FLUSH"): flush2.sum()      # leave L2 full of CLEAN lines (no write-back during the kernel)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); r.sample_points_device(d_pts.data_ptr(), npts, 0, d_val.data_ptr(), d_ok.data_ptr()); e1.record(stream)
        torch.cuda.synchronize()
        if i >= 3: ts.append(e0.elapsed_time(e1))
    pb = 16 * npts + 44 * pst["segments"] + 36 * pst["brickVisits"] + 4 * pst["cornerLoads"]
    out["points"] = {"us": round(1e3 * float(np.mean(ts)), 2), "alg_GBs": round(pb / np.mean(ts) / 1e6, 1), "visits_per_pt": round(pst["brickVisits"] / npts, 3), "corners_per_pt": round(pst["cornerLoads"] / npts, 3)}
    out["checksum"] = float(d_val.double().sum().item())
    hbuf = (np.zeros(npts, np.float32), np.zeros(npts, np.uint8))    # the caller's result arrays, reused (touched once)
    r.sample_points(pts, 0, out=hbuf)                                # host buffers in and out (first call allocates the staging)
    t = time.perf_counter()
    for _ in range(5): hv, hok = r.sample_points(pts, 0, out=hbuf)
    hs = (time.perf_counter() - t) / 5
    t = time.perf_counter()
    for _ in range(5): r.sample_points(pts, 0)
    out["points_host_fresh_arrays_ms"] = round((time.perf_counter() - t) / 5 * 1e3, 3)
    out["points_host"] = {"ms": round(hs * 1e3, 3), "Gsamples_s": round(npts / hs / 1e9, 3), "equals_device": bool(np.array_equal(hv, d_val.cpu().numpy()))}
    if os.environ.get("PQ_DIAG"):
        # diagnostics: (a) points confined to a 128^3 box (data L2-resident), (b) the same random points sorted along x-major cells
        def run(p_np):
            dp = torch.from_numpy(np.ascontiguousarray(p_np)).cuda(); tt = []
            for i in range(8):
                flush.fill_(i); flush2.sum()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); r.sample_points_device(dp.data_ptr(), npts, 0, d_val.data_ptr(), d_ok.data_ptr()); e1.record(stream)
                torch.cuda.synchronize()
                if i >= 3: tt.append(e0.elapsed_time(e1))
            return round(1e3 * float(np.mean(tt)), 2)
        c = min(bench.dom3(domain)) / 2
        out["points_local128_us"] = run(synth.random_points(npts, [c - 64] * 3, [c + 64] * 3, seed=2))
        key = (pts[:, 2] // 16).astype(np.int64) * 4096 * 4096 + (pts[:, 1] // 16).astype(np.int64) * 4096 + (pts[:, 0] // 16).astype(np.int64)
        out["points_sorted_us"] = run(pts[np.argsort(key, kind="stable")])
        # no flush at all (everything that can be L2-resident is)
        tt = []
        for i in range(8):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream); r.sample_points_device(d_pts.data_ptr(), npts, 0, d_val.data_ptr(), d_ok.data_ptr()); e1.record(stream)
            torch.cuda.synchronize(); tt.append(e0.elapsed_time(e1))
        out["points_noflush_us"] = round(1e3 * float(np.mean(tt[3:])), 2)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
