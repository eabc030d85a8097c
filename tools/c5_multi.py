#!/usr/bin/env python
"""Config C5 (BASELINE.json configs[4]: molecular-cloud-shaped 3-channel AMR, multichannel DVR at 8K + 1000 x 100
streamlines) on N GPUs through the sort-last pipeline with the DATA partitioned (GPU box only; run under torchrun):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/c5_multi.py --scale 1

scale k: synth.cloud_shaped(k) (k = 1 is the frozen C5: 2 068 348 416 cells, whose 3 x 66 GB of dual-cell records do not
fit one GPU).  Rank 0 generates the cells and the three velocity channels, builds the bricks with the GPU builder and
leaves bricks + fields in /dev/shm; every rank maps them and uploads ONLY the bricks of its kd part
(exa_renderer_create_partitioned).  Measured, device-timed, max over ranks: the 3-channel DVR frame (default XF and opacity
x 0.02), end to end with every rank copying its span to the shared host frame, the frame with the streamlines drawn as
cones (surface records composited too).  The RK4 tracer needs the velocity field along whole streamlines, i.e. every part:
it runs on rank 0 over a FLAT-layout replica of the scene (3 x 4 B per cell fit one GPU), which also renders the same
frame on one GPU: that is the parity reference of the composited frame and the 1-GPU time it is compared with.
Prints one JSON line on rank 0."""
import argparse, ctypes as C, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=int, default=4)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--image", default="7680x4320")
    ap.add_argument("--no-replica", action="store_true", help="skip the flat 1-GPU replica (parity reference + tracer)")
    ap.add_argument("--modes", default="sortlast", help="comma list: sortlast (data partitioned, dual-cell records), tiles (flat layout replicated on every GPU, image tiles)")
    a = ap.parse_args()
    if os.environ.get("OMP_NUM_THREADS") == "1":
        os.environ.pop("OMP_NUM_THREADS")                          # torchrun's default; rank 0 builds with all cores
    import torch
    import torch.distributed as dist
    import bench, scenes
    from concurrent.futures import ThreadPoolExecutor
    from owlexabrick_b200 import capi, multigpu, renderer as R, synth
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, H = (int(x) for x in a.image.split("x"))
    log = lambda *x: print("[c5 r%d]" % rank, *x, file=sys.stderr, flush=True)
    base = "/dev/shm/exa_c5_s%d" % a.scale
    info = {}
    avail_gb = 0.0
    try:
        avail_gb = [int(l.split()[1]) for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0] / 1048576.0
    except Exception:
        pass
    need_gb = 2.07e9 / a.scale ** 3 * (16 + 12 + 4 + 4 + 4 * (world - 1)) / 1e9 * 1.15      # cells, fields, ids (+ file), ids per other rank
    if rank == 0:
        log("host memory available %.0f GB, estimated need %.0f GB" % (avail_gb, need_gb))
    if avail_gb and avail_gb < need_gb:
        if rank == 0:
            print(json.dumps({"workload": "c5 scale %d" % a.scale, "skipped": "host memory: %.0f GB available, %.0f GB needed" % (avail_gb, need_gb)}), flush=True)
        dist.destroy_process_group()
        return
    if rank == 0:
        t = time.time()
        cells, dom = synth.cloud_shaped(a.scale)
        n = len(cells)
        fields = [np.lib.format.open_memmap("%s_f%d.npy" % (base, c), mode="w+", dtype=np.float32, shape=(n,)) for c in range(3)]
        chunk = 1 << 22

        def work(b):
            v = synth.cloud_velocity(cells[b:b + chunk], dom)
            for c in range(3):
                fields[c][b:b + chunk] = v[c]
        with ThreadPoolExecutor(max_workers=min(32, os.cpu_count() or 1)) as ex:
            list(ex.map(work, range(0, n, chunk)))
        info["generate_s"] = round(time.time() - t, 1)
        t = time.time()
        try:
            bricks = R.ExaBricks.build_gpu(cells, device=local)
            info["builder"] = dict(kind="gpu", **{k: round(v, 3) if isinstance(v, float) else v for k, v in bricks.timings.items()})
        except Exception as e:                                     # the builder's working set (≈ 55 B per cell at its peak) exceeds the HBM: host builder
            log("GPU builder failed (%s): host builder" % e)
            if avail_gb and avail_gb < need_gb + 70 * n / 1e9:
                raise
            torch.cuda.empty_cache()
            bricks = R.ExaBricks.build(cells)
            info["builder"] = {"kind": "host"}
        info["builder"]["wall_s"] = round(time.time() - t, 1)
        del cells
        bricks.save(base + ".bricks")
        xfd = [(float(f.min()), float(f.max())) for f in fields]
        meta = torch.tensor([n, dom] + [x for p in xfd for x in p], dtype=torch.float64, device="cuda")
        log("scene: %d cells, domain %d^3, %d bricks; %s" % (n, dom, bricks.num_bricks, info))
    else:
        meta = torch.zeros(8, dtype=torch.float64, device="cuda")
    dist.broadcast(meta, 0)
    n, dom = int(meta[0].item()), int(meta[1].item())
    xfd = [(float(meta[2 + 2 * c].item()), float(meta[3 + 2 * c].item())) for c in range(3)]
    if rank != 0:
        bricks = R.ExaBricks.load(base + ".bricks")
        fields = [np.load("%s_f%d.npy" % (base, c), mmap_mode="r") for c in range(3)]
    modes = a.modes.split(",")
    if "tiles" in modes:
        run_tiles(a, dist, torch, capi, multigpu, R, scenes, bench, bricks, fields, n, dom, xfd, rank, world, local, W, H, log, info)
    if "sortlast" not in modes:
        cleanup(rank, base); dist.destroy_process_group(); return
    t = time.time()
    r = R.Renderer(bricks, fields, device=local, part=(rank, world))
    torch.cuda.synchronize()
    nb_res, nc_res = r.residency()
    lay, nbytes = r.scalar_layout()
    log("partitioned renderer in %.1fs: %d of %d bricks, %d cells resident, scalar layout %d (%.1f GB)" % (time.time() - t, nb_res, bricks.num_bricks, nc_res, lay, nbytes / 1e9))
    create_s = torch.tensor([time.time() - t], dtype=torch.float64, device="cuda")
    dist.all_reduce(create_s, op=dist.ReduceOp.MAX)
    stream = torch.cuda.current_stream(); r.set_stream(stream.cuda_stream)
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H), 3)
    fd.opacity_scale, fd.dt, fd.xf_domain = bench.OPACITY_SCALE, bench.DT, xfd
    fd.apply(r)

    def gather(v, dtype=torch.float64):
        t_ = torch.tensor([v], dtype=dtype, device="cuda")
        out = [torch.zeros(1, dtype=dtype, device="cuda") for _ in range(world)]
        dist.all_gather(out, t_)
        return [x.item() for x in out]

    def set_opacity(rr, scale):
        for c in range(3):
            rr.updateXF(c, fd.xf[c][0], fd.xf[c][1], fd.xf_domain[c], scale)

    def timed(pipe, k, readback=False, rr=None):
        rr = rr or r
        for _ in range(3):
            pipe.step(readback)
        pipe.drain(); torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(k):
            if readback:
                rr.updateCamera(*fd.cam); set_opacity(rr, cur_scale[0])
            pipe.step(readback)
        pipe.drain(); e1.record(stream)
        if readback:
            pipe.finish()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        return max(gather(e0.elapsed_time(e1) / k)), max(gather(wall / k * 1e3))

    def own_kernel_ms(k=5):
        p = pipe.partials.ptr[rank][0]
        for _ in range(2):
            r.render_partial(p)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); dist.barrier()
        e0.record(stream)
        for _ in range(k):
            r.render_partial(p)
        e1.record(stream); torch.cuda.synchronize()
        return gather(e0.elapsed_time(e1) / k)

    cur_scale = [fd.opacity_scale]
    pipe = multigpu.SortLastPipeline(dist, torch, capi, np, r, rank, world, local, W, H, stream, tag="c5_%d" % os.getppid())
    out = {"workload": "c5 (synth.cloud_shaped(%d)): %d cells, domain %d^3, 3 channels, multichannel DVR %dx%d + streamlines" % (a.scale, n, dom, W, H),
           "cells": n, "bricks": bricks.num_bricks, "n_gpus": world, "mode": "sort-last, data partitioned along the ABR kd-tree",
           "resident_cells_per_rank": [int(x) for x in gather(nc_res, torch.int64)], "resident_bricks_per_rank": [int(x) for x in gather(nb_res, torch.int64)],
           "scalar_bytes_per_rank": [int(x) for x in gather(nbytes, torch.int64)], "scalar_layout": "dual-cell records" if lay == capi.LAYOUT_DUAL else "flat",
           "renderer_create_s_max": round(create_s.item(), 2), "build_times_rank0": r.build_times(), "scene": info}
    st = r.render_stats()
    samples = sum(gather(st["samples"], torch.int64))
    ms, _ = timed(pipe, a.steps)
    out["dvr_3ch"] = {"ms_per_frame": round(ms, 4), "frames_per_s": round(1e3 / ms, 2), "samples_per_frame_all_parts": int(samples), "samples_per_s": samples / ms * 1e3,
                      "own_partial_kernel_ms_per_rank": [round(x, 3) for x in own_kernel_ms()]}
    ems, wall = timed(pipe, a.steps, readback=True)
    out["dvr_3ch"]["e2e_frames_per_s"] = round(1e3 / wall, 2)
    out["dvr_3ch"]["e2e_d2h_bytes_per_rank"] = pipe.d2h_bytes
    b = pipe.step(readback=True); pipe.drain(); pipe.finish(); torch.cuda.synchronize(); dist.barrier()
    frame_opaque = pipe.host_frame(b).copy() if rank == 0 else None
    cur_scale[0] = 0.02
    set_opacity(r, 0.02)
    st = r.render_stats()
    samples = sum(gather(st["samples"], torch.int64))
    ms, _ = timed(pipe, max(3, a.steps // 2))
    out["dvr_3ch_translucent"] = {"xf_opacity_scale": 0.02, "ms_per_frame": round(ms, 4), "frames_per_s": round(1e3 / ms, 2), "samples_per_frame_all_parts": int(samples),
                                  "samples_per_s": samples / ms * 1e3, "own_partial_kernel_ms_per_rank": [round(x, 3) for x in own_kernel_ms(3)]}
    b = pipe.step(readback=True); pipe.drain(); pipe.finish(); torch.cuda.synchronize(); dist.barrier()
    frame_trans = pipe.host_frame(b).copy() if rank == 0 else None
    cur_scale[0] = fd.opacity_scale
    set_opacity(r, fd.opacity_scale)

    # ---- rank 0: flat 1-GPU replica = parity reference, 1-GPU time, tracer ----
    NT, T = 1000, 100
    traces = torch.zeros((NT, T, 3), dtype=torch.float32, device="cuda")
    if rank == 0 and not a.no_replica:
        t = time.time()
        full = R.Renderer(bricks, fields, device=local, layout=capi.LAYOUT_FLAT)
        torch.cuda.synchronize()
        full.set_stream(stream.cuda_stream); fd.apply(full)
        regions, leaf = full.get_abr()
        out["regions"] = len(regions); out["kd"] = full.kd_info()
        log("flat replica in %.1fs (%d regions)" % (time.time() - t, len(regions)))

        def one_gpu(scale, k):
            set_opacity(full, scale)
            for _ in range(2): full.render(sync=False)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(k): full.render(sync=False)
            e1.record(stream); torch.cuda.synchronize()
            return e0.elapsed_time(e1) / k, full.read_rgba8().copy(), full.render_stats()["samples"]
        for key, scale, frame in (("dvr_3ch", fd.opacity_scale, frame_opaque), ("dvr_3ch_translucent", 0.02, frame_trans)):
            ms1, ref, s1 = one_gpu(scale, 3)
            d = np.abs(scenes.unpack_rgba8(frame).astype(np.int32) - scenes.unpack_rgba8(ref).astype(np.int32))
            out[key]["one_gpu_flat_layout"] = {"ms_per_frame": round(ms1, 3), "samples_per_frame": int(s1)}
            out[key]["speedup_vs_one_gpu_flat"] = round(ms1 / out[key]["ms_per_frame"], 2)
            out[key]["vs_one_gpu_frame"] = {"max_abs_rgba8_diff": int(d.max()), "pixels_differing": int((d.max(axis=-1) > 0).sum()), "pixels_over_2": int((d.max(axis=-1) > 2).sum()),
                                            "note": "per-part early termination at 0.98 (H1) bounds the deviation"}
        set_opacity(full, fd.opacity_scale)
        full.setTracerParams(NT, T, dom / 100.0, (0, 1, 2)); full.setTracerEnabled(True)
        ts = []
        for i in range(3):
            full.resetTracer()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream); full.traceAll(); e1.record(stream); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        tr = full.readTraces()
        alive = int((tr[..., 0] < 2e10).sum())
        out["streamlines"] = {"traces": NT, "timesteps": T, "ms_trace_all": round(float(np.mean(ts[1:])), 4), "alive_vertices": alive, "where": "rank 0, flat-layout replica"}
        traces.copy_(torch.from_numpy(tr))
        full.close(); del full
        torch.cuda.empty_cache()
    dist.broadcast(traces, 0)
    if not a.no_replica:
        # the traces as cones through the sort-last path: every rank draws them (replicated surface) and clips its samples at them
        r.load_traces(traces.cpu().numpy(), T)
        r.setTracerEnabled(True)
        pipe2 = multigpu.SortLastPipeline(dist, torch, capi, np, r, rank, world, local, W, H, stream, surf=True, tag="c5b_%d" % os.getppid())
        ms, _ = timed(pipe2, max(3, a.steps // 2))
        out["dvr_3ch_with_cones"] = {"ms_per_frame": round(ms, 4), "frames_per_s": round(1e3 / ms, 2), "note": "1000 x 100 streamline vertices as rounded cones, replicated; surface records composited"}
        pipe2.close()
    if rank == 0:
        print(json.dumps(out), flush=True)
    pipe.close()
    cleanup(rank, base)
    dist.destroy_process_group()


def cleanup(rank, base):
    if rank == 0:
        for f in ["%s_f%d.npy" % (base, c) for c in range(3)] + [base + ".bricks"]:
            try: os.unlink(f)
            except OSError: pass


def run_tiles(a, dist, torch, capi, multigpu, R, scenes, bench, bricks, fields, n, dom, xfd, rank, world, local, W, H, log, info):
    """C5 with the scene REPLICATED in the flat layout (3 x 4 B per cell: 24.8 GB per GPU) and the frame split into image
    tiles: every GPU renders its cost-balanced share of the tiles into the peer-mapped frame strips"""
    t = time.time()
    full = R.Renderer(bricks, fields, device=local, layout=capi.LAYOUT_FLAT)
    torch.cuda.synchronize()
    create_s = time.time() - t
    stream = torch.cuda.current_stream(); full.set_stream(stream.cuda_stream)
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H), 3)
    fd.opacity_scale, fd.dt, fd.xf_domain = bench.OPACITY_SCALE, bench.DT, xfd
    fd.apply(full)
    lay, nbytes = full.scalar_layout()

    def gmax(v):
        t_ = torch.tensor([v], dtype=torch.float64, device="cuda"); dist.all_reduce(t_, op=dist.ReduceOp.MAX); return float(t_.item())

    def set_opacity(scale):
        for c in range(3):
            full.updateXF(c, fd.xf[c][0], fd.xf[c][1], fd.xf_domain[c], scale)

    def one_gpu(k):
        for _ in range(2): full.render(sync=False)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(k): full.render(sync=False)
        e1.record(stream); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / k, full.read_rgba8().copy(), full.render_stats()["samples"]
    ref = {}
    for key, scale in (("dvr_3ch", fd.opacity_scale), ("dvr_3ch_translucent", 0.02)):
        set_opacity(scale)
        ref[key] = one_gpu(3)                                        # every rank holds the whole scene: the same 1-GPU frame everywhere
    set_opacity(fd.opacity_scale)
    regions, leaf = full.get_abr()
    pipe = multigpu.TilePipeline(dist, torch, capi, np, full, rank, world, local, W, H, stream, tag="c5t_%d" % os.getppid())

    def timed(k, scale, readback=False):
        for _ in range(3): pipe.step(readback)
        pipe.drain(); torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record(stream)
        for _ in range(k):
            if readback:
                full.updateCamera(*fd.cam); set_opacity(scale)
            pipe.step(readback)
        pipe.drain(); e1.record(stream)
        if readback: pipe.finish()
        torch.cuda.synchronize()
        return gmax(e0.elapsed_time(e1) / k), gmax((time.perf_counter() - t0) / k * 1e3)
    out = {"workload": "c5 (synth.cloud_shaped(%d)): %d cells, domain %d^3, 3 channels, multichannel DVR %dx%d + streamlines" % (a.scale, n, dom, W, H),
           "cells": n, "bricks": bricks.num_bricks, "regions": len(regions), "kd": full.kd_info(), "n_gpus": world,
           "mode": "image tiles, scene replicated in the flat layout", "scalar_layout": "dual-cell records" if lay == capi.LAYOUT_DUAL else "flat",
           "scalar_bytes_per_rank": nbytes, "renderer_create_s_max": round(gmax(create_s), 2), "build_times_rank0": full.build_times(), "scene": info}
    for key, scale, k in (("dvr_3ch", fd.opacity_scale, a.steps), ("dvr_3ch_translucent", 0.02, max(3, a.steps // 2))):
        set_opacity(scale)
        bal = pipe.rebalance()
        ms, _ = timed(k, scale)
        _, wall = timed(k, scale, readback=True)
        b = pipe.step(readback=True); pipe.drain(); pipe.finish(); torch.cuda.synchronize(); dist.barrier()
        ms1, img1, s1 = ref[key]
        o = {"ms_per_frame": round(ms, 4), "frames_per_s": round(1e3 / ms, 2), "samples_per_frame": int(s1), "samples_per_s": s1 / ms * 1e3,
             "e2e_frames_per_s": round(1e3 / wall, 2), "e2e_d2h_bytes_per_rank": pipe.d2h_bytes, "tile_balance_max_over_mean": bal,
             "one_gpu_flat_layout": {"ms_per_frame": round(ms1, 3)}, "speedup_vs_one_gpu_flat": round(ms1 / ms, 2)}
        if rank == 0:
            d = np.abs(scenes.unpack_rgba8(pipe.host_frame(b)) - scenes.unpack_rgba8(img1))
            o["vs_one_gpu_frame"] = {"max_abs_rgba8_diff": int(d.max()), "pixels_differing": int((d.max(axis=-1) > 0).sum())}
        out[key] = o
    # streamlines: every GPU holds the whole field, so every GPU advects all seeds itself (7-8 ms) and draws the cones in its tiles
    set_opacity(fd.opacity_scale)
    NT, T = 1000, 100
    full.setTracerParams(NT, T, dom / 100.0, (0, 1, 2)); full.setTracerEnabled(True)
    ts = []
    for i in range(3):
        full.resetTracer()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); full.traceAll(); e1.record(stream); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    out["streamlines"] = {"traces": NT, "timesteps": T, "ms_trace_all": round(gmax(float(np.mean(ts[1:]))), 4), "where": "every rank, replicated"}
    pipe.rebalance()
    ms, _ = timed(max(3, a.steps // 2), fd.opacity_scale)
    out["dvr_3ch_with_cones"] = {"ms_per_frame": round(ms, 4), "frames_per_s": round(1e3 / ms, 2)}
    if rank == 0:
        print(json.dumps(out), flush=True)
    pipe.close()
    full.close()
    torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
