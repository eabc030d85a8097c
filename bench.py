#!/usr/bin/env python
"""bench.py — the hot path's headline benchmark (BASELINE.json: reconstructed samples/s and
4K DVR frames/s at 1/2/4/8 B200; % of HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c3|c2|c1|small]

A "step" is one DVR frame of the workload through the fused frame kernel.  The default
workload is BASELINE.json configs[2] ("c3": exajet-shaped synthetic 4-level AMR, ~650 M
cells, one scalar, DVR 3840x2160) — the configuration the 4K frames/s metric and the
8-GPU target are quoted on; it fits one GPU, so the SAME frame is measured at every N
(scaling "strong").  configs[1] (~95 M cells, DVR 1024^2, 1 GPU) is `--workload c2`; its
round-1 lines are under profiles/bench_c2_r01_*.json.  N>1 renders the frame as image tiles
(data replicated, tile t -> rank t mod N, pixels stored straight into rank 0's frame buffer
over NVLink) or, with --partition sortlast, as a kd spatial partition with an NCCL binary
swap.  Prints ONE JSON line (see the task contract for the keys).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

# torchrun exports OMP_NUM_THREADS=1 unless the user set it; the scene set-up (host builder, rank 0)
# and the CPU reference arm (rank 0 alone) are OpenMP code, so undo that default before any
# OpenMP runtime is loaded: rank 0 gets the cores the other ranks do not need
if os.environ.get("OMP_NUM_THREADS") == "1" and "TORCHELASTIC_RUN_ID" in os.environ:
    _w = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
    _n = os.cpu_count() or 1
    if "--impl" in sys.argv and "reference" in sys.argv:
        # the CPU arm: ranks > 0 exit at once, so rank 0 has the whole host at EVERY N (same thread count for N = 1, 2, 4, 8)
        os.environ["OMP_NUM_THREADS"] = str(_n)
    else:
        os.environ["OMP_NUM_THREADS"] = str(max(1, _n - 2 * (_w - 1)) if os.environ.get("RANK", "0") == "0" else 2)

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (domain, shape, image)           SURVEY.md §8(d)
    # shape = list of nested boxes (synth.nested_boxes) or ("exajet", scale) (synth.exajet_shaped)
    "c3": ((4096, 1024, 2048), ("exajet", 1), (3840, 2160)),
    "c3s2": ((2048, 512, 1024), ("exajet", 2), (1920, 1080)),
    "c3s4": ((1024, 256, 512), ("exajet", 4), (960, 540)),
    # C4 landing-gear-shaped: 7 levels, small deeply nested off-centre level-0 core -> 262 277 632 cells; iso + DVR
    "c4": (8192, [((2048, 1536, 2560), (2608, 2080, 3104)), ((1920, 1408, 2432), (2752, 2208, 3232)), ((1664, 1152, 2176), (3008, 2496, 3520)),
                  ((1280, 768, 1792), (3328, 2816, 3840)), ((768, 256, 1280), (3840, 3328, 4352)), ((0, 0, 0), (5120, 4096, 5120))], (3840, 2160)),
    "c4s4": (2048, [((512, 384, 640), (652, 520, 776)), ((480, 352, 608), (688, 552, 808)), ((416, 288, 544), (752, 624, 880)),
                    ((320, 192, 448), (832, 704, 960)), ((192, 64, 320), (960, 832, 1088)), ((0, 0, 0), (1280, 1024, 1280))], (960, 540)),
    "c2": (1024, [(320, 704), (192, 832)], (1024, 1024)),
    "c1": (256, [(72, 184), (32, 224)], (1024, 1024)),
    "small": (64, [(18, 46), (8, 56)], (256, 256)),
}
OPACITY_SCALE = 1.0      # viewer default xfOpacityScale (viewer.cpp:113)
DT = 0.5                 # viewer default (viewer.cpp:115)
SCENE_CACHE_VERSION = 1  # bump when a generator changes


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def dom3(domain):
    return [float(domain)] * 3 if np.isscalar(domain) else [float(x) for x in domain]


def field_threaded(cells, domain, channel=0, chunk=1 << 20):
    """synth.analytic_field over chunks on a thread pool (numpy releases the GIL inside the
    ufuncs); same values as the single call — the noise counter is the global cell index"""
    from concurrent.futures import ThreadPoolExecutor
    from owlexabrick_b200 import synth
    out = np.empty(len(cells), np.float32)

    def work(b):
        out[b:b + chunk] = synth.analytic_field_chunked(cells[b:b + chunk], domain, channel, chunk=chunk, first=b)
    with ThreadPoolExecutor(max_workers=min(32, os.cpu_count() or 1)) as ex:
        list(ex.map(work, range(0, len(cells), chunk)))
    return out


def make_workload(name, with_field=True):
    from owlexabrick_b200 import synth
    domain, shape, image = WORKLOADS[name]
    t = time.time()
    if isinstance(shape, tuple) and shape[0] == "exajet":
        cells, _ = synth.exajet_shaped(shape[1])
    else:
        cells = synth.nested_boxes(domain, shape)
    field = field_threaded(cells, domain, 0) if with_field else None
    log("workload %s: %d cells generated in %.1fs" % (name, len(cells), time.time() - t))
    return cells, field, domain, image


BUILDER_INFO = {}


def load_or_build_scene(name, R, cache_dir=None, gpu_builder=False, time_builder=False):
    """-> (ExaBricks, field).  Cells -> bricks with the library's builder: the GPU builder (exa_bricks_build_gpu) on the
    product arm, the OpenMP host builder (exa_bricks_build) on the CPU arm; both bit-exact with the reference's exaBuilder
    (tests/test_host.py, tests/test_gpu_builder.py).  The result is cached in /dev/shm so that the two bench arms and the
    ranks of one box build once; time_builder: when the cache exists already, the cells are generated again and the GPU
    build is run and timed anyway (and compared with the cached bricks)."""
    cache_dir = cache_dir or os.environ.get("EXA_SCENE_CACHE", "/dev/shm")
    base = os.path.join(cache_dir, "exa_scene_%s_v%d" % (name, SCENE_CACHE_VERSION))

    def gpu_build(cells):
        t = time.time()
        b = R.ExaBricks.build_gpu(cells, device=int(os.environ.get("LOCAL_RANK", "0")))
        BUILDER_INFO.update({"exa_builder_gpu": {**{k: (round(v, 3) if isinstance(v, float) else v) for k, v in b.timings.items()},
                                                  "wall_s": round(time.time() - t, 3), "cells": int(len(cells)), "bricks": b.num_bricks,
                                                  "cells_per_s": round(len(cells) / max(b.timings["total_s"], 1e-9))}})
        log("exa_bricks_build_gpu: %d cells -> %d bricks, %s" % (len(cells), b.num_bricks, b.timings))
        return b
    if os.path.exists(base + ".ok"):
        if gpu_builder and time_builder:
            try:
                cells, _, _, _ = make_workload(name, with_field=False)
                b = gpu_build(cells)
                del cells
                cached = R.ExaBricks.load(base + ".bricks")
                BUILDER_INFO["exa_builder_gpu"]["identical_to_cached_bricks"] = bool(b.headers.tobytes() == cached.headers.tobytes() and np.array_equal(b.cell_ids, cached.cell_ids))
                del b, cached
            except Exception as e:
                BUILDER_INFO["exa_builder_gpu"] = {"error": repr(e)[:300]}
                log("GPU builder timing failed: %r" % e)
        t = time.time()
        bricks = R.ExaBricks.load(base + ".bricks")
        field = np.fromfile(base + ".scalars", np.float32)
        log("scene %s loaded from %s in %.1fs" % (name, base, time.time() - t))
        return bricks, field
    cells, field, _, _ = make_workload(name)
    t = time.time()
    if gpu_builder:
        bricks = gpu_build(cells)
    else:
        bricks = R.ExaBricks.build(cells)
        log("exa_bricks_build: %d cells -> %d bricks in %.1fs" % (len(cells), bricks.num_bricks, time.time() - t))
    del cells
    try:
        bricks.save(base + ".bricks")
        field.tofile(base + ".scalars")
        open(base + ".ok", "w").write("ok")
    except OSError as e:
        log("scene cache not written (%s)" % e)
    return bricks, field


def ncu_traffic(kernel, workload="c2"):
    """DRAM bytes per launch of `kernel` on `workload` from the newest committed ncu capture that has it
    (profiles/traffic_r02.json, else traffic_r01.json), or None"""
    for f in ("traffic_r02.json", "traffic_r01.json"):
        try:
            t = json.load(open(os.path.join(ROOT, "profiles", f)))
            t = t[workload][kernel] if workload in t else (t[kernel] if workload == "c2" else None)
            return int(t["dram_bytes_read"]) + int(t["dram_bytes_write"])
        except Exception:
            continue
    return None


def dram_frac(traffic, ms, peak):
    """measured DRAM bytes of one launch / the launch's duration here / HBM peak (None without an ncu capture)"""
    return None if not traffic or not ms else traffic / (ms * 1e-3) / 1e9 / peak


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.proc = index, [], False, None

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
                if self.stop_flag:
                    break
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def frame_desc(R, domain, image, field_range, iso=False):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import scenes
    W, H = image
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, dom3(domain), W, H))
    fd.opacity_scale = OPACITY_SCALE
    fd.dt = DT
    fd.xf_domain = [(float(field_range[0]), float(field_range[1]))]
    # the viewer's default transfer function: alpha ramp i/127 over colour map 0, "Paraview Cool Warm"
    # (exa/viewer.cpp:526-568: xfColorMap[c] = 0, xfColor[j][i] = colormaps[j](i/127))
    fd.xf = [scenes.xf_ramp(colors=R.ColorMapper.embedded(0).table())]
    if iso:                                  # C4: one implicit iso-surface at the XF-domain midpoint, channel 0, + DVR
        fd.iso = [(1, 0.5 * (float(field_range[0]) + float(field_range[1])), 0), (0, 0.0, 0)]
    return fd


def has_iso(name):
    return name.startswith("c4")


def algorithmic_bytes(st, W, H):
    """B_frame of BASELINE.md §2 / SURVEY.md §8(d)"""
    return 36 * st["brickVisits"] + 4 * st["cornerLoads"] + 44 * st["segments"] + 20 * W * H


# ---------------------------------------------------------------------------------------------
# reference arm: the reference's CPU implementation of the path (oracle port; the reference's
# device code needs OptiX and cannot be compiled here)
# ---------------------------------------------------------------------------------------------
def cpu_scene(orc, cells, field):
    raw = orc.build_bricks(cells)
    bricks, ids = orc.parse_bricks(raw)
    sc = orc.flatten_scalars(ids, [field])
    regions, leaf = orc.build_abr(bricks, sc)
    return orc.Scene(bricks, sc, regions, leaf)


def cpu_frame_sample(orc, scene, fd, window):
    t = time.time()
    _, _, st = scene.render(fd.to_orc(orc), window=window)
    return st, time.time() - t


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import orc
    from owlexabrick_b200 import renderer as R      # host-side camera helper only
    orc.build()
    name = args.workload
    domain, _, image = WORKLOADS[name]
    t = time.time()
    if name in ("c3", "c3s2", "c3s4", "c4"):
        # 260-650 M cells: the serial restatement of exaBuilder needs > 10 min, so the (untimed) scene
        # set-up takes the bricks from the OpenMP host builder, whose output is bit-exact with
        # it (tests/test_host.py); flatten, ABR build and the timed frames are the oracle's
        bricks, field = load_or_build_scene(name, R)
        hdr = bricks.headers.view(orc.BRICK_DT)
        sc = orc.flatten_scalars(bricks.cell_ids, [field])
        regions, leaf = orc.build_abr(hdr, sc)
        scene = orc.Scene(hdr, sc, regions, leaf)
    else:
        cells, field, _, _ = make_workload(name)
        log("reference arm: building the CPU scene (oracle builder, serial)")
        scene = cpu_scene(orc, cells, field)
        del cells
    log("cpu scene built in %.1fs" % (time.time() - t))
    W, H = image
    fd = frame_desc(R, domain, image, (float(field.min()), float(field.max())), iso=has_iso(name))
    # one step = the WHOLE frame on all host cores (the oracle's region query is BVH-accelerated,
    # oracle.cpp:closestRegion, so a C2 frame is a fraction of a second)
    window = (0, 0, W, H)
    for _ in range(args.warmup):
        cpu_frame_sample(orc, scene, fd, (W // 2 - 64, H // 2 - 64, W // 2 + 64, H // 2 + 64))
    tot_s, tot_t = 0, 0.0
    for _ in range(args.steps):
        st, dt = cpu_frame_sample(orc, scene, fd, window)
        tot_s += st["samples"]
        tot_t += dt
    value = tot_s / tot_t
    sample = "the whole %dx%d %s frame per step, %d steps, %.1f s of CPU time" % (W, H, name, args.steps, tot_t)
    line = {"impl": "reference", "metric": "dvr_reconstructed_samples_per_s", "value": value, "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(name), "image": list(image), "dt": DT},
            "cpu_baseline": {"value": value, "unit": "samples/s", "cores": orc.num_threads(), "kind": "port", "sample": sample, "pinned": "the port is bit-identical to the reference's own device source programs/exabrick.cu compiled as C++ (oracle/_ref/ref_shader, tests/test_oracle_pinned.py); that build is not timed here: its stand-in for the OptiX BVH is a brute-force loop over all regions"},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def reference_host_stages(name, bricks, field):
    """BASELINE.md rows B1/B2 beside the GPU numbers: wall time of the reference's OWN host stages (oracle/_ref, the
    reference's sources compiled in place) on this box: exaBuilder on config C1 (serial, as shipped) and
    ExaBrickRegions::buildFrom on the benchmarked bricks.  File I/O on tmpfs is inside their wall time."""
    import tempfile
    from owlexabrick_b200 import synth
    ref = os.path.join(ROOT, "oracle", "_ref")
    out = {}
    if not os.path.exists(os.path.join(ref, "exaBuilder")):
        return {"reference_host_stages": "oracle/_ref not built on this box"}
    with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as d:
        try:
            cells = synth.nested_boxes(256, [(72, 184), (32, 224)])
            cf, bf = os.path.join(d, "c1.cells"), os.path.join(d, "c1.bricks")
            np.ascontiguousarray(cells, np.int32).tofile(cf)
            t = time.time()
            subprocess.check_call([os.path.join(ref, "exaBuilder"), cf, "-o", bf], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=300)
            dt = time.time() - t
            out["reference_exaBuilder_c1"] = {"wall_s": round(dt, 3), "cells": int(len(cells)), "cells_per_s": round(len(cells) / dt), "threads": 1,
                                              "note": "the reference's builder/builder.cpp compiled in place, serial as shipped, incl. .cells/.bricks I/O on tmpfs"}
        except Exception as e:
            out["reference_exaBuilder_c1"] = {"error": str(e)[:200]}
        base = os.path.join(os.environ.get("EXA_SCENE_CACHE", "/dev/shm"), "exa_scene_%s_v%d" % (name, SCENE_CACHE_VERSION))
        if os.path.exists(base + ".bricks") and os.path.exists(base + ".scalars") and os.path.exists(os.path.join(ref, "ref_abr")):
            try:
                of = os.path.join(d, "w.abr")
                t = time.time()
                subprocess.check_call([os.path.join(ref, "ref_abr"), base + ".bricks", of, base + ".scalars"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=600)
                dt = time.time() - t
                nr = int(np.frombuffer(open(of, "rb").read(8), "<i4")[0])
                out["reference_buildFrom"] = {"wall_s": round(dt, 2), "regions": nr, "regions_per_s": round(nr / dt),
                                              "note": "ExaBrickRegions::buildFrom (exa/Regions.cpp) on the benchmarked bricks incl. ExaBricks::load and the scalar gather; serial build (-DOWL_DISABLE_TBB)"}
            except Exception as e:
                out["reference_buildFrom"] = {"error": str(e)[:200]}
    return out


def workload_name(name):
    domain, shape, image = WORKLOADS[name]
    if isinstance(shape, tuple):
        return "%s: synthetic exajet-shaped 4-level AMR (fuselage/wing/tail/wake boxes), domain %dx%dx%d, single scalar, DVR %dx%d, default camera/XF" % (
            (name,) + tuple(int(x) for x in domain) + (image[0], image[1]))
    if has_iso(name):
        return "%s: synthetic landing-gear-shaped %d-level nested AMR (off-centre core), domain %d^3, single scalar, implicit iso-surface (XF-domain midpoint) + DVR %dx%d, default camera/XF" % (
            name, len(shape) + 1, domain, image[0], image[1])
    return "%s: synthetic 3-level nested-box AMR, domain %d^3, single scalar, DVR %dx%d, default camera/XF" % (name, domain, image[0], image[1])


def measure_sortlast(args, torch, dist, capi, R, multigpu, r, sl_r, sl_pipe, bricks, field, fd, name, rank, world, local, W, H, stream,
                     compare_with_single, frame_via_pipeline):
    """sort-last sub-object of the N>1 line: data-partitioned renderer per rank, direct-send composite through peer memory"""
    if args.no_sortlast:
        return None
    own = sl_pipe is None
    try:
        if own:
            t = time.time()
            sl_r = R.Renderer(bricks, [field], device=local, part=(rank, world))
            sl_r.set_stream(stream.cuda_stream)
            fd.apply(sl_r)
            sl_pipe = multigpu.SortLastPipeline(dist, torch, capi, np, sl_r, rank, world, local, W, H, stream, nccl_barrier=args.nccl_barrier,
                                                surf=has_iso(name), tag="%s_sl" % os.environ.get("MASTER_PORT", "0"))
            log("rank %d: sort-last renderer in %.1fs, resident %s of %d bricks" % (rank, time.time() - t, sl_r.residency(), bricks.num_bricks))
        nb, ncell = sl_r.residency()
        res = torch.tensor([nb, ncell, sl_r.scalar_layout()[1]], dtype=torch.int64, device="cuda")
        allres = [torch.zeros(3, dtype=torch.int64, device="cuda") for _ in range(world)]
        dist.all_gather(allres, res)
        out = {"mode": "data partitioned along the ABR kd-tree (balanced by estimated lattice samples), float4 partials in peer-mapped buffers, "
                       "one barrier, fused composite+finish kernel reading all ranks' partials over NVLink (direct send), triple-buffered",
               "resident_bricks_per_rank": [int(x[0]) for x in allres], "resident_cells_per_rank": [int(x[1]) for x in allres],
               "scalar_bytes_per_rank": [int(x[2]) for x in allres], "bricks_total": bricks.num_bricks, "cells_total": int(bricks.total_num_cells),
               "peer_bytes_read_per_rank_per_frame": sl_pipe.peer_read_bytes, "barrier": "NCCL all-reduce" if sl_pipe.bar.nccl else "peer memory"}
        alpha, rgb = fd.xf[0]
        for label, scale in (("opaque", fd.opacity_scale), ("translucent", 0.02)):
            sl_r.updateXF(0, alpha, rgb, fd.xf_domain[0], scale)
            st = sl_r.render_stats()
            stt = torch.tensor([st["samples"]], dtype=torch.int64, device="cuda"); dist.all_reduce(stt)
            for _ in range(3):
                sl_pipe.step()
            sl_pipe.drain(); torch.cuda.synchronize(); dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            for _ in range(args.steps):
                sl_pipe.step()
            sl_pipe.drain(); b.record(stream); torch.cuda.synchronize()
            ms = torch.tensor([a.elapsed_time(b) / args.steps], dtype=torch.float64, device="cuda"); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            # end to end: every rank's span to the shared host frame
            dist.barrier(); t = time.perf_counter()
            for _ in range(args.steps):
                sl_r.updateCamera(*fd.cam)
                sl_pipe.step(readback=True)
            sl_pipe.drain(); sl_pipe.finish(); torch.cuda.synchronize()
            te = torch.tensor([time.perf_counter() - t], dtype=torch.float64, device="cuda"); dist.all_reduce(te, op=dist.ReduceOp.MAX)
            ent = {"ms_per_frame": float(ms.item()), "frames_per_s": 1e3 / float(ms.item()), "samples_per_frame": int(stt.item()),
                   "samples_per_s": int(stt.item()) / (float(ms.item()) * 1e-3), "e2e_frames_per_s": args.steps / float(te.item())}
            # deviation from the single-GPU frame: with the reference's termination threshold, and with it off
            for ert_label, ert in (("ert_0.98", 0.98), ("ert_off", 2.0)):
                sl_r.setErtThreshold(ert)
                img_n = frame_via_pipeline(sl_pipe)
                if rank == 0:
                    r.updateXF(0, alpha, rgb, fd.xf_domain[0], scale)
                    ent["vs_1gpu_" + ert_label] = compare_with_single(img_n, ert=ert)
                dist.barrier()
            sl_r.setErtThreshold(0.98)
            out[label] = ent
        if rank == 0:
            r.updateXF(0, alpha, rgb, fd.xf_domain[0], fd.opacity_scale)
            out["barrier_timed_out"] = sl_pipe.bar.timed_out()
            log("sort-last:", json.dumps(out))
        if own:
            sl_pipe.close()
            sl_r.close()
        return out
    except Exception as e:                                            # the secondary measurement must not take the headline line down
        log("rank %d: sort-last measurement failed: %r" % (rank, e))
        return {"error": repr(e)[:300]}


def run_product(args):
    import torch
    import torch.distributed as dist
    from owlexabrick_b200 import build
    build.build_lib()
    from owlexabrick_b200 import capi, renderer as R, synth
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        log("WORLD_SIZE %d != --gpus %d; using WORLD_SIZE" % (world, args.gpus))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    assert capi.lib.exa_cuda_available() == 1, capi.last_error()

    name = args.workload
    domain, boxes, image = WORKLOADS[name]
    W, H = image
    # ---- setup (untimed): rank 0 generates + builds, the others load from /dev/shm ----
    t0 = time.time()
    if rank == 0:
        bricks, field = load_or_build_scene(name, R, gpu_builder=True, time_builder=not args.no_builder_timing)
    if world > 1:
        dist.barrier()
        if rank != 0:
            bricks, field = load_or_build_scene(name, R)              # from rank 0's /dev/shm cache
    ncells = bricks.total_num_cells
    t = time.time()
    r = R.Renderer(bricks, [field], device=local, layout=capi.LAYOUT_FLAT if args.flat_layout else None)
    torch.cuda.synchronize()
    regions, leaf = r.get_abr()
    create_s = time.time() - t
    bt = r.build_times()
    build_info = {**BUILDER_INFO, "renderer_create_wall_s": round(create_s, 3), **{k: round(v, 3) for k, v in bt.items()},
                  "abr_build_ms": round(bt["abr_geometry_ms"] + bt["value_ranges_ms"], 3), "abr_regions": int(len(regions)),
                  "abr_regions_per_s": round(len(regions) / max(1e-9, 1e-3 * (bt["abr_geometry_ms"] + bt["value_ranges_ms"]))),
                  "note": "device time (CUDA events): ABR geometry = kd build + regions + leaf lists (ExaBrickRegions::buildFrom, exa/Regions.cpp:242-287), value ranges = Regions.cpp:182-240; the reference's buildFrom on the same bricks is timed in cpu_baseline.reference_buildFrom"}
    log("rank %d: renderer created in %.1fs (%d regions, L=%d, kd %s) %s" % (rank, create_s, len(regions), len(leaf), r.kd_info(), build_info))
    stream = torch.cuda.current_stream()
    r.set_stream(stream.cuda_stream)
    fd = frame_desc(R, domain, image, (float(field.min()), float(field.max())), iso=has_iso(name))
    fd.apply(r)
    npix = W * H
    mode = "none" if world == 1 else args.partition
    pipe, sl_r = None, None
    if world > 1:
        from owlexabrick_b200 import multigpu
        tag = "%s_%s" % (os.environ.get("MASTER_PORT", "0"), mode)
        if mode == "tiles":
            # data replicated; the frame is N row strips, strip k on GPU k: every rank's frame kernel stores its tiles into
            # the strips over NVLink (peer pointers), a peer-memory barrier orders "frame complete", and rank k reads strip k
            # back over its own PCIe link into a pinned host frame shared by the ranks (owlexabrick_b200/multigpu.py)
            pipe = multigpu.TilePipeline(dist, torch, capi, np, r, rank, world, local, W, H, stream, nccl_barrier=args.nccl_barrier, tag=tag, weighted=not args.uniform_strips)
            log("rank %d: strips %s rows, copy bandwidth per rank %s GB/s" % (rank, pipe.rows, [round(x, 1) for x in pipe.copy_gbs]))
        else:
            # sort-last with the DATA partitioned: a second renderer that holds only this rank's kd subtree
            sl_r = R.Renderer(bricks, [field], device=local, part=(rank, world))
            sl_r.set_stream(stream.cuda_stream)
            fd.apply(sl_r)
            pipe = multigpu.SortLastPipeline(dist, torch, capi, np, sl_r, rank, world, local, W, H, stream, nccl_barrier=args.nccl_barrier, surf=has_iso(name), tag=tag)
        log("rank %d: %s pipeline, barrier: %s" % (rank, mode, "NCCL all-reduce" if pipe.bar.nccl else "peer memory"))

    def drain():
        """main stream waits for the completion barriers still in flight on the side stream"""
        if pipe is not None:
            pipe.drain()

    overlap = world > 1 and mode == "tiles" and not args.static_tiles and not args.no_frame_overlap

    def step(readback=False, overlapped=False):
        if pipe is None:
            r.render(sync=False)
            return 1
        if overlapped and overlap:
            pipe.step(readback, overlap=True)                     # frames on two alternating streams (no device-state update in between)
        else:
            pipe.step(readback)
        return 2 + (1 if mode == "sortlast" else 0)              # frame kernel + barrier kernel (+ composite kernel)

    rr = sl_r if sl_r is not None else r                          # the renderer the timed frames run on
    # instrumented, untimed: counters of the algorithmic-bytes model
    st = rr.render_stats() if sl_r is None else sl_r.render_stats()
    stats_t = torch.tensor([st["samples"], st["brickVisits"], st["cornerLoads"], st["segments"]], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(stats_t)
    tot = dict(zip(["samples", "brickVisits", "cornerLoads", "segments"], [int(x) for x in stats_t.tolist()]))
    sampler = ClockSampler(local)          # samples clocks from the warm-up to the end of the e2e loop
    if rank == 0:
        sampler.start()
    if args.profile_range:                  # ncu --profile-from-start off: launch list of warm-up + timed steps only
        torch.cuda.profiler.start()
    for _ in range(args.warmup):
        step()
    balance = None
    if world > 1 and mode == "tiles" and not args.static_tiles:
        # tiles dealt by measured cost (part of the warm-up, as a viewer would do every few dozen frames)
        balance = {"max_over_mean_rank_load_after_lpt": pipe.rebalance()}
        for _ in range(3):
            step(overlapped=True)
        drain()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    l0 = rr.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record(stream)
    launches = 0
    for _ in range(args.steps):
        launches += step(overlapped=True)
    drain()
    ev1.record(stream)
    torch.cuda.synchronize()
    if args.profile_range:
        torch.cuda.profiler.stop()
    if world > 1:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    tms = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms = float(tms.item())
    lib_launches = rr.launch_count() - l0
    ms_step = ms / args.steps
    value = tot["samples"] / (ms_step * 1e-3)

    # ---- end to end through the C ABI with host buffers (N=1 path; at N>1 rank 0's read-back) ----
    host_img = torch.empty((H, W), dtype=torch.int32).pin_memory()
    host_ring = [host_img, torch.empty((H, W), dtype=torch.int32).pin_memory()]
    alpha, rgb = fd.xf[0]
    e2e_count = [0]
    def e2e_step():
        rr.updateCamera(*fd.cam)
        rr.updateXF(0, alpha, rgb, fd.xf_domain[0], fd.opacity_scale)      # host -> device: 128 float4 (stream-ordered, pinned ring)
        if world == 1:
            # every frame is rendered from this step's inputs and read back to (pinned) host memory;
            # the copy of frame i runs on the copy stream while frame i+1 renders into the second buffer
            r.render(sync=False)
            capi.check(capi.lib.exa_renderer_read_rgba8_async(r.h, host_ring[e2e_count[0] & 1].data_ptr()))   # device -> host: W*H*4
            e2e_count[0] += 1
        else:
            step(readback=True)                                       # every rank copies ITS strip / span of frame i to the shared pinned host frame while frame i+1 renders
    for _ in range(2):
        e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    drain()
    r.sync()                                                         # render stream + copy stream: the last frame is in host memory
    if pipe is not None:
        pipe.finish()                                                # this rank's strips / spans have landed in the shared host frames
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())
    e2e_value = tot["samples"] / (e2e_s / args.steps)
    # what bounds e2e: the frame's device->host copy alone (pinned, device-timed)
    d2h_ms = None
    if rank == 0:
        dev_img = torch.zeros((H, W), dtype=torch.int32, device="cuda")
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        host_img.copy_(dev_img, non_blocking=True); torch.cuda.synchronize()
        c0.record(stream)
        for _ in range(10):
            host_img.copy_(dev_img, non_blocking=True)
        c1.record(stream); torch.cuda.synchronize()
        d2h_ms = c0.elapsed_time(c1) / 10
        del dev_img
    # N>1: what bounds e2e on each rank — its own strip / span copy alone (side stream, into the shared host frame), and
    # the e2e loop without the copies (host-side issue rate + barrier chain)
    e2e_diag = None
    if world > 1 and pipe is not None and pipe.d2h_bytes:
        import ctypes as C
        src = pipe.strips.ptr[rank][0] if hasattr(pipe, "strips") else pipe.out.data_ptr() + pipe.lo * 4
        dst_off = (pipe.y0 * W * 4) if hasattr(pipe, "strips") else pipe.lo * 4
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); dist.barrier()
        c0.record(pipe.side)
        for _ in range(10):
            capi.check(capi.lib.exa_copy_async(C.c_void_p(pipe.host.ptr(0, dst_off)), C.c_void_p(src), pipe.d2h_bytes, C.c_void_p(pipe.side.cuda_stream)))
        c1.record(pipe.side); torch.cuda.synchronize()
        tcopy = torch.tensor([c0.elapsed_time(c1) / 10], dtype=torch.float64, device="cuda")
        all_copy = [torch.zeros(1, dtype=torch.float64, device="cuda") for _ in range(world)]
        dist.all_gather(all_copy, tcopy)
        dist.barrier()
        t = time.perf_counter()
        for _ in range(args.steps):
            rr.updateCamera(*fd.cam)
            rr.updateXF(0, alpha, rgb, fd.xf_domain[0], fd.opacity_scale)
            step(readback=False)
        drain(); torch.cuda.synchronize()
        tn = torch.tensor([time.perf_counter() - t], dtype=torch.float64, device="cuda")
        dist.all_reduce(tn, op=dist.ReduceOp.MAX)
        # the same e2e loop at 5x the steps, with the host's enqueue time apart: separates the per-frame rate from the
        # fixed cost of filling and draining the pipeline (the contract's K = 20 frames of ~0.1-0.3 ms is a short run)
        dist.barrier()
        t = time.perf_counter()
        for _ in range(5 * args.steps):
            e2e_step()
        t_enq = time.perf_counter() - t
        drain(); r.sync(); pipe.finish(); torch.cuda.synchronize()
        tl = torch.tensor([time.perf_counter() - t, t_enq], dtype=torch.float64, device="cuda")
        dist.all_reduce(tl, op=dist.ReduceOp.MAX)
        e2e_diag = {"own_copy_alone_ms_per_rank": [round(float(x.item()), 4) for x in all_copy], "copy_bytes": pipe.d2h_bytes,
                    "e2e_without_copies_frames_per_s": args.steps / float(tn.item()),
                    "e2e_frames_per_s_at_5x_steps": 5 * args.steps / float(tl[0].item()), "host_enqueue_ms_per_frame_at_5x_steps": float(tl[1].item()) / (5 * args.steps) * 1e3}
        log("rank %d: e2e diagnostics %s" % (rank, e2e_diag))
    # keep the GPU(s) under the same load until nvidia-smi has produced a few samples
    t = time.time()
    while True:
        need = 1 if (rank == 0 and len(sampler.rows) < 5 and time.time() - t < 3.0) else 0
        if world > 1:
            nt = torch.tensor([need], device="cuda"); dist.broadcast(nt, 0); need = int(nt.item())
        if not need:
            break
        for _ in range(50):
            step()
        torch.cuda.synchronize()
    clocks = sampler.finish() if rank == 0 else None
    if world > 1:
        dist.barrier()

    # ---- secondary measurement: the same frame with a translucent XF (opacity scale 0.02), which
    # makes rays traverse the whole volume instead of terminating after a few samples ----
    # ---- N>1 diagnostics: this rank's frame kernel alone (no exchange), max over ranks; the barrier alone ----
    render_only_ms, barrier_ms = None, None
    if world > 1:
        def alone():
            if mode == "tiles":
                r.render(sync=False)
            else:
                sl_r.render_partial(pipe.partials.ptr[rank][0], pipe.surfs.ptr[rank][0] if pipe.surfs else None)
        for _ in range(3):
            alone()
        torch.cuda.synchronize(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(args.steps):
            alone()
        b.record(stream); torch.cuda.synchronize()
        mine_ms = a.elapsed_time(b) / args.steps
        ro = torch.tensor([mine_ms], dtype=torch.float64, device="cuda")
        allr = [torch.zeros(1, dtype=torch.float64, device="cuda") for _ in range(world)]
        dist.all_gather(allr, ro)
        per_rank = [float(x.item()) for x in allr]
        render_only_ms = max(per_rank)
        dist.barrier()
        a.record(stream)
        for _ in range(args.steps):
            pipe.bar.wait(stream)
        b.record(stream); torch.cuda.synchronize()
        barrier_ms = a.elapsed_time(b) / args.steps
        log("rank %d: own frame kernel %.4f ms/step; per rank %s; barrier alone %.4f ms (%s)" % (
            rank, mine_ms, ["%.4f" % x for x in per_rank], barrier_ms, "NCCL all-reduce" if pipe.bar.nccl else "peer memory"))
        dist.barrier()

    # ---- multi-GPU check (untimed): the N-GPU frame as it arrives in the shared host frame (every rank's strip / span
    # copied over its own PCIe link) against the same frame rendered by rank 0 alone ----
    mg_check = None

    def frame_via_pipeline(p):
        """one frame through pipeline p with read-back -> the host frame (valid on every rank after the barrier)"""
        p.drain(); torch.cuda.synchronize(); dist.barrier()
        bsel = p.j % 3
        if rank == 0:
            for k in range(3):
                p.host_frame(k)[:] = 0                                # stale pixels must not be able to hide a missing tile
        dist.barrier()
        if overlap and p is pipe:
            # the checked frame is the LAST of three overlapped ones: it ran while its predecessor was still finishing
            for _ in range(3):
                bsel = p.j % 3
                p.step(readback=True, overlap=True)
            p.drain(); p.finish(); torch.cuda.synchronize(); dist.barrier()
            return p.host_frame(bsel).reshape(-1).copy()
        p.step(readback=True); p.drain(); p.finish(); torch.cuda.synchronize(); dist.barrier()
        return p.host_frame(bsel).reshape(-1).copy()

    def compare_with_single(img_n, ert=None):
        """rank 0: the same frame on the replicated renderer alone"""
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import scenes
        r.set_partition(capi.PARTITION_NONE, 0, 1)
        capi.check(capi.lib.exa_renderer_set_color_strips(r.h, None, 0))
        if ert is not None:
            r.setErtThreshold(ert)
        r.render()
        img_1 = r.read_rgba8().reshape(-1)
        if ert is not None:
            r.setErtThreshold(0.98)
        dif = np.abs(scenes.unpack_rgba8(img_n) - scenes.unpack_rgba8(img_1))
        return {"max_abs_rgba8_diff_vs_1gpu": int(dif.max()), "pixels_differing": int((dif.max(axis=-1) > 0).sum())}

    if world > 1:
        img_n = frame_via_pipeline(pipe)
        if rank == 0:
            mg_check = compare_with_single(img_n)
            mg_check["note"] = "image tiles: bit-identical expected; sort-last: per-rank early termination (H1) bounds the deviation"
            mg_check["barrier_timed_out"] = pipe.bar.timed_out()
            log("multi-GPU check:", mg_check)
        if mode == "tiles":                                           # rank 0 left tile mode for the comparison: back to the static deal on EVERY rank (consistent)
            r.set_partition(capi.PARTITION_IMAGE_TILES, rank, world)
        dist.barrier()

    # ---- sort-last (north_star's primary multi-GPU mode), measured inside the default --gpus N run: data partitioned
    # along the ABR kd-tree, partials composited through peer memory; opaque and translucent frame, deviation vs 1 GPU ----
    sortlast = None
    if world > 1:
        sortlast = measure_sortlast(args, torch, dist, capi, R, multigpu, r, sl_r, pipe if mode == "sortlast" else None, bricks, field, fd, name,
                                    rank, world, local, W, H, stream, compare_with_single, frame_via_pipeline)

    translucent = None
    if world == 1:
        r.updateXF(0, alpha, rgb, fd.xf_domain[0], 0.02)
        tst = r.render_stats()
        for _ in range(3):
            r.render(sync=False)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(args.steps):
            r.render(sync=False)
        b.record(stream)
        torch.cuda.synchronize()
        tms_ = a.elapsed_time(b) / args.steps
        tb = algorithmic_bytes(tst, W, H)
        translucent = {"xf_opacity_scale": 0.02, "ms_per_frame": tms_, "samples_per_frame": tst["samples"], "segments_per_frame": tst["segments"],
                       "samples_per_s": tst["samples"] / (tms_ * 1e-3), "algorithmic_gbs": tb / (tms_ * 1e-3) / 1e9,
                       "frac": tb / (tms_ * 1e-3) / 1e9 / peaks()[0], "frac_kind": "algorithmic bytes / HBM peak", "limiter": "issue",
                       "traffic": ncu_traffic("k_renderFrame_translucent", name),
                       "dram_frac": dram_frac(ncu_traffic("k_renderFrame_translucent", name), tms_, peaks()[0])}
        r.updateXF(0, alpha, rgb, fd.xf_domain[0], fd.opacity_scale)

    # ---- reconstruction sampling (K4): 1 M random points, device resident, + host-buffer call ----
    recon = None
    if rank == 0:
        npts = 1 << 20
        pts = synth.random_points(npts, [0] * 3, dom3(domain), seed=1)
        d_pts = torch.from_numpy(pts).cuda()
        d_val = torch.empty(npts, dtype=torch.float32, device="cuda")
        d_ok = torch.empty(npts, dtype=torch.uint8, device="cuda")
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        flush2 = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        pst = r.sample_points_stats(d_pts.data_ptr(), npts, 0)
        times = []
        for i in range(args.warmup + max(args.steps, 5)):
            flush.fill_(i & 0xff)                                            # L2 flush between iterations: write 256 MiB,
            flush2.sum()                                                     # then read 256 MiB so that L2 holds clean lines (no write-back inside the timed kernel)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            r.sample_points_device(d_pts.data_ptr(), npts, 0, d_val.data_ptr(), d_ok.data_ptr())
            b.record(stream)
            torch.cuda.synchronize()
            if i >= args.warmup:
                times.append(a.elapsed_time(b))
        pms = float(np.mean(times))
        pbytes = 16 * npts + 44 * pst["segments"] + 36 * pst["brickVisits"] + 4 * pst["cornerLoads"]
        hbuf = (np.zeros(npts, np.float32), np.zeros(npts, np.uint8))  # the caller's result arrays, reused from call to call
        r.sample_points(pts, 0, out=hbuf)                             # first call allocates the staging buffers
        t = time.perf_counter()
        for _ in range(5):
            r.sample_points(pts, 0, out=hbuf)
        h_s = (time.perf_counter() - t) / 5
        peak, _ = peaks()
        recon = {"points": npts, "ms": pms, "samples_per_s": npts / (pms * 1e-3), "algorithmic_bytes": pbytes,
                 "achieved_gbs": pbytes / (pms * 1e-3) / 1e9, "frac": pbytes / (pms * 1e-3) / 1e9 / peak,
                 "frac_kind": "algorithmic bytes / HBM peak", "traffic": ncu_traffic("k_samplePoints", name),
                 "dram_frac": dram_frac(ncu_traffic("k_samplePoints", name), pms, peak), "e2e_samples_per_s": npts / h_s, "l2_flush": "256 MiB write + 256 MiB read between iterations"}

    # ---- CPU baseline + parity (rank 0, N=1 only): the oracle builds ITS OWN scene from the same bricks
    # (flatten, ABR build), renders the benchmarked frames and samples the benchmarked points; the GPU results
    # are compared with it here, on the full-size workload ----
    cpu, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import orc
        orc.build()
        t = time.time()
        hdr = bricks.headers.view(orc.BRICK_DT)
        sc = orc.flatten_scalars(bricks.cell_ids, [field])
        t_fl = time.time() - t
        t = time.time()
        o_regions, o_leaf = orc.build_abr(hdr, sc)                   # the oracle's own ABRs (not the product's)
        t_oabr = time.time() - t
        oscene = orc.Scene(hdr, sc, o_regions, o_leaf)
        log("cpu baseline scene: flatten %.1fs, oracle ABR build %.1fs (%d regions)" % (t_fl, t_oabr, len(o_regions)))
        parity = {"abr_bit_exact": bool(regions.tobytes() == o_regions.tobytes() and leaf.tobytes() == o_leaf.tobytes()),
                  "abr_regions": int(len(o_regions)), "abr_leaf_list": int(len(o_leaf)),
                  "against": "oracle port with its own flatten + ABR build over the same bricks (bit-identical to the reference's device source on the pinned set, tests/test_oracle_pinned.py)"}
        # the timed frame (opaque XF): GPU accumulation buffer vs the oracle's
        r.updateXF(0, alpha, rgb, fd.xf_domain[0], fd.opacity_scale)
        r.render()
        g_acc = r.read_accum()
        t = time.time()
        o_acc, _, cst = oscene.render(fd.to_orc(orc), window=(0, 0, W, H))
        cdt = time.time() - t
        def frame_parity(g, o):
            e = np.abs(g[..., :3] - o[..., :3]).max(axis=-1)
            worst = np.argsort(e.reshape(-1))[-3:][::-1]
            return {"rgba_max_abs": float(e.max()), "pixels_over_1e-3": int((e > 1e-3).sum()), "pixels_over_1e-2": int((e > 1e-2).sum()), "pixels": int(e.size),
                    "p99.99_abs": float(np.quantile(e, 0.9999)),
                    "worst": [{"xy": [int(i % W), int(i // W)], "gpu": [round(float(v), 5) for v in g.reshape(-1, 4)[i, :3]],
                               "oracle": [round(float(v), 5) for v in o.reshape(-1, 4)[i, :3]]} for i in worst]}
        parity["frame"] = frame_parity(g_acc, o_acc)
        parity["frame"]["samples_gpu_vs_oracle"] = [tot["samples"], int(cst["samples"])]
        reps = int(min(200, max(1, round(12.0 / max(cdt, 1e-3)))))  # aim at ~12 s of CPU work
        tot_s, tot_t = int(cst["samples"]), cdt
        for _ in range(reps - 1):
            cst, cdt = cpu_frame_sample(orc, oscene, fd, (0, 0, W, H))
            tot_s += cst["samples"]; tot_t += cdt
        # the translucent frame (rays cross the whole volume)
        if translucent is not None and not args.quick_parity:
            import copy
            fd2 = copy.copy(fd); fd2.opacity_scale = 0.02
            r.updateXF(0, alpha, rgb, fd.xf_domain[0], 0.02)
            r.render()
            g2 = r.read_accum()
            t = time.time()
            o2, _, cst2 = oscene.render(fd2.to_orc(orc), window=(0, 0, W, H))
            parity["frame_translucent"] = frame_parity(g2, o2)
            parity["frame_translucent"]["cpu_s"] = round(time.time() - t, 1)
            translucent["cpu_samples_per_s"] = cst2["samples"] / (time.time() - t)
            r.updateXF(0, alpha, rgb, fd.xf_domain[0], fd.opacity_scale)
        # reconstruction at the benchmarked 1 M points
        if recon is not None:
            t = time.time()
            ov, ook, _ = oscene.sample(pts, 0)
            t_pts = time.time() - t
            gv, gok = d_val.cpu().numpy(), d_ok.cpu().numpy()
            m = ook == 1
            fmax = float(np.abs(field).max())
            rel = np.abs(gv[m].astype(np.float64) - ov[m]) / np.maximum(np.abs(ov[m]), 1e-3 * fmax)
            parity["recon"] = {"points": int(len(pts)), "valid": int(m.sum()), "validity_mismatches": int((gok != ook).sum()),
                               "max_rel_err": float(rel.max()) if m.any() else 0.0, "gate": 1e-5}
            recon["cpu_samples_per_s"] = len(pts) / t_pts
        cpu = {"value": tot_s / tot_t, "unit": "samples/s", "frames_per_s": reps / tot_t, "cores": orc.num_threads(), "kind": "port", "pinned": "the port is bit-identical to the reference's own device source programs/exabrick.cu compiled as C++ (oracle/_ref/ref_shader, tests/test_oracle_pinned.py); that build is not timed here: its stand-in for the OptiX BVH is a brute-force loop over all regions",
               "sample": "the same whole frame, %d times, %.1f s (oracle port, OpenMP over pixel rows, BVH region query)" % (reps, tot_t),
               "abr_build": {"oracle_port_s": round(t_oabr, 2), "regions": int(len(o_regions))}}
        cpu.update(reference_host_stages(name, bricks, field))
        log("parity:", json.dumps(parity))

    if rank == 0:
        peak, peak_src = peaks()
        bframe = algorithmic_bytes(tot, W, H)
        achieved = bframe / (ms_step * 1e-3) / 1e9 / world                  # per GPU
        line = {"metric": "dvr_reconstructed_samples_per_s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "frames_per_s": 1e3 / ms_step, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": workload_name(name), "cells": int(ncells), "bricks": bricks.num_bricks, "regions": int(len(regions)),
                           "image": [W, H], "dt": DT, "xf": "alpha ramp i/127 over colour map 0 (Paraview Cool Warm), opacity scale %.2f" % OPACITY_SCALE,
                           "parallelism": "1 GPU" if world == 1 else ("sort-last x%d: data partitioned along the ABR kd-tree, float4 partials composited through NVLink peer memory inside the composite kernel (direct send), %s barrier per frame" % (world, "NCCL" if pipe.bar.nccl else "peer-memory") if mode == "sortlast" else
                                           "image tiles x%d (data replicated): RGBA8 tiles stored into N peer-mapped row strips (strip k on GPU k) over NVLink, %s barrier per frame, triple-buffered" % (world, "NCCL" if pipe.bar.nccl else "peer-memory")),
                           "scalar_layout": {1: "flat", 2: "dual-cell records"}[r.scalar_layout()[0]], "scalar_bytes": r.scalar_layout()[1],
                           "l2": "inputs larger than L2 (%.0f MB of scalars vs 126 MB), no flush" % (4e-6 * ncells),
                           "samples_per_frame": tot["samples"], "segments_per_frame": tot["segments"],
                           "step": "eager launches" + (", consecutive frames on two alternating streams: the tail of frame i (its last long-ray tiles) overlaps the head of frame i+1; no device state changes between the timed frames" if overlap else "") + (", frames triple-buffered: the completion barrier (and the composite) of frame i runs on a side stream while frame i+1 renders" if world > 1 else "")},
                "e2e": {"value": e2e_value, "unit": "samples/s", "frames_per_s": args.steps / e2e_s,
                        "h2d_bytes_per_step": (128 * 16 + 48) * world, "d2h_bytes_per_step": W * H * 4,
                        "d2h_bytes_per_step_per_rank": (W * H * 4) if world == 1 else pipe.d2h_bytes,
                        "d2h_copy_alone_ms": d2h_ms, "diagnostics": e2e_diag,
                        "strip_rows": getattr(pipe, "rows", None), "d2h_gbs_per_rank_all_copying": [round(x, 1) for x in pipe.copy_gbs] if hasattr(pipe, "copy_gbs") else None, "pipeline": "frame i+1 renders while frame i is copied to pinned host memory" + (" (exa_renderer_read_rgba8_async)" if world == 1 else
                                     ": every rank copies its own strip / span over its own PCIe link into one pinned host frame shared by the ranks (POSIX shared memory, scan-line layout)")},
                "gpu_launches": int(lib_launches), "clocks": clocks,
                "roofline": {"bound": "hbm", "kernel": "k_renderFrame", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": ncu_traffic("k_renderFrame", name) if world == 1 else None, "algorithmic_bytes_per_launch": bframe // world, "peak_source": peak_src,
                             "frac_kind": "algorithmic bytes / HBM peak", "limiter": "issue",
                             "dram_frac": dram_frac(ncu_traffic("k_renderFrame", name), ms_step, peak) if world == 1 else None,
                             "note": "achieved / frac are ALGORITHMIC (SURVEY 8d: every brick-header / scalar load of addBasisFunctions counted, divided by the HBM peak, as the bench contract defines them); almost all of those loads are served by L1/L2. dram_frac = measured DRAM bytes of one launch (traffic, ncu) / launch time / peak: the frame kernel is bound by instruction issue (limiter), not by HBM"},
                "dvr_translucent": translucent, "recon": recon, "multi_gpu_check": mg_check, "frame_kernel_only_ms_per_step": render_only_ms,
                "barrier_alone_ms": barrier_ms, "tile_balance": balance, "sortlast": sortlast, "cpu_baseline": cpu,
                "parity": parity, "build": build_info}
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        if pipe is not None:
            pipe.close()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="product", choices=["product", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--flat-layout", action="store_true", help="keep the scalars in the reference's flat layout (4 B per cell) instead of the dual-cell records (32 B per cell): what a scene too large for the records runs on")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-builder-timing", action="store_true", help="do not re-run the GPU exaBuilder for its timing when the scene cache exists")
    ap.add_argument("--quick-parity", action="store_true", help="skip the CPU render of the translucent frame (about 30 s on C3)")
    ap.add_argument("--profile-range", action="store_true", help="cudaProfilerStart/Stop around warm-up + timed steps (for ncu --profile-from-start off)")
    ap.add_argument("--nccl-barrier", action="store_true", help="N>1: 4-byte NCCL all-reduce per frame instead of the peer-memory barrier kernel")
    ap.add_argument("--no-frame-overlap", action="store_true", help="N>1 image tiles: every frame on the one render stream (no overlap of a frame's tail with the next frame's head)")
    ap.add_argument("--uniform-strips", action="store_true", help="N>1 image tiles: equal frame strips instead of strips sized by each GPU's measured copy bandwidth")
    ap.add_argument("--static-tiles", action="store_true", help="N>1 image tiles: keep the static round-robin deal (no cost-balanced tile lists)")
    ap.add_argument("--no-sortlast", action="store_true", help="N>1: skip the secondary sort-last measurement")
    ap.add_argument("--partition", default="tiles", choices=["tiles", "sortlast"], help="multi-GPU mode (N>1)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "product" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_product(args)


if __name__ == "__main__":
    main()
