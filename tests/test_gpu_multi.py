"""GPU tests of the multi-GPU pieces, emulated on ONE device (run with -m gpu): N renderers stand for the N
ranks.  Sort-last with the data partitioned (exa_renderer_create_partitioned) + the fused composite kernel
against the single-renderer frame; frame strips in image-tile mode; the tracer-shape fix."""
import ctypes as C

import numpy as np
import pytest

from owlexabrick_b200 import multigpu, synth
import scenes

pytestmark = pytest.mark.gpu

RGBA_TOL = 1e-3


@pytest.fixture(scope="module")
def R(built_lib):
    assert built_lib.lib.exa_cuda_available() == 1, built_lib.last_error()
    from owlexabrick_b200 import renderer
    return renderer


def _scene(R, kind="nb64_off", channels=1, mlw=8):
    cells, fields, dom = scenes.small_scene(kind, channels)
    b = R.ExaBricks.build(cells, max_leaf_width=mlw)
    return b, fields, dom


def _sort_last_frame(R, torch, b, fields, fd, world, data_partitioned, surf=False):
    """-> (accum float32[H,W,4], rgba8 uint32[H,W], residency list) of the composited frame"""
    W, H = fd.width, fd.height
    npix = W * H
    parts, surfs, rs = [], [], []
    for k in range(world):
        if data_partitioned:
            r = R.Renderer(b, fields, part=(k, world))
        else:
            r = R.Renderer(b, fields)
        fd.apply(r)
        if not data_partitioned:
            from owlexabrick_b200 import capi
            r.set_partition(capi.PARTITION_SORT_LAST, k, world)
        # NaN-filled: a part renders only the tile rectangle of its box, and the composite must not read the rest
        p = torch.full((npix, 4), float("nan"), dtype=torch.float32, device="cuda")
        sv = torch.full((npix, 4), float("nan"), dtype=torch.float32, device="cuda") if surf else None
        r.render_partial(p.data_ptr(), sv.data_ptr() if surf else None)
        r.sync()
        parts.append(p); surfs.append(sv); rs.append(r)
    order, nparts = rs[0].partition_order()
    assert sorted(order) == list(range(nparts)) and nparts <= world
    for r in rs[1:]:
        assert r.partition_order()[0] == order                      # every rank derives the same order
    out = torch.zeros(npix, dtype=torch.int32, device="cuda")
    acc = np.zeros((H, W, 4), np.float32)
    for k in range(world):                                           # rank k composites span k
        lo, hi = multigpu.span_of(k, world, npix)
        rs[k].composite_partials([parts[j].data_ptr() for j in order], [surfs[j].data_ptr() for j in order] if surf else None, lo, hi - lo, out.data_ptr())
        rs[k].sync()
        a = rs[k].read_accum().reshape(-1, 4)
        acc.reshape(-1, 4)[lo:hi] = a[lo:hi]
    res = [r.residency() for r in rs]
    assert np.isfinite(acc).all()
    host_parts = [p.cpu().numpy() for p in parts]
    unwritten = [float(np.isnan(h[:, 3]).mean()) for h in host_parts]
    if world >= 4:
        assert max(unwritten) > 0.2, unwritten                      # the rectangles do restrict the work
    return acc, out.cpu().numpy().view(np.uint32).reshape(H, W), res, [np.nan_to_num(h, nan=0.0) for h in host_parts], order


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("data_partitioned", [True, False])
def test_sort_last_equals_single_gpu_without_ert(R, orc, world, data_partitioned):
    import torch
    b, fields, dom = _scene(R)
    W, H = 160, 96
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H))
    fd.opacity_scale = 0.05
    fd.ert = 2.0                                                     # threshold off: the composite must be exact up to fp32 re-association
    r1 = R.Renderer(b, fields)
    fd.apply(r1)
    r1.render()
    ref_acc, ref_rgba = r1.read_accum(), r1.read_rgba8()
    acc, rgba, res, parts, order = _sort_last_frame(R, torch, b, fields, fd, world, data_partitioned)
    assert np.abs(acc[..., :3] - ref_acc[..., :3]).max() <= RGBA_TOL
    assert np.abs(scenes.unpack_rgba8(rgba) - scenes.unpack_rgba8(ref_rgba)).max() <= 1
    assert ref_acc[..., :3].max() > 0.05
    # the numpy restatement of the composite loop agrees with the kernel (premultiplied colour before the frame tail)
    comp = multigpu.composite_reference(np, parts, order, ert=2.0)
    col = comp[:, 3:4] * comp[:, :3]
    assert np.abs(col.reshape(H, W, 3) - acc[..., :3]).max() <= 1e-5
    if data_partitioned:
        total_cells = b.total_num_cells
        assert sum(c for _, c in res) < (2.5 if world <= 4 else 4.0) * total_cells   # straddling bricks are duplicated (this scene has 9 192 cells in a few bricks), nothing more
        assert max(c for _, c in res) < total_cells                  # no rank holds everything
        assert all(nb <= b.num_bricks for nb, _ in res)


def test_sort_last_with_ert_is_bounded(R, orc):
    """per-rank early termination (threshold 0.98 on local alpha, H1): bounded deviation from the single-GPU frame"""
    import torch
    b, fields, dom = _scene(R)
    W, H = 160, 96
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H))
    fd.opacity_scale = 0.6
    r1 = R.Renderer(b, fields)
    fd.apply(r1)
    r1.render()
    ref = r1.read_accum()
    acc, _, _, _, _ = _sort_last_frame(R, torch, b, fields, fd, 4, True)
    assert np.abs(acc[..., :3] - ref[..., :3]).max() <= 0.03        # (1 - 0.98) of a colour, plus a sample


def test_sort_last_composites_surfaces(R, orc):
    """iso-surface + DVR through the sort-last path: the nearest surface of all ranks is the background"""
    import torch
    b, fields, dom = _scene(R)
    W, H = 160, 96
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H))
    fd.iso = [(1, 0.5, 0), (0, 0.0, 0)]
    fd.opacity_scale = 0.02
    fd.ert = 2.0
    r1 = R.Renderer(b, fields)
    fd.apply(r1)
    r1.render()
    ref = r1.read_accum()
    acc, _, _, _, _ = _sort_last_frame(R, torch, b, fields, fd, 4, True, surf=True)
    err = np.abs(acc[..., :3] - ref[..., :3]).max(axis=-1)
    # an iso crossing bracketed by the last sample of one part and the first of the next is not seen (the
    # functor state does not travel between ranks): a thin seam of pixels may differ
    assert (err > RGBA_TOL).sum() <= 0.02 * err.size, "%d pixels differ" % int((err > RGBA_TOL).sum())
    assert ref[..., :3].max() > 0.2
    # a surface without a surface buffer is refused
    r = R.Renderer(b, fields, part=(0, 2))
    fd.apply(r)
    p = torch.zeros((W * H, 4), dtype=torch.float32, device="cuda")
    with pytest.raises(RuntimeError, match="surface"):
        r.render_partial(p.data_ptr())


def test_partitioned_renderer_samples_only_its_part(R, orc):
    """points inside a rank's part reconstruct exactly as on the full renderer; points of other parts are invalid
    (their regions are pruned), never garbage"""
    b, fields, dom = _scene(R, channels=2)
    full = R.Renderer(b, fields)
    pts = synth.random_points(100000, [0.0] * 3, [float(dom)] * 3, seed=5)
    fv, fok = full.sample_points(pts, 1)
    seen = np.zeros(len(pts), int)
    for k in range(4):
        r = R.Renderer(b, fields, part=(k, 4))
        v, ok = r.sample_points(pts, 1)
        m = ok == 1
        assert (fok[m] == 1).all() and np.array_equal(v[m], fv[m])
        seen += m
    assert ((seen == 1) == (fok == 1)).all()                        # every valid point belongs to exactly one part
    with pytest.raises(Exception):
        r.set_partition(0, 0, 1)                                     # the data of the other parts is not resident


@pytest.mark.parametrize("world", [2, 3, 8])
def test_frame_strips_union_is_the_frame(R, orc, built_lib, world):
    """image-tile mode with the frame as N row strips: each 'rank' stores its tiles into the strips they fall in;
    the strips, concatenated, are the single-renderer frame bit for bit"""
    import torch
    b, fields, dom = _scene(R, mlw=127)
    W, H = 200, 116                                                 # not a multiple of the tile size
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H))
    r = R.Renderer(b, fields)
    fd.apply(r)
    r.render()
    ref = r.read_rgba8().copy()
    strips = []
    for k in range(world):
        y0, y1 = multigpu.strip_rows(k, world, H)
        cy0, cy1 = C.c_int(), C.c_int()
        built_lib.check(built_lib.lib.exa_strip_rows(k, world, H, C.byref(cy0), C.byref(cy1)))
        assert (cy0.value, cy1.value) == (y0, y1)
        strips.append(torch.zeros((max(y1 - y0, 1), W), dtype=torch.int32, device="cuda"))
    ptrs = (C.c_void_p * world)(*[s.data_ptr() for s in strips])
    built_lib.check(built_lib.lib.exa_renderer_set_color_strips(r.h, ptrs, world))
    for k in range(world):
        r.set_partition(built_lib.PARTITION_IMAGE_TILES, k, world)
        r.render()
    got = np.concatenate([strips[k].cpu().numpy().view(np.uint32)[:multigpu.strip_rows(k, world, H)[1] - multigpu.strip_rows(k, world, H)[0]] for k in range(world)])
    assert got.shape == ref.shape and (got == ref).all()
    built_lib.check(built_lib.lib.exa_renderer_set_color_strips(r.h, None, 0))
    r.set_partition(built_lib.PARTITION_NONE, 0, 1)
    r.render()
    assert (r.read_rgba8() == ref).all()


def test_weighted_frame_strips(R, orc, built_lib):
    """strips of caller-chosen heights (sized by copy bandwidth in the bench): same frame bit for bit; bad tables refused"""
    import torch
    b, fields, dom = _scene(R, mlw=127)
    W, H = 200, 116
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H))
    r = R.Renderer(b, fields)
    fd.apply(r)
    r.render()
    ref = r.read_rgba8().copy()
    for weights in ([1.0, 1.0, 2.0, 2.0], [12.0, 12.0, 21.0], [1.0, 0.0, 1.0], [3.0]):
        world = len(weights)
        rows = multigpu.weighted_strip_rows(weights, H)
        assert rows[0] == 0 and rows[-1] == H and all(a <= c for a, c in zip(rows, rows[1:])) and all(y % 8 == 0 or y == H for y in rows)
        strips = [torch.zeros((max(rows[k + 1] - rows[k], 1), W), dtype=torch.int32, device="cuda") for k in range(world)]
        ptrs = (C.c_void_p * world)(*[s.data_ptr() for s in strips])
        built_lib.check(built_lib.lib.exa_renderer_set_color_strips_rows(r.h, ptrs, (C.c_int * (world + 1))(*rows), world))
        for k in range(world):
            r.set_partition(built_lib.PARTITION_IMAGE_TILES, k, world)
            r.render()
        got = np.concatenate([strips[k].cpu().numpy().view(np.uint32)[:rows[k + 1] - rows[k]] for k in range(world)])
        assert got.shape == ref.shape and (got == ref).all()
    assert multigpu.weighted_strip_rows([1.0] * 3, H) == [multigpu.strip_rows(k, 3, H)[0] for k in range(3)] + [H]
    assert multigpu.weighted_strip_rows([2.0, 1.0], 64) == [0, 40, 64]      # 8 tile rows: 5.33 / 2.67 -> 5 + 3
    two = (C.c_void_p * 2)(strips[0].data_ptr(), strips[0].data_ptr())
    for bad in ([0, 4, H], [0, 64, H - 1], [8, 64, H], [0, 64, 32]):
        assert built_lib.lib.exa_renderer_set_color_strips_rows(r.h, two, (C.c_int * 3)(*bad), 2) == built_lib.EXA_ERR_INVALID
    built_lib.check(built_lib.lib.exa_renderer_set_color_strips(r.h, None, 0))
    r.set_partition(built_lib.PARTITION_NONE, 0, 1)
    r.render()
    assert (r.read_rgba8() == ref).all()


def test_tracer_params_take_effect_at_reset(R, orc):
    """growing numTraces / numTimesteps without a reset must not change the shape the kernels index
    the trace buffer with (the reference pushes them inside resetTracer, OptixRenderer.cpp:450-474)"""
    cells, _, dom = scenes.small_scene("nb64_off")
    c = cells[:, :3] + 0.5 * (1 << cells[:, 3:4])
    vx = (0.2 + 0 * c[:, 0]).astype(np.float32); vy = (0.1 + 0 * c[:, 0]).astype(np.float32); vz = (0.3 + 0 * c[:, 0]).astype(np.float32)
    b = R.ExaBricks.build(cells)
    r = R.Renderer(b, [vx, vy, vz])
    seeds = synth.random_points(16, [10, 10, 10], [50, 50, 50], seed=3)
    r.setTracerParams(16, 10, 0.5, (0, 1, 2))
    r.setTracerSeeds(seeds)
    r.setTracerEnabled(True)
    r.traceAll()
    before = r.readTraces()
    assert before.shape == (16, 10, 3)
    r.setTracerParams(4000, 500, 0.5, (0, 1, 2))                    # 12 500x larger, no reset
    r.advanceTracer()
    W, H = 64, 48
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H), 3)
    fd.apply(r)
    r.render()                                                       # cones of the OLD shape
    assert np.array_equal(r.readTraces(), before)
    r.resetTracer()                                                  # now the new shape
    assert r.readTraces().shape == (4000, 500, 3)


@pytest.mark.parametrize("world", [2, 5])
def test_cost_balanced_tile_lists_partition_the_frame(R, orc, built_lib, world):
    """image tiles dealt by measured cost (exa_balance_tiles + exa_renderer_set_tile_list): the lists of the ranks are a
    partition of the tiles, balanced in cost, and the union of what they render is the single-renderer frame bit for bit"""
    import torch
    lib, check = built_lib.lib, built_lib.check
    b, fields, dom = _scene(R, mlw=127)
    W, H = 200, 116
    nt = ((W + 15) // 16) * ((H + 7) // 8)
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H))
    fd.opacity_scale = 0.1
    r = R.Renderer(b, fields)
    fd.apply(r)
    r.render()
    ref = r.read_rgba8().copy()
    check(lib.exa_renderer_enable_tile_costs(r.h, 1))
    r.render()
    costs = np.zeros(nt, np.uint32)
    check(lib.exa_renderer_get_tile_costs(r.h, costs.ctypes.data_as(C.c_void_p), nt))
    check(lib.exa_renderer_enable_tile_costs(r.h, 0))
    assert (costs > 0).all()                                        # every tile was rendered and timed
    fb = torch.zeros((H, W), dtype=torch.int32, device="cuda")
    check(lib.exa_renderer_set_color_buffer(r.h, fb.data_ptr()))
    seen, loads = np.zeros(nt, int), []
    for k in range(world):
        mine, n = np.zeros(nt, np.uint32), C.c_int()
        check(lib.exa_balance_tiles(costs.ctypes.data_as(C.c_void_p), nt, world, k, mine.ctypes.data_as(C.c_void_p), C.byref(n)))
        mine = mine[:n.value]
        assert (np.diff(costs[mine].astype(np.int64)) <= 0).all()   # most expensive first
        seen[mine] += 1
        loads.append(int(costs[mine].astype(np.int64).sum()))
        r.set_partition(built_lib.PARTITION_IMAGE_TILES, k, world)
        check(lib.exa_renderer_set_tile_list(r.h, mine.ctypes.data_as(C.c_void_p), len(mine)))
        r.render()
    assert (seen == 1).all()
    assert max(loads) <= 1.05 * (sum(loads) / world) + int(costs.max())
    assert (fb.cpu().numpy().view(np.uint32) == ref).all()
    with pytest.raises(Exception):
        bad = np.array([nt], np.uint32)
        check(lib.exa_renderer_set_tile_list(r.h, bad.ctypes.data_as(C.c_void_p), 1))
    r.set_partition(built_lib.PARTITION_NONE, 0, 1)
    check(lib.exa_renderer_set_color_buffer(r.h, None))
    r.render()
    assert (r.read_rgba8() == ref).all()


@pytest.mark.parametrize("mode_name,parts", [("tiles", 1), ("tiles", 3), ("sortlast", 1), ("sortlast", 4)])
def test_renderer_group_behind_the_c_abi(R, orc, built_lib, mode_name, parts):
    """exa_group_*: one host thread, N devices (here one GPU listed `parts` times; with real peers the same code path
    uses cudaDeviceEnablePeerAccess): image tiles reproduce the single-renderer frame bit for bit, sort-last within
    1 RGBA8 level with the termination threshold off"""
    b, fields, dom = _scene(R)
    W, H = 200, 116
    fd = scenes.FrameDesc(W, H, R.default_camera([0] * 3, [dom] * 3, W, H))
    fd.opacity_scale = 0.05
    fd.ert = 2.0
    r1 = R.Renderer(b, fields)
    fd.apply(r1)
    r1.render()
    ref = r1.read_rgba8()
    mode = built_lib.PARTITION_IMAGE_TILES if mode_name == "tiles" else built_lib.PARTITION_SORT_LAST
    g = R.RendererGroup(b, fields, [0] * parts, mode)
    g.apply(fd)
    for _ in range(2):                                               # second frame: cost-ordered tile schedule
        img = g.render()
    if mode_name == "tiles":
        assert (img == ref).all()
    else:
        assert np.abs(scenes.unpack_rgba8(img) - scenes.unpack_rgba8(ref)).max() <= 1
    # tiles dealt by measured cost: the same frame, a balanced deal (no-op for sort-last)
    ratio = g.rebalance()
    assert 1.0 <= ratio < 1.2
    img = g.render()
    if mode_name == "tiles":
        assert (img == ref).all()
    # a per-device setter through the group's renderer handles, then a per-frame setter through the group
    for r in g.renderers:
        r.setGradientShadingDVR(False)
    alpha, rgb = fd.xf[0]
    built_lib.check(built_lib.lib.exa_group_update_xf(g.h, 0, alpha.ctypes.data_as(C.c_void_p), np.ascontiguousarray(rgb).ctypes.data_as(C.c_void_p), 0.0, 1.0, 0.2))
    r1.setGradientShadingDVR(False)
    r1.updateXF(0, alpha, rgb, (0.0, 1.0), 0.2)
    r1.render()
    img2 = g.render()
    assert np.abs(scenes.unpack_rgba8(img2) - scenes.unpack_rgba8(r1.read_rgba8())).max() <= (0 if mode_name == "tiles" else 1)
    assert (img2 != img).any()
    g.close()
