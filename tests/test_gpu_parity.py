"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI,
against the CPU oracle on the same seeded inputs and against the golden fixtures made from
the reference's own code.

Bars (BASELINE.json north_star): ABR/brick/cell-ID construction bit-exact; reconstructed
scalars within 1e-5 relative; rendered RGBA within 1e-3 max absolute per channel at an
identical step size, jitter and transfer function.
"""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_golden
from owlexabrick_b200 import synth
import scenes

pytestmark = pytest.mark.gpu

REL_TOL = 1e-5          # reconstruction: |gpu-oracle| <= REL_TOL * max(|oracle|, 1e-3 * max|field|)
RGBA_TOL = 1e-3         # frame: max abs per channel of the linear float colour


@pytest.fixture(scope="module")
def R(built_lib):
    assert built_lib.lib.exa_cuda_available() == 1, built_lib.last_error()
    from owlexabrick_b200 import renderer
    return renderer


@pytest.fixture(params=["flat", "dual"], autouse=True)
def scalar_layout(request, built_lib):
    """every GPU parity test runs against both device layouts of the scalars (flat reference
    layout, dual-cell records); results must be identical"""
    mode = built_lib.LAYOUT_FLAT if request.param == "flat" else built_lib.LAYOUT_DUAL
    built_lib.check(built_lib.lib.exa_set_scalar_layout(mode))
    yield request.param
    built_lib.lib.exa_set_scalar_layout(built_lib.LAYOUT_AUTO)


def make_pair(R, orc, cells, fields, mlw=127):
    """(product renderer, oracle scene) over the same input"""
    b = R.ExaBricks.build(cells, max_leaf_width=mlw)
    r = R.Renderer(b, fields)
    bricks, ids = b.headers, b.cell_ids
    sc = orc.flatten_scalars(ids, fields)
    regions, leaf = orc.build_abr(bricks.view(orc.BRICK_DT), sc)
    return r, orc.Scene(bricks.view(orc.BRICK_DT), sc, regions, leaf), b


# ------------------------------------------------------------------------------------------
# ABR build (exa/Regions.cpp) — bit-exact
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["ex0", "ex1", "ex2", "ex3", "ex4", "nb64", "nb64_mlw8", "nb64_dups"])
def test_abr_bit_exact_vs_reference_golden(R, orc, name):
    g = load_golden(name)
    b = R.ExaBricks.from_bytes(g["bricks"].tobytes())
    r = R.Renderer(b, list(g["fields"]))
    regions, leaf = r.get_abr()
    rr, rl = orc.parse_abr(g["abr"].tobytes())
    assert len(regions) == len(rr) and len(leaf) == len(rl)
    assert leaf.tobytes() == rl.tobytes()
    assert regions.tobytes() == rr.tobytes()


@pytest.mark.parametrize("kind,mlw,channels", [("nb32", 127, 1), ("nb64_off", 127, 3), ("nb64_off", 6, 2), ("nb64", 4, 1)])
def test_abr_bit_exact_vs_oracle(R, orc, kind, mlw, channels):
    cells, fields, _ = scenes.small_scene(kind, channels)
    r, sc, _ = make_pair(R, orc, cells, fields, mlw)
    regions, leaf = r.get_abr()
    assert leaf.tobytes() == sc.leaf.tobytes()
    assert regions.tobytes() == sc.regions.tobytes()
    nodes, depth = r.kd_info()
    assert nodes >= 2 * len(regions) - 1 and depth >= 1


def test_abr_c1_hash_vs_reference(R, orc):
    """config C1 (2.27 M cells): GPU ABR output == reference buildFrom output (SHA-256)"""
    hashes = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))
    cells = synth.nested_boxes(256, [(72, 184), (32, 224)])
    b = R.ExaBricks.build(cells)
    r = R.Renderer(b, [synth.analytic_field(cells, 256, 0)])
    regions, leaf = r.get_abr()
    raw = np.array([len(regions), len(leaf)], "<i4").tobytes() + regions.tobytes() + leaf.tobytes()
    assert hashlib.sha256(raw).hexdigest() == hashes["C1"]["abr_sha256"]


def test_abr_c5_quarter_scale_hash_vs_reference(R, orc):
    """C5 shape at 1/4 scale (26.6 M cells, 3 channels, 4 021 bricks): GPU ABR output (251 744 regions) ==
    reference buildFrom output (SHA-256; the oracle's equals it: tests/test_host.py)"""
    h = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))["C5_scale4"]
    cells, dom = synth.cloud_shaped(4)
    fields = [np.empty(len(cells), np.float32) for _ in range(3)]
    for s0 in range(0, len(cells), 1 << 22):
        for c, v in enumerate(synth.cloud_velocity(cells[s0:s0 + (1 << 22)], dom)):
            fields[c][s0:s0 + (1 << 22)] = v
    b = R.ExaBricks.build(cells)
    assert hashlib.sha256(b.to_bytes()).hexdigest() == h["bricks_sha256"]
    r = R.Renderer(b, fields)
    regions, leaf = r.get_abr()
    assert len(regions) == h["regions"]
    assert hashlib.sha256(regions.tobytes() + leaf.tobytes()).hexdigest() == h["abr_sha256"]


# ------------------------------------------------------------------------------------------
# tent-basis reconstruction (exabrick.cu:620-816, 883-942) — 1e-5 relative
# ------------------------------------------------------------------------------------------
def _check_samples(gv, gok, ov, ook, fmax):
    assert (gok == ook).all(), "validity differs at %d points" % int((gok != ook).sum())
    m = ook == 1
    err = np.abs(gv[m].astype(np.float64) - ov[m])
    bound = REL_TOL * np.maximum(np.abs(ov[m]), 1e-3 * fmax)
    assert (err <= bound).all(), "max rel err %.3g" % float((err / np.maximum(np.abs(ov[m]), 1e-3 * fmax)).max())


@pytest.mark.parametrize("kind,mlw,channels", [("nb32", 127, 1), ("nb64_off", 127, 3), ("nb64_off", 6, 2), ("nb64", 4, 1)])
def test_reconstruction_vs_oracle(R, orc, kind, mlw, channels):
    cells, fields, dom = scenes.small_scene(kind, channels)
    r, sc, _ = make_pair(R, orc, cells, fields, mlw)
    pts = synth.random_points(200000, [-1.0] * 3, [dom + 1.0] * 3, seed=1)
    for c in range(channels):
        gv, gok = r.sample_points(pts, c)
        ov, ook, _ = sc.sample(pts, c)
        _check_samples(gv, gok, ov, ook, float(np.abs(fields[c]).max()))
    gv, gok, gg = r.sample_points(pts, 0, grad=True)
    ov, ook, og, _ = sc.sample(pts, 0, grad=True)
    _check_samples(gv, gok, ov, ook, float(np.abs(fields[0]).max()))
    m = ook == 1
    gscale = np.abs(og[m]).max()
    assert np.abs(gg[m] - og[m]).max() <= 1e-4 * gscale


def test_reconstruction_1m_points_c1(R, orc):
    """north_star gate: 1 M random points on config C1 within 1e-5 relative"""
    cells = synth.nested_boxes(256, [(72, 184), (32, 224)])
    f = synth.analytic_field(cells, 256, 0)
    r, sc, _ = make_pair(R, orc, cells, [f])
    pts = synth.random_points(1000000, [0.0] * 3, [256.0] * 3, seed=1)
    gv, gok = r.sample_points(pts, 0)
    ov, ook, st = sc.sample(pts, 0)
    _check_samples(gv, gok, ov, ook, float(np.abs(f).max()))
    assert gok.sum() > 990000
    # instrumented run: the counters behind the algorithmic-bytes model equal the oracle's
    import torch
    d = torch.from_numpy(pts).cuda()
    gs = r.sample_points_stats(d.data_ptr(), len(pts), 0)
    assert gs["brickVisits"] == st["brickVisits"] and gs["cornerLoads"] == st["cornerLoads"]


def test_reconstruction_edge_cases(R, orc):
    cells, fields, dom = scenes.small_scene("nb32")
    r, sc, _ = make_pair(R, orc, cells, fields)
    gv, gok = r.sample_points(np.zeros((0, 3), np.float32))
    assert len(gv) == 0
    # points on region faces, on the outer boundary, far outside, NaN
    faces = np.unique(np.concatenate([sc.regions["lo"].ravel(), sc.regions["hi"].ravel()]))
    rng = np.random.RandomState(2)
    pts = rng.uniform(0, dom, size=(3000, 3)).astype(np.float32)
    pts[:1000, 0] = rng.choice(faces, 1000)
    pts[1000:2000, 1] = rng.choice(faces, 1000)
    pts[2000:, 2] = rng.choice(faces, 1000)
    extra = np.array([[-0.5, -0.5, -0.5], [dom + 0.5] * 3, [1e9, 0, 0], [np.nan, 1, 1], [-1e9, -1e9, -1e9]], np.float32)
    pts = np.concatenate([pts, extra])
    gv, gok = r.sample_points(pts)
    ov, ook, _ = sc.sample(pts)
    _check_samples(gv, gok, ov, ook, 1.0)
    with pytest.raises(Exception):
        r.sample_points(pts, channel=5)


# ------------------------------------------------------------------------------------------
# frames: DVR, gradient shading, clip box, accumulation, iso-surfaces — 1e-3 abs
# ------------------------------------------------------------------------------------------
def _frame_pair(R, orc, kind="nb32", channels=1, size=(96, 64), mlw=127):
    cells, fields, dom = scenes.small_scene(kind, channels)
    r, sc, _ = make_pair(R, orc, cells, fields, mlw)
    cam = R.default_camera([0] * 3, [dom] * 3, size[0], size[1])
    return r, sc, scenes.FrameDesc(size[0], size[1], cam, channels), dom


def _render_both(r, sc, fd, orc, accum=None):
    fd.apply(r)
    r.render()
    g_acc, g_rgba = r.read_accum(), r.read_rgba8()
    o_acc, o_rgba, o_st = sc.render(fd.to_orc(orc), accum=accum)
    return g_acc, g_rgba, o_acc, o_rgba, o_st


def _assert_frame(g_acc, g_rgba, o_acc, o_rgba):
    err = np.abs(g_acc[..., :3] - o_acc[..., :3]).max()
    assert err <= RGBA_TOL, "max abs colour error %.3g" % err
    d8 = np.abs(scenes.unpack_rgba8(g_rgba) - scenes.unpack_rgba8(o_rgba)).max()
    assert d8 <= 1, "RGBA8 differs by %d levels" % d8
    assert o_acc[..., :3].max() > 0.05, "test frame is empty"


@pytest.mark.parametrize("grad", [True, False])
@pytest.mark.parametrize("kind,channels,mlw", [("nb32", 1, 127), ("nb64_off", 2, 127), ("nb64", 1, 6)])
def test_dvr_frame_vs_oracle(R, orc, kind, channels, mlw, grad):
    r, sc, fd, dom = _frame_pair(R, orc, kind, channels, mlw=mlw)
    fd.grad_dvr = grad
    fd.opacity_scale = 0.3
    _assert_frame(*_render_both(r, sc, fd, orc)[:4])


def test_dvr_stats_match_oracle(R, orc):
    r, sc, fd, dom = _frame_pair(R, orc, "nb32", 1)
    fd.opacity_scale = 0.2
    fd.apply(r)
    gs = r.render_stats()
    _, _, os_ = sc.render(fd.to_orc(orc))
    for k in ("samples", "brickVisits", "cornerLoads", "segments"):
        assert abs(gs[k] - os_[k]) <= 2e-3 * os_[k] + 32 * fd.width * fd.height * (k != "segments"), (k, gs[k], os_[k])


def test_dvr_no_ert_no_skipping_dt(R, orc):
    r, sc, fd, dom = _frame_pair(R, orc, "nb32", 1)
    fd.ert = 2.0
    fd.space_skipping = False
    fd.dt = 0.8
    fd.opacity_scale = 0.05
    _assert_frame(*_render_both(r, sc, fd, orc)[:4])


def test_dvr_space_skipping_partial_xf(R, orc):
    """XF that is transparent below 0.6: most regions must be skipped, result unchanged"""
    r, sc, fd, dom = _frame_pair(R, orc, "nb64", 1, mlw=8)
    alpha, rgb = fd.xf[0]
    alpha = np.where(np.arange(128) / 127.0 < 0.6, 0.0, alpha).astype(np.float32)
    fd.xf[0] = (alpha, rgb)
    g_acc, g_rgba, o_acc, o_rgba, _ = _render_both(r, sc, fd, orc)
    _assert_frame(g_acc, g_rgba, o_acc, o_rgba)
    fd.space_skipping = False
    g2, _, _, _, _ = _render_both(r, sc, fd, orc)
    assert np.abs(g2 - g_acc).max() <= 1e-5


def test_dvr_clip_box_and_transform(R, orc):
    r, sc, fd, dom = _frame_pair(R, orc, "nb32", 1)
    # world = voxel/16 - 1  <=>  voxel = 16*(world+1)
    fd.xfm = np.array([16, 0, 0, 0, 16, 0, 0, 0, 16, 16, 16, 16], np.float32)
    fd.cam = R.default_camera([-1] * 3, [1] * 3, fd.width, fd.height)
    fd.clip = ([-0.5, -1, -1], [0.7, 1, 0.4])
    fd.opacity_scale = 0.3
    _assert_frame(*_render_both(r, sc, fd, orc)[:4])


def test_progressive_accumulation(R, orc):
    r, sc, fd, dom = _frame_pair(R, orc, "nb32", 1, size=(48, 32))
    fd.opacity_scale = 0.3
    o_acc = None
    for frame in range(3):
        fd.frame_id = frame
        g_acc, g_rgba, o_acc, o_rgba, _ = _render_both(r, sc, fd, orc, accum=o_acc)
        assert np.abs(g_acc[..., :3] - o_acc[..., :3]).max() <= RGBA_TOL * (frame + 1)
        assert np.abs(scenes.unpack_rgba8(g_rgba) - scenes.unpack_rgba8(o_rgba)).max() <= 1


def test_clock_scale_heat_map(R, orc):
    """frameState.clockScale > 0 (exabrick.cu:1703-1707): the red channel becomes min(clockScale * cycles / 1e6, 1);
    green and blue, and the frame with clockScale back at 0, are untouched"""
    r, sc, fd, dom = _frame_pair(R, orc, "nb32", 1, size=(48, 32))
    fd.opacity_scale = 0.3
    fd.apply(r)
    r.render()
    ref = r.read_accum().copy()
    r.setClockScale(10.0)
    r.render()
    hot = r.read_accum().copy()
    assert np.array_equal(hot[..., 1:3], ref[..., 1:3])
    assert (hot[..., 0] > 0).all() and (hot[..., 0] <= 1.0).all() and not np.array_equal(hot[..., 0], ref[..., 0])
    r.setClockScale(1e9)
    r.render()
    assert (r.read_accum()[..., 0] == 1.0).all()                    # saturates
    r.setClockScale(0.0)
    r.render()
    assert np.array_equal(r.read_accum(), ref)


@pytest.mark.parametrize("grad_iso", [True, False])
def test_iso_surface_plus_dvr(R, orc, grad_iso):
    r, sc, fd, dom = _frame_pair(R, orc, "nb64_off", 1)
    fd.iso = [(1, 0.5, 0), (0, 0.0, 0)]
    fd.grad_iso = grad_iso
    fd.opacity_scale = 0.02
    g_acc, g_rgba, o_acc, o_rgba, _ = _render_both(r, sc, fd, orc)
    # an iso hit decided by samples that bracket the value to within rounding may flip
    # between implementations: allow a handful of pixels, bound everything else
    err = np.abs(g_acc[..., :3] - o_acc[..., :3]).max(axis=-1)
    assert (err > RGBA_TOL).sum() <= 0.002 * err.size, "%d pixels differ" % int((err > RGBA_TOL).sum())
    assert o_acc[..., :3].max() > 0.2


def test_two_iso_surfaces(R, orc):
    r, sc, fd, dom = _frame_pair(R, orc, "nb32", 2)
    fd.iso = [(1, 0.35, 0), (1, 0.65, 1)]
    fd.xf[0] = (np.zeros(128, np.float32), fd.xf[0][1])      # no DVR contribution from channel 0
    fd.xf[1] = (np.zeros(128, np.float32), fd.xf[1][1])
    g_acc, g_rgba, o_acc, o_rgba, _ = _render_both(r, sc, fd, orc)
    err = np.abs(g_acc[..., :3] - o_acc[..., :3]).max(axis=-1)
    assert (err > RGBA_TOL).sum() <= 0.005 * err.size, "%d pixels differ" % int((err > RGBA_TOL).sum())


def test_pipelined_readback_equals_blocking(R, orc):
    """update -> render_async -> read_rgba8_async over 5 frames with a different camera and
    transfer function each == the same frames rendered and read one by one"""
    import torch
    r, sc, fd, _ = _frame_pair(R, orc, "nb64_off", 1, size=(160, 96))
    W, H = fd.width, fd.height
    fd.apply(r)
    alpha, rgb = fd.xf[0]
    def inputs(i):
        pos = np.asarray(fd.cam[0], np.float32) + np.float32(0.7 * i)
        a = np.clip(np.asarray(alpha, np.float32) * (1.0 - 0.15 * i), 0, 1)
        return (pos,) + tuple(fd.cam[1:]), a
    ref = []
    for i in range(5):
        cam, a = inputs(i)
        r.updateCamera(*cam); r.updateXF(0, a, rgb, fd.xf_domain[0], fd.opacity_scale)
        r.render()
        ref.append(r.read_rgba8().copy())
    assert any((ref[i] != ref[0]).any() for i in range(1, 5))
    host = [torch.empty((H, W), dtype=torch.int32).pin_memory() for _ in range(5)]
    for i in range(5):
        cam, a = inputs(i)
        r.updateCamera(*cam); r.updateXF(0, a, rgb, fd.xf_domain[0], fd.opacity_scale)
        r.render(sync=False)
        r.read_rgba8_async(host[i].data_ptr())
    r.sync()
    for i in range(5):
        assert (host[i].numpy().view(np.uint32) == ref[i]).all(), i
    # a caller-owned frame buffer cannot be double-buffered: loud error
    fb = torch.empty((H, W), dtype=torch.int32, device="cuda")
    r.resizeFrameBuffer(fb.data_ptr(), (W, H))
    with pytest.raises(RuntimeError):
        r.read_rgba8_async(host[0].data_ptr())


@pytest.mark.parametrize("world", [2, 3, 8])
def test_image_tile_partition_union_is_the_frame(R, orc, built_lib, world):
    """multi-GPU image-tile mode, emulated on one GPU: each 'rank' renders exactly its
    multigpu.owned_tiles into one shared frame buffer; the union is the 1-rank frame, bit for bit"""
    import torch
    from owlexabrick_b200 import multigpu
    r, sc, fd, _ = _frame_pair(R, orc, "nb64_off", 1, size=(200, 116))      # not a multiple of the 16x8 tile
    fd.apply(r)
    r.render()
    ref = r.read_rgba8().copy()
    W, H = fd.width, fd.height
    assert (ref != 0).all()                                                 # alpha byte: a written pixel is never 0
    fb = torch.zeros((H, W), dtype=torch.int32, device="cuda")
    built_lib.check(built_lib.lib.exa_renderer_set_color_buffer(r.h, fb.data_ptr()))
    tiles_x = (W + 15) // 16
    seen = np.zeros((H, W), bool)
    for rank in range(world):
        r.set_partition(built_lib.PARTITION_IMAGE_TILES, rank, world)
        r.render(); r.render()                                              # second frame: cost-ordered tile schedule
        now = fb.cpu().numpy().view(np.uint32) != 0
        mine = np.zeros((H, W), bool)
        for t in multigpu.owned_tiles(rank, world, W, H):
            ty, tx = divmod(t, tiles_x)
            mine[ty * 8:ty * 8 + 8, tx * 16:tx * 16 + 16] = True
        assert ((now & ~seen) == mine).all(), "rank %d wrote outside its tiles / missed some" % rank
        seen = now
    assert (fb.cpu().numpy().view(np.uint32) == ref).all()
    r.set_partition(built_lib.PARTITION_NONE, 0, 1)
    built_lib.check(built_lib.lib.exa_renderer_set_color_buffer(r.h, None))


# ------------------------------------------------------------------------------------------
# full size (BASELINE.json configs[1], ~95 M cells): size-independent properties, no oracle
# ------------------------------------------------------------------------------------------
_C2 = {}


def _c2_scene(R):
    """bricks + three fields of config C2, built once per session: f (SURVEY 8d analytic field),
    the constant 1, and 2 f + 3"""
    if not _C2:
        cells = synth.nested_boxes(1024, [(320, 704), (192, 832)])
        assert len(cells) == 94994432
        f = np.empty(len(cells), np.float32)
        for b in range(0, len(cells), 1 << 23):
            f[b:b + (1 << 23)] = synth.analytic_field_chunked(cells[b:b + (1 << 23)], 1024, 0, first=b)
        _C2["bricks"] = R.ExaBricks.build(cells)
        _C2["fields"] = [f, np.ones_like(f), (2.0 * f.astype(np.float64) + 3.0).astype(np.float32)]
    return _C2["bricks"], _C2["fields"]


def test_full_size_c2_properties(R, orc, scalar_layout):
    """at full size: partition of unity (constant field reconstructs to itself at every one of
    1 M points), linearity across channels, every point of the domain is covered, the frame
    is reproducible launch to launch (tile schedule changes between frames), multi-channel DVR
    of identical channels is well defined"""
    bricks, fields = _c2_scene(R)
    r = R.Renderer(bricks, fields)
    assert r.scalar_layout()[0] == (1 if scalar_layout == "flat" else 2)
    regions, leaf = r.get_abr()
    assert leaf.min() >= 0 and leaf.max() < bricks.num_bricks and len(set(leaf.tolist())) == bricks.num_bricks
    pts = synth.random_points(1 << 20, [0.0] * 3, [1024.0] * 3, seed=7)
    v0, ok0 = r.sample_points(pts, 0)
    v1, ok1 = r.sample_points(pts, 1)
    v2, ok2 = r.sample_points(pts, 2)
    assert ok0.all() and ok1.all() and ok2.all()                       # the cells tile the whole domain
    assert np.abs(v1 - 1.0).max() <= 2e-6                              # sum w * 1 / sum w
    lin = 2.0 * v0.astype(np.float64) + 3.0
    assert (np.abs(v2 - lin) <= REL_TOL * np.abs(lin)).all()
    fmin, fmax = float(fields[0].min()), float(fields[0].max())
    assert v0.min() >= fmin - 1e-5 and v0.max() <= fmax + 1e-5        # convex combination of cell values
    # the frame: same bits from consecutive launches although the tile order changes
    cam = R.default_camera([0] * 3, [1024.0] * 3, 1024, 1024)
    fd = scenes.FrameDesc(1024, 1024, cam, 3)                           # multi-field DVR over the three channels
    fd.xf_domain = [(fmin, fmax), (0.0, 2.0), (2 * fmin + 3, 2 * fmax + 3)]
    fd.apply(r)
    frames = []
    for _ in range(3):
        r.render()
        frames.append(r.read_rgba8().copy())
    assert (frames[0] == frames[1]).all() and (frames[1] == frames[2]).all()
    st = r.render_stats()
    assert st["samples"] > 10 ** 6 and st["segments"] > 10 ** 5
    # transparent transfer function => every pixel is the background, whatever the data
    alpha, rgb = fd.xf[0]
    for c in range(3):
        r.updateXF(c, np.zeros_like(alpha), rgb, fd.xf_domain[c], 1.0)
    r.render()
    acc = r.read_accum()
    r2 = R.Renderer(bricks, [fields[1]])                                # constant field, same camera: no structure at all
    fd1 = scenes.FrameDesc(1024, 1024, cam, 1)
    fd1.xf_domain = [(0.0, 2.0)]
    fd1.apply(r2)
    r2.updateXF(0, np.zeros_like(alpha), rgb, (0.0, 2.0), 1.0)
    r2.render()
    assert np.abs(acc - r2.read_accum()).max() == 0.0


def test_render_errors(R, orc):
    g = load_golden("ex4")
    b = R.ExaBricks.from_bytes(g["bricks"].tobytes())
    r = R.Renderer(b, [g["fields"][0]])
    with pytest.raises(Exception, match="resizeFrameBuffer"):
        r.render()
    with pytest.raises(RuntimeError, match="mismatching xf size"):
        r.updateXF(0, np.zeros(100, np.float32), np.zeros((100, 3), np.float32), (0, 1), 1.0)
    with pytest.raises(Exception):
        R.Renderer(b, [g["fields"][0][:50]])                 # "invalid cell ID" (OptixRenderer.cpp:125)
    lp, fs = r.launch_params()
    assert len(lp) == 240 and len(fs) == 408
    assert r.launch_count() > 0


# ------------------------------------------------------------------------------------------
# streamlines (exabrick.cu:1531-1574)
# ------------------------------------------------------------------------------------------
def test_streamlines_vs_oracle(R, orc):
    cells, _, dom = scenes.small_scene("nb64_off")
    c = cells[:, :3] + 0.5 * (1 << cells[:, 3:4])
    ctr = np.array([32, 32, 32.0])
    # swirling flow around the z axis plus drift
    vx = (-(c[:, 1] - ctr[1]) * 0.05 + 0.2).astype(np.float32)
    vy = ((c[:, 0] - ctr[0]) * 0.05).astype(np.float32)
    vz = (0.3 + 0.002 * c[:, 2]).astype(np.float32)
    r, sc, _ = make_pair(R, orc, cells, [vx, vy, vz])
    seeds = synth.random_points(200, [10, 10, 4], [54, 54, 20], seed=11)
    T = 60
    r.setTracerParams(len(seeds), T, 0.7, (0, 1, 2))
    r.setTracerSeeds(seeds)
    r.setTracerEnabled(True)
    r.traceAll()
    g = r.readTraces()
    o = sc.trace(seeds, T, (0, 1, 2), 0.7, [0, 0, 0], [dom] * 3)
    dead_g, dead_o = g[..., 0] >= 2e10, o[..., 0] >= 2e10
    # a trace may die one step earlier/later when it grazes the boundary
    assert (dead_g != dead_o).sum() <= 3
    m = ~dead_g & ~dead_o
    assert m.sum() > 0.5 * m.size
    assert np.abs(g[m] - o[m]).max() <= 1e-5 * dom * 4       # <= 1e-5 relative to the domain, accumulated over 60 RK4 steps
    # per-frame stepping (advanceTracer) gives the same vertices as traceAll
    r.setTracerSeeds(seeds)
    for _ in range(T + 2):
        r.advanceTracer()
    assert np.array_equal(r.readTraces(), g)
