"""GPU parity tests of the surface pass of the frame kernel (run with -m gpu on a B200): contour
planes, streamline cones and ambient occlusion (programs/exabrick.cu:441-573, 1267-1406,
1476-1529, 1604-1652) through the C ABI against the CPU oracle.  Bar: RGBA within 1e-3 max
absolute per channel; a surface hit decided to within rounding may flip between the two
implementations, so a small fraction of pixels is exempt (as in the iso-surface tests)."""
import numpy as np
import pytest

import scenes
from test_gpu_parity import R, scalar_layout, make_pair, _frame_pair, _render_both, RGBA_TOL  # noqa: F401 (fixtures)

pytestmark = pytest.mark.gpu

OFF = (0, (1.0, 0.0, 0.0), 0.5, 0)


def _close(g_acc, o_acc, frac=0.003):
    err = np.abs(g_acc[..., :3] - o_acc[..., :3]).max(axis=-1)
    bad = int((err > RGBA_TOL).sum())
    assert bad <= frac * err.size, "%d of %d pixels differ by more than %g (max %.3g)" % (bad, err.size, RGBA_TOL, err.max())


def _world(r, fd):
    lp, _ = r.launch_params()                            # worldSpaceBounds_lo/_hi at offset 48 (LaunchParams.h:37)
    b = np.frombuffer(lp[48:72], np.float32)
    fd.world = (tuple(float(x) for x in b[:3]), tuple(float(x) for x in b[3:]))


@pytest.mark.parametrize("normal,offset", [((1.0, 0.0, 0.0), 0.5), ((0.0, 0.0, 2.0), 0.4), ((1.0, 1.0, 1.0), 0.9)])
def test_contour_plane_vs_oracle(R, orc, normal, offset):
    r, sc, fd, dom = _frame_pair(R, orc, "nb64_off", 2)
    _world(r, fd)
    fd.opacity_scale = 0.05
    fd.contour = [(1, normal, offset, 1), OFF, OFF]
    g_acc, g_rgba, o_acc, o_rgba, _ = _render_both(r, sc, fd, orc)
    _close(g_acc, o_acc)
    # the plane is there: the frame differs from the one without it
    fd.contour = [OFF] * 3
    fd.apply(r)
    r.render()
    assert np.abs(r.read_accum()[..., :3] - g_acc[..., :3]).max() > 0.05


def test_three_contour_planes_and_iso(R, orc):
    r, sc, fd, dom = _frame_pair(R, orc, "nb32", 1)
    _world(r, fd)
    fd.opacity_scale = 0.02
    fd.iso = [(1, 0.5, 0), (0, 0.0, 0)]
    fd.contour = [(1, (1.0, 0.0, 0.0), 0.3, 0), (1, (0.0, 1.0, 0.0), 0.6, 0), (1, (0.2, 0.1, 1.0), 0.5, 0)]
    g_acc, _, o_acc, _, _ = _render_both(r, sc, fd, orc)
    _close(g_acc, o_acc, 0.006)
    assert o_acc[..., :3].max() > 0.2


def _flow_pair(R, orc):
    cells, _, dom = scenes.small_scene("nb64_off")
    c = cells[:, :3] + 0.5 * (1 << cells[:, 3:4])
    vx = (-(c[:, 1] - 32) * 0.05 + 0.2).astype(np.float32)
    vy = ((c[:, 0] - 32) * 0.05).astype(np.float32)
    vz = (0.3 + 0.002 * c[:, 2]).astype(np.float32)
    r, sc, _ = make_pair(R, orc, cells, [vx, vy, vz])
    cam = R.default_camera([0] * 3, [dom] * 3, 96, 64)
    fd = scenes.FrameDesc(96, 64, cam, 3)
    fd.xf_domain = [(-2.0, 2.0)] * 3
    fd.opacity_scale = 0.01
    return r, sc, fd, dom


def _start_tracer(r, dom, n=60, T=40):
    from owlexabrick_b200 import synth
    seeds = synth.random_points(n, [12, 12, 6], [52, 52, 30], seed=7)
    r.setTracerParams(len(seeds), T, 0.9, (0, 1, 2))
    r.setTracerSeeds(seeds)
    r.setTracerEnabled(True)
    return T


@pytest.mark.parametrize("steps", [1, 2, 17, 45])
def test_streamline_cones_vs_oracle(R, orc, steps):
    """cones drawn over the traces the GPU tracer produced; the oracle renders the SAME vertices
    (read back), so this checks visibility (k < currentTimestep-1, dead ends), the per-trace boxes
    and the cone intersection, not the advection (test_streamlines_vs_oracle does that)"""
    r, sc, fd, dom = _flow_pair(R, orc)
    _world(r, fd)
    T = _start_tracer(r, dom)
    for _ in range(steps):
        r.advanceTracer()
    fd.traces = r.readTraces()
    fd.current_timestep = min(steps, T)
    g_acc, _, o_acc, _, _ = _render_both(r, sc, fd, orc)
    _close(g_acc, o_acc)
    if steps >= 17:
        fd2 = scenes.FrameDesc(96, 64, fd.cam, 3)         # same frame without cones
        fd2.xf_domain, fd2.opacity_scale, fd2.world = fd.xf_domain, fd.opacity_scale, fd.world
        o2, _, _ = sc.render(fd2.to_orc(orc))
        assert (np.abs(o2[..., :3] - o_acc[..., :3]).max(axis=-1) > 0.05).sum() > 50, "no cones in the test frame"


def test_ambient_occlusion_vs_oracle(R, orc):
    """two AO rays per shaded hit (cones + iso-surface as occluders); the AO direction uses
    cosf/sinf, whose last-bit differences between host and device can flip an occlusion test:
    a slightly larger share of pixels is exempt"""
    r, sc, fd, dom = _flow_pair(R, orc)
    _world(r, fd)
    _start_tracer(r, dom)
    r.traceAll()
    fd.traces = r.readTraces()
    fd.current_timestep = fd.traces.shape[1]
    fd.iso = [(1, 0.0, 0), (0, 0.0, 0)]
    fd.ao = (1, 12.0)
    g_acc, _, o_acc, _, _ = _render_both(r, sc, fd, orc)
    _close(g_acc, o_acc, 0.01)
    fd.ao = (0, 0.0)
    plain, _, _ = sc.render(fd.to_orc(orc))
    assert (np.abs(plain[..., :3] - o_acc[..., :3]).max(axis=-1) > 0.02).sum() > 50, "AO changed nothing in the test frame"


def test_surface_setters_errors(R, orc):
    r, sc, fd, dom = _frame_pair(R, orc, "nb32", 1)
    with pytest.raises(Exception):
        r.updateContourPlanes([(1, 0, 0)] * 3, [0.5] * 3, [3, 0, 0], [1, 0, 0])     # channel out of range
    r.updateContourPlanes([(1, 0, 0)] * 3, [0.5] * 3, [3, 0, 0], [0, 0, 0])         # ... ignored while disabled


def test_c5_shaped_multichannel_dvr_and_streamlines(R, orc):
    """BASELINE.json configs[4] at 1/16 scale (synth.cloud_shaped: block-structured octree from a
    noise threshold, 4 levels, 3 velocity channels): ABR bit-exact, 3-channel DVR within 1e-3,
    RK4 streamlines within 1e-5 of the domain, and the frame with the cones drawn"""
    from owlexabrick_b200 import synth
    cells, dom = synth.cloud_shaped(16)
    fields = synth.cloud_velocity(cells, dom)
    r, sc, _ = make_pair(R, orc, cells, fields)
    regions, leaf = r.get_abr()
    assert leaf.tobytes() == sc.leaf.tobytes() and regions.tobytes() == sc.regions.tobytes()
    cam = R.default_camera([0] * 3, [dom] * 3, 128, 96)
    fd = scenes.FrameDesc(128, 96, cam, 3)
    _world(r, fd)
    fd.xf_domain = [(-1.2, 1.2)] * 3
    fd.opacity_scale = 0.1
    g_acc, g_rgba, o_acc, o_rgba, o_st = _render_both(r, sc, fd, orc)
    assert np.abs(g_acc[..., :3] - o_acc[..., :3]).max() <= RGBA_TOL
    assert o_acc[..., :3].max() > 0.3
    st = r.render_stats()
    for k in ("samples", "brickVisits", "cornerLoads", "segments"):      # the counters behind the algorithmic-bytes model
        assert abs(st[k] - o_st[k]) <= 1e-3 * o_st[k], (k, st[k], o_st[k])
    seeds = synth.random_points(100, [0.25 * dom] * 3, [0.75 * dom, 0.75 * dom, 0.5 * dom], seed=5)
    T = 50
    r.setTracerParams(len(seeds), T, dom / 64.0, (0, 1, 2))
    r.setTracerSeeds(seeds)
    r.setTracerEnabled(True)
    r.traceAll()
    g = r.readTraces()
    o = sc.trace(seeds, T, (0, 1, 2), dom / 64.0, [0, 0, 0], [dom] * 3)
    dead_g, dead_o = g[..., 0] >= 2e10, o[..., 0] >= 2e10
    assert (dead_g != dead_o).sum() <= 3
    m = ~dead_g & ~dead_o
    assert m.sum() > 0.5 * m.size and np.abs(g[m] - o[m]).max() <= 1e-5 * dom * 4
    fd.opacity_scale = 0.01                              # thin volume, or the cones inside it stay hidden
    _, _, o_thin, _, _ = _render_both(r, sc, fd, orc)
    fd.traces = g
    fd.current_timestep = T
    g_acc, _, o_acc2, _, _ = _render_both(r, sc, fd, orc)
    _close(g_acc, o_acc2)
    assert (np.abs(o_acc2[..., :3] - o_thin[..., :3]).max(axis=-1) > 0.05).sum() > 50, "no cones in the test frame"


@pytest.mark.parametrize("ao", [False, True])
def test_triangle_mesh_vs_oracle(R, orc, ao):
    """world-space triangle meshes (the `surfaces` of the OptixRenderer ctor): BVH traversal in the
    frame kernel against the oracle's brute-force closest hit; with AO the mesh and an iso-surface
    occlude each other"""
    r, sc, fd, dom = _frame_pair(R, orc, "nb64_off", 1)
    _world(r, fd)
    fd.opacity_scale = 0.03
    v1, i1 = scenes.uv_sphere([dom * 0.45, dom * 0.5, dom * 0.55], dom * 0.2, 40, 20)
    v2, i2 = scenes.uv_sphere([dom * 0.8, dom * 0.7, dom * 0.3], dom * 0.12, 16, 8)
    fd.mesh = (np.concatenate([v1, v2]), np.concatenate([i1, i2 + len(v1)]))
    if ao:
        fd.iso = [(1, 0.85, 0), (0, 0.0, 0)]
        fd.ao = (1, 10.0)
    g_acc, _, o_acc, _, _ = _render_both(r, sc, fd, orc)
    _close(g_acc, o_acc, 0.01 if ao else 0.003)
    fd.mesh = None                                       # removing the meshes gives the plain frame again
    g2, _, o2, _, _ = _render_both(r, sc, fd, orc)
    _close(g2, o2, 0.01 if ao else 0.003)
    assert (np.abs(o2[..., :3] - o_acc[..., :3]).max(axis=-1) > 0.05).sum() > 100, "no mesh in the test frame"
