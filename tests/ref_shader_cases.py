"""Frames on which the CPU oracle is pinned against the reference's OWN device source
(programs/exabrick.cu compiled as plain C++: oracle/ref_shader_main.cpp + oracle/ref_shim/shader_shim.h).
Shared by tests/golden/make_golden.py (writes tests/golden/ref_shader_frames.npz from
oracle/_ref/ref_shader) and tests/test_oracle_pinned.py."""
import numpy as np

import scenes
from owlexabrick_b200.renderer import default_camera, simple_camera

W, H = 48, 32
OFF = (0, (1.0, 0.0, 0.0), 0.5, 0)


def _traces(dom):
    """a few helical traces through the volume, one with a dead tail"""
    T, n = 14, 5
    tr = np.zeros((n, T, 3), np.float32)
    for i in range(n):
        t = np.linspace(0, 1, T)
        tr[i, :, 0] = dom * (0.15 + 0.7 * t)
        tr[i, :, 1] = dom * (0.5 + 0.25 * np.sin(4 * t + i))
        tr[i, :, 2] = dom * (0.5 + 0.25 * np.cos(4 * t + i))
    tr[3, 9:] = 2e10
    return tr


def cases(extra=False):
    """-> name, scene kind, channels, max leaf width, configure(fd, dom).  extra=True adds the frames that pin the
    oracle only (tests/test_oracle_pinned.py); tests/test_gpu_pinned_cases.py renders the first 15."""
    def base(fd, dom):
        pass

    def nograd(fd, dom):
        fd.grad_dvr = False

    def thin(fd, dom):
        fd.opacity_scale = 0.03
        fd.dt = 0.3

    def noskip_noert(fd, dom):
        fd.space_skipping = False
        fd.ert = 2.0
        fd.opacity_scale = 0.05

    def partial_xf(fd, dom):
        a, rgb = fd.xf[0]
        a = a.copy()
        a[:70] = 0.0
        fd.xf[0] = (a, rgb)

    def clip_xfm(fd, dom):
        fd.xfm = np.array([2, 0, 0, 0, 2, 0, 0, 0, 2, 1.5, -2.0, 0.5], np.float32)    # world = (voxel - p) / 2
        fd.cam = default_camera([-0.75, 1.0, -0.25], [dom / 2 - 0.75, dom / 2 + 1.0, dom / 2 - 0.25], W, H)
        fd.world = ((-0.75, 1.0, -0.25), (dom / 2 - 0.75, dom / 2 + 1.0, dom / 2 - 0.25))
        fd.clip = ((2.0, 3.0, 1.0), (dom / 2 - 3.0, dom / 2 - 1.0, dom / 2 - 2.0))
        fd.opacity_scale = 0.1

    def inside(fd, dom):
        c = np.array([dom / 2] * 3, np.float32)
        fd.cam = simple_camera(c + np.array([1.0, 2.0, 3.0], np.float32), c + np.array([9.0, 4.0, -6.0], np.float32), [0, 1, 0], 70.0, W, H)
        fd.opacity_scale = 0.05

    def iso(fd, dom):
        fd.iso = [(1, 0.5, 0), (0, 0.0, 0)]
        fd.opacity_scale = 0.02

    def iso_flat(fd, dom):
        iso(fd, dom)
        fd.grad_iso = False

    def two_iso(fd, dom):
        fd.iso = [(1, 0.35, 0), (1, 0.65, 1)]
        fd.opacity_scale = 0.01

    def contour(fd, dom):
        fd.contour = [(1, (1.0, 0.0, 0.0), 0.45, 0), (1, (0.3, 1.0, 0.2), 0.6, 0), OFF]
        fd.opacity_scale = 0.03

    def cones(fd, dom):
        fd.traces = _traces(dom)
        fd.current_timestep = 12
        fd.opacity_scale = 0.01

    def cones_ao_iso(fd, dom):
        cones(fd, dom)
        fd.iso = [(1, 0.6, 0), (0, 0.0, 0)]
        fd.ao = (1, 9.0)

    def progressive(fd, dom):
        fd.frame_id = 3
        fd.opacity_scale = 0.2

    yield "dvr", "nb32", 1, 127, base
    yield "dvr_nograd", "nb32", 1, 127, nograd
    yield "dvr_thin_dt", "nb64_off", 1, 127, thin
    yield "dvr_2ch_small_bricks", "nb64_off", 2, 6, thin
    yield "dvr_noskip_noert", "nb32", 1, 127, noskip_noert
    yield "dvr_partial_xf", "nb32", 1, 127, partial_xf
    yield "dvr_clip_transform", "nb32", 1, 127, clip_xfm
    yield "dvr_camera_inside", "nb32", 1, 127, inside
    yield "iso", "nb64_off", 1, 127, iso
    yield "iso_flat", "nb32", 1, 127, iso_flat
    yield "two_iso_2ch", "nb32", 2, 127, two_iso
    yield "contour_planes", "nb64_off", 1, 127, contour
    yield "cones", "nb32", 1, 127, cones
    yield "cones_ao_iso", "nb32", 1, 127, cones_ao_iso
    yield "progressive", "nb32", 1, 127, progressive
    if not extra:
        return

    def xfm_world(fd, dom):
        fd.xfm = np.array([2, 0, 0, 0, 2, 0, 0, 0, 2, 1.5, -2.0, 0.5], np.float32)
        fd.cam = default_camera([-0.75, 1.0, -0.25], [dom / 2 - 0.75, dom / 2 + 1.0, dom / 2 - 0.25], W, H)
        fd.world = ((-0.75, 1.0, -0.25), (dom / 2 - 0.75, dom / 2 + 1.0, dom / 2 - 0.25))

    def contour_ch1_xfm(fd, dom):                        # plane classified by channel 1, world != voxel space
        xfm_world(fd, dom)
        fd.contour = [(1, (0.2, 1.0, 0.1), 0.55, 1), OFF, (1, (0.0, 0.0, 1.0), 0.35, 1)]
        fd.opacity_scale = 0.03

    def iso_ch1_ao(fd, dom):                             # AO with the iso-surface as the only occluder
        fd.iso = [(0, 0.0, 0), (1, 0.55, 1)]
        fd.ao = (1, 7.0)
        fd.opacity_scale = 0.02

    def contour_ao_clip(fd, dom):
        fd.contour = [(1, (1.0, 0.2, 0.0), 0.5, 0), OFF, OFF]
        fd.ao = (1, 5.0)
        fd.iso = [(1, 0.7, 0), (0, 0.0, 0)]
        fd.clip = ((dom * 0.1, dom * 0.15, dom * 0.05), (dom * 0.9, dom * 0.8, dom * 0.95))
        fd.opacity_scale = 0.02

    def cones_xfm_2ch(fd, dom):                          # cones live in world space
        xfm_world(fd, dom)
        fd.traces = ((_traces(dom) - np.array([1.5, -2.0, 0.5], np.float32)) / 2).astype(np.float32)
        fd.traces[3, 9:] = 2e10
        fd.current_timestep = 9
        fd.opacity_scale = 0.01

    def iso_inside(fd, dom):
        inside(fd, dom)
        fd.iso = [(1, 0.45, 0), (0, 0.0, 0)]
        fd.opacity_scale = 0.01

    def progressive7_dt(fd, dom):
        fd.frame_id = 7
        fd.dt = 0.8
        fd.grad_dvr = False
        fd.opacity_scale = 0.1

    def mesh(fd, dom):                                   # `surfaces` triangle meshes, closest-hit program SurfaceBVH
        v1, i1 = scenes.uv_sphere([dom * 0.45, dom * 0.5, dom * 0.55], dom * 0.2, 24, 12)
        v2, i2 = scenes.uv_sphere([dom * 0.8, dom * 0.7, dom * 0.3], dom * 0.12, 12, 6)
        fd.mesh = (np.concatenate([v1, v2]), np.concatenate([i1, i2 + len(v1)]))
        fd.opacity_scale = 0.03

    def mesh_ao_iso_cones(fd, dom):                      # every surface type at once, occluding each other
        mesh(fd, dom)
        cones(fd, dom)
        fd.iso = [(1, 0.8, 0), (0, 0.0, 0)]
        fd.contour = [OFF, (1, (0.0, 1.0, 0.3), 0.3, 0), OFF]
        fd.ao = (1, 8.0)
        fd.opacity_scale = 0.02

    def colormap_second(fd, dom):                        # !multiFieldDvr: iso on channel 0 coloured by channel 1 through XF 0
        fd.colormap_second = True
        fd.iso = [(1, 0.5, 0), (1, 0.75, 0)]
        fd.opacity_scale = 0.02

    yield "x_mesh", "nb64_off", 1, 127, mesh
    yield "x_mesh_ao_iso_cones_contour", "nb32", 1, 127, mesh_ao_iso_cones
    yield "x_contour_ch1_xfm", "nb32", 2, 127, contour_ch1_xfm
    yield "x_iso_ch1_ao", "nb64_off", 2, 127, iso_ch1_ao
    yield "x_contour_ao_clip", "nb32", 1, 127, contour_ao_clip
    yield "x_cones_xfm_2ch", "nb32", 2, 127, cones_xfm_2ch
    yield "x_iso_camera_inside", "nb64_off", 1, 127, iso_inside
    yield "x_progressive7_dt", "nb64_off", 1, 9, progressive7_dt
    yield "x_iso_colormap_second_channel", "nb32", 2, 127, colormap_second


CD_CASES = ["dvr", "dvr_thin_dt", "dvr_2ch_small_bricks", "dvr_partial_xf", "dvr_clip_transform", "dvr_camera_inside",
            "iso", "two_iso_2ch", "x_iso_ch1_ao", "x_iso_camera_inside"]


def cd_cases():
    """the frames with gradient shading among the cases above, with gradients by central differences (the reference
    built with -DANALYTIC_GRADIENTS=0, oracle/_ref/ref_shader_cd): name "cd_<case>" """
    for name, kind, channels, mlw, configure in cases(extra=True):
        if name not in CD_CASES:
            continue

        def cfg(fd, dom, configure=configure):
            configure(fd, dom)
            fd.central_differences = True
        yield "cd_" + name, kind, channels, mlw, cfg


def build(orc, kind, channels, mlw):
    cells, fields, dom = scenes.small_scene(kind, channels)
    raw = orc.build_bricks(cells, mlw)
    bricks, ids = orc.parse_bricks(raw)
    sc = orc.flatten_scalars(ids, fields)
    regions, leaf = orc.build_abr(bricks, sc)
    return raw, fields, dom, orc.Scene(bricks, sc, regions, leaf)


def frame(dom, channels, configure):
    fd = scenes.FrameDesc(W, H, default_camera([0] * 3, [dom] * 3, W, H), channels)
    fd.world = ((0.0,) * 3, (float(dom),) * 3)
    configure(fd, dom)
    return fd


def start_accum(name):
    """the accumulation buffer a progressive frame starts from (frameID > 0)"""
    rng = np.random.RandomState(5)
    a = rng.uniform(0, 3, (H, W, 4)).astype(np.float32)
    a[..., 3] = 1.0
    return a


def tracer_case():
    """swirling flow on the 4-level off-centre scene: cells, 3 velocity channels, seeds, T, steplen"""
    from owlexabrick_b200 import synth
    cells, _, dom = scenes.small_scene("nb64_off")
    c = cells[:, :3] + 0.5 * (1 << cells[:, 3:4])
    vx = (-(c[:, 1] - 32) * 0.05 + 0.2).astype(np.float32)
    vy = ((c[:, 0] - 32) * 0.05).astype(np.float32)
    vz = (0.3 + 0.002 * c[:, 2]).astype(np.float32)
    seeds = synth.random_points(64, [10, 10, 4], [54, 54, 20], seed=11)
    return cells, [vx, vy, vz], dom, seeds, 60, 1.3


def tracer_case_b():
    """same flow: permuted tracer channels, backward integration (negative step), world bounds inside the volume
    -> channels, steplen, lo, hi"""
    return (2, 0, 1), -0.7, [4, 6, 2], [60, 58, 40]
