"""Shared test helpers: the same scene / frame description fed to the CPU oracle (oracle/)
and to the CUDA path (through the C ABI)."""
import ctypes as C
import numpy as np

from owlexabrick_b200 import synth


def xf_ramp(alpha_scale=1.0, colors=None):
    """default viewer XF: alpha ramp i/127 (viewer.cpp:558-566) + a simple colour map"""
    t = np.arange(128, dtype=np.float32) / np.float32(127)
    alpha = (t * np.float32(alpha_scale)).astype(np.float32)
    if colors is None:
        rgb = np.stack([t, 1 - np.abs(2 * t - 1), 1 - t], axis=1).astype(np.float32)
    else:
        rgb = np.asarray(colors, np.float32)
    return alpha, rgb


class FrameDesc:
    """everything that defines one frame; applied to both implementations"""

    def __init__(self, width, height, cam, num_channels=1):
        self.width, self.height = width, height
        self.cam = cam                                  # pos, dir00, du, dv
        self.xfm = np.array([1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0], np.float32)
        self.frame_id = 0
        self.dt = 0.5
        self.num_channels = num_channels
        self.grad_dvr = True
        self.grad_iso = True
        self.space_skipping = True
        self.opacity_scale = 1.0
        self.xf = [xf_ramp() for _ in range(num_channels)]
        self.xf_domain = [(0.0, 1.0)] * num_channels
        self.iso = [(0, 0.0, 0), (0, 0.0, 0)]           # (enabled, value, channel)
        self.clip = None
        self.ert = 0.98
        # surfaces: contour planes (enabled, normal, offset, channel) x 3, AO (enabled, length),
        # streamline cones (traces[numTraces, numTimesteps, 3], currentTimestep), world bounds
        self.contour = [(0, (1.0, 0.0, 0.0), 0.5, 0)] * 3
        self.ao = (0, 0.0)
        self.traces = None
        self.current_timestep = 0
        self.world = ((0.0, 0.0, 0.0), (1.0, 1.0, 1.0))
        self.mesh = None                                 # (vertices float32[n,3], indices int32[m,3]), world space
        self.colormap_second = False                     # oracle only: the reference's !multiFieldDvr mode (OptixRenderer.cpp:277-283)
        self.central_differences = False                 # gradients as in the reference built with ANALYTIC_GRADIENTS=0

    def to_orc(self, orc):
        f = orc.OrcFrame()
        for name, v in zip(("cam_pos", "cam_dir00", "cam_du", "cam_dv"), self.cam):
            getattr(f, name)[:] = [float(x) for x in v]
        f.xf_vx[:] = [float(x) for x in self.xfm[0:3]]
        f.xf_vy[:] = [float(x) for x in self.xfm[3:6]]
        f.xf_vz[:] = [float(x) for x in self.xfm[6:9]]
        f.xf_p[:] = [float(x) for x in self.xfm[9:12]]
        f.width, f.height, f.frameID, f.dt = self.width, self.height, self.frame_id, self.dt
        f.numPrimaryChannels, f.colormapChannel = (1, 1) if self.colormap_second else (self.num_channels, 0)
        f.gradientShadingDVR, f.gradientShadingISO = int(self.grad_dvr), int(self.grad_iso)
        f.spaceSkipping = int(self.space_skipping)
        f.centralDifferences = int(self.central_differences)
        f.xfOpacityScale = self.opacity_scale
        for c in range(self.num_channels):
            f.xfDomain[c][0], f.xfDomain[c][1] = self.xf_domain[c]
            alpha, rgb = self.xf[c]
            tab = np.concatenate([rgb, alpha[:, None]], axis=1).astype(np.float32)
            C.memmove(C.addressof(f.xf[c]), tab.ctypes.data, 128 * 16)
        for i, (en, val, ch) in enumerate(self.iso):
            f.isoEnabled[i], f.isoValue[i], f.isoChannel[i] = int(en), float(val), int(ch)
        if self.clip is not None:
            f.clipEnabled = 1
            f.clipLo[:] = [float(x) for x in self.clip[0]]
            f.clipHi[:] = [float(x) for x in self.clip[1]]
        f.ertThreshold = self.ert
        f.texFilter8bit = 0
        f.partial = 0
        f.traces = None
        if self.traces is not None:
            tr = np.ascontiguousarray(self.traces, dtype=np.float32)
            f._keep = tr                                  # keep the buffer alive with the frame
            f.numTraces, f.numTimesteps, f.currentTimestep = tr.shape[0], tr.shape[1], int(self.current_timestep)
            f.traces = tr.ctypes.data_as(C.POINTER(C.c_float))
        if self.mesh is not None:
            mv = np.ascontiguousarray(self.mesh[0], dtype=np.float32).reshape(-1, 3)
            mi = np.ascontiguousarray(self.mesh[1], dtype=np.int32).reshape(-1, 3)
            f._keep_mesh = (mv, mi)
            f.numMeshVerts, f.numMeshTris = len(mv), len(mi)
            f.meshVerts = mv.ctypes.data_as(C.POINTER(C.c_float))
            f.meshIndex = mi.ctypes.data_as(C.POINTER(C.c_int32))
        for i, (en, nrm, off, ch) in enumerate(self.contour):
            n = np.asarray(nrm, np.float32)
            n = n * (np.float32(1) / np.sqrt(np.float32(n[0] * n[0] + n[1] * n[1] + n[2] * n[2])))   # OptixRenderer.cpp:511
            f.cpEnabled[i], f.cpOffset[i], f.cpChannel[i] = int(en), float(off), int(ch)
            f.cpNormal[i][:] = [float(x) for x in n]
        f.aoEnabled, f.aoLength = int(self.ao[0]), float(self.ao[1])
        f.worldLo[:] = [float(x) for x in self.world[0]]
        f.worldHi[:] = [float(x) for x in self.world[1]]
        return f

    def apply(self, r):
        r.setVoxelSpaceTransform(self.xfm)
        if r.fb_size != (self.width, self.height):
            r.resizeFrameBuffer(None, (self.width, self.height))
        r.updateCamera(*self.cam)
        for c in range(self.num_channels):
            alpha, rgb = self.xf[c]
            r.updateXF(c, alpha, rgb, self.xf_domain[c], self.opacity_scale)
        r.updateIsoValues([i[1] for i in self.iso], [i[2] for i in self.iso], [i[0] for i in self.iso])
        r.updateFrameID(self.frame_id)
        r.updateDt(self.dt)
        r.setSpaceSkipping(self.space_skipping)
        r.setGradientShadingDVR(self.grad_dvr)
        r.setGradientShadingISO(self.grad_iso)
        if self.clip is not None:
            r.setClipBox(True, self.clip[0], self.clip[1])
        else:
            r.setClipBox(False)
        r.setErtThreshold(self.ert)
        r.setGradientMode(self.central_differences)
        r.updateContourPlanes([c[1] for c in self.contour], [c[2] for c in self.contour], [c[3] for c in self.contour],
                              [c[0] for c in self.contour])
        r.setAO(self.ao[0], self.ao[1])
        if self.mesh is not None:
            r.setTriangleMeshes(self.mesh[0], self.mesh[1])
        else:
            r.setTriangleMeshes(None, None)


def uv_sphere(centre, radius, nu=24, nv=12):
    """triangle mesh of a sphere: (vertices float32[n,3], indices int32[m,3])"""
    th = np.linspace(0, np.pi, nv + 1)
    ph = np.linspace(0, 2 * np.pi, nu, endpoint=False)
    v = np.array([[np.sin(t) * np.cos(p), np.cos(t), np.sin(t) * np.sin(p)] for t in th for p in ph])
    v = (np.asarray(centre)[None, :] + radius * v).astype(np.float32)
    idx = []
    for j in range(nv):
        for i in range(nu):
            a, b = j * nu + i, j * nu + (i + 1) % nu
            c, d = a + nu, b + nu
            if j > 0:
                idx.append((a, b, c))
            if j < nv - 1:
                idx.append((b, d, c))
    return v, np.array(idx, np.int32)


def unpack_rgba8(img):
    img = np.asarray(img, np.uint32)
    return np.stack([(img >> s) & 0xff for s in (0, 8, 16, 24)], axis=-1).astype(np.int32)


def small_scene(kind="nb32", channels=1):
    """cells + fields for small parity scenes"""
    if kind == "nb32":
        cells = synth.nested_boxes(32, [(10, 22), (4, 28)])
        dom = 32
    elif kind == "nb64":
        cells = synth.nested_boxes(64, [(18, 46), (8, 56)])
        dom = 64
    elif kind == "nb64_off":              # off-centre refinement, 4 levels
        cells = synth.nested_boxes(64, [((8, 16, 24), (24, 32, 40)), ((4, 8, 16), (36, 40, 48)), ((0, 0, 8), (48, 48, 56))])
        dom = 64
    else:
        raise ValueError(kind)
    fields = [synth.analytic_field(cells, dom, c) for c in range(channels)]
    return cells, fields, dom


def random_amr(seed):
    """random octree AMR for the builder / ABR fuzz (tests/test_fuzz_vs_reference.py): domain 16^3 or 32^3 of
    level-1..3 root cells, each refined recursively with a random probability; optionally ragged (missing root
    cells), shifted to negative coordinates, shuffled, with duplicated cells -> cells int32[n,4], max leaf width,
    fields"""
    rng = np.random.RandomState(seed)
    dom_log, max_level = int(rng.choice([4, 5])), int(rng.choice([1, 2, 3]))
    cells = []

    def rec(x, y, z, l, p):
        if l > 0 and rng.rand() < p:
            h = 1 << (l - 1)
            for dz in (0, h):
                for dy in (0, h):
                    for dx in (0, h):
                        rec(x + dx, y + dy, z + dz, l - 1, p * 0.8)
        else:
            cells.append((x, y, z, l))
    top = 1 << max_level
    n = (1 << dom_log) // top
    holes = rng.rand() < 0.5
    for z in range(n):
        for y in range(n):
            for x in range(n):
                if holes and rng.rand() < 0.15:
                    continue
                rec(x * top, y * top, z * top, max_level, rng.uniform(0.2, 0.8))
    c = np.array(cells, np.int32)
    c[:, :3] += (rng.randint(-3, 4, 3) * top).astype(np.int32)
    c = c[rng.permutation(len(c))]
    if rng.rand() < 0.5:
        c = np.concatenate([c, c[rng.randint(0, len(c), 50)]])
    mlw = [127, 3, 8, 17][rng.randint(4)]
    fields = [rng.uniform(-1, 1, len(c)).astype(np.float32) for _ in range(rng.randint(1, 3))]
    return c, mlw, fields


def random_amr_frame(seed, W=48, H=32):
    """a frame over random_amr(seed) with smooth fields: DVR, + iso-surface (odd seeds), + contour plane (seed % 3 == 0),
    + AO (seed % 4 == 1) -> cells, max leaf width, fields, FrameDesc"""
    from owlexabrick_b200.renderer import default_camera
    cells, mlw, fields = random_amr(seed)
    c = cells[:, :3] + 0.5 * (1 << cells[:, 3:4])
    lo, hi = cells[:, :3].min(0).astype(np.float32), (cells[:, :3] + (1 << cells[:, 3:4])).max(0).astype(np.float32)
    u = (c - lo) / (hi - lo)
    fields = [(0.5 + 0.5 * np.sin(3 * u[:, 0] + seed) * np.cos(2.5 * u[:, 1]) * np.cos(2 * u[:, 2])).astype(np.float32) for _ in fields]
    fd = FrameDesc(W, H, default_camera(lo, hi, W, H), len(fields))
    fd.world = (tuple(float(x) for x in lo), tuple(float(x) for x in hi))
    fd.opacity_scale = 0.05
    if seed % 2:
        fd.iso = [(1, 0.5, 0), (0, 0.0, 0)]
    if seed % 3 == 0:
        fd.contour = [(1, (1.0, 0.3, 0.1), 0.5, 0), (0, (1.0, 0.0, 0.0), 0.5, 0), (0, (1.0, 0.0, 0.0), 0.5, 0)]
    if seed % 4 == 1:
        fd.ao = (1, 4.0)
    return cells, mlw, fields, fd


def random_amr_points(seed, n=20000):
    """points for the reconstruction fuzz: uniform over the bounds of random_amr(seed) grown by one unit (holes and
    the outside included), sampled on the last channel -> cells, max leaf width, fields, points, channel"""
    from owlexabrick_b200 import synth
    cells, mlw, fields, fd = random_amr_frame(seed)
    lo, hi = fd.world
    return cells, mlw, fields, synth.random_points(n, [x - 1 for x in lo], [x + 1 for x in hi], seed=seed), len(fields) - 1
