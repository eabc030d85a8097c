"""ctypes front-end of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module (see oracle/oracle.h).  The product package never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_DIR = os.path.join(HERE, "_ref")


def build():
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])


def _load():
    if not os.path.exists(LIB_PATH):
        build()
    return C.CDLL(LIB_PATH)


BRICK_DT = np.dtype([("lower", "<i4", 3), ("size", "<i4", 3), ("level", "<i4"), ("begin", "<u4")])
REGION_DT = np.dtype([("lo", "<f4", 3), ("hi", "<f4", 3), ("vlo", "<f4"), ("vhi", "<f4"),
                      ("leafListBegin", "<i4"), ("leafListSize", "<i4"), ("fcw", "<f4")])
assert BRICK_DT.itemsize == 32 and REGION_DT.itemsize == 44


class OrcFrame(C.Structure):
    _fields_ = [
        ("cam_pos", C.c_float * 3), ("cam_dir00", C.c_float * 3), ("cam_du", C.c_float * 3), ("cam_dv", C.c_float * 3),
        ("xf_vx", C.c_float * 3), ("xf_vy", C.c_float * 3), ("xf_vz", C.c_float * 3), ("xf_p", C.c_float * 3),
        ("width", C.c_int32), ("height", C.c_int32), ("frameID", C.c_int32), ("dt", C.c_float),
        ("numPrimaryChannels", C.c_int32), ("colormapChannel", C.c_int32),
        ("gradientShadingDVR", C.c_int32), ("gradientShadingISO", C.c_int32), ("spaceSkipping", C.c_int32),
        ("xfOpacityScale", C.c_float), ("xfDomain", (C.c_float * 2) * 10), ("xf", ((C.c_float * 4) * 128) * 10),
        ("isoEnabled", C.c_int32 * 2), ("isoValue", C.c_float * 2), ("isoChannel", C.c_int32 * 2),
        ("clipEnabled", C.c_int32), ("clipLo", C.c_float * 3), ("clipHi", C.c_float * 3),
        ("ertThreshold", C.c_float), ("texFilter8bit", C.c_int32), ("partial", C.c_int32),
        ("numTraces", C.c_int32), ("numTimesteps", C.c_int32), ("currentTimestep", C.c_int32),
        ("traces", C.POINTER(C.c_float)),
        ("cpEnabled", C.c_int32 * 3), ("cpNormal", (C.c_float * 3) * 3), ("cpOffset", C.c_float * 3), ("cpChannel", C.c_int32 * 3),
        ("aoEnabled", C.c_int32), ("aoLength", C.c_float),
        ("worldLo", C.c_float * 3), ("worldHi", C.c_float * 3),
        ("numMeshVerts", C.c_int32), ("numMeshTris", C.c_int32),
        ("meshVerts", C.POINTER(C.c_float)), ("meshIndex", C.POINTER(C.c_int32)),
        ("centralDifferences", C.c_int32),
    ]


class OrcStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("brickVisits", C.c_uint64), ("cornerLoads", C.c_uint64), ("segments", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = _load()
        _lib.orc_scene_create.restype = C.c_void_p
        _lib.orc_free.argtypes = [C.c_void_p]
        _lib.orc_scene_destroy.argtypes = [C.c_void_p]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def num_threads():
    return lib().orc_num_threads()


def parse_bricks(buf):
    """.bricks bytes -> (bricks[BRICK_DT], cellIDs int32) (exa/ExaBricks.cpp:27-38)."""
    buf = np.frombuffer(buf, dtype=np.uint8)
    bricks, ids, off, begin = [], [], 0, 0
    while off + 28 <= len(buf):
        hdr = np.frombuffer(buf[off:off + 28].tobytes(), dtype="<i4")
        n = int(hdr[0]) * int(hdr[1]) * int(hdr[2])
        off += 28
        ids.append(np.frombuffer(buf[off:off + 4 * n].tobytes(), dtype="<i4"))
        off += 4 * n
        bricks.append((tuple(hdr[3:6]), tuple(hdr[0:3]), int(hdr[6]), begin))
        begin += n
    b = np.array(bricks, dtype=BRICK_DT) if bricks else np.zeros(0, BRICK_DT)
    i = np.concatenate(ids) if ids else np.zeros(0, np.int32)
    return b, i


def build_bricks(cells, max_leaf_width=127):
    cells = np.ascontiguousarray(cells, dtype=np.int32).reshape(-1, 4)
    out = C.c_void_p()
    n = C.c_int64()
    rc = lib().orc_build_bricks(_p(cells), C.c_int64(len(cells)), C.c_int(max_leaf_width), C.byref(out), C.byref(n))
    if rc:
        raise RuntimeError("orc_build_bricks failed")
    raw = C.string_at(out, n.value)
    lib().orc_free(out)
    return raw


def flatten_scalars(cell_ids, fields):
    """scalar[i] = field[cellID[i]] per channel (exa/OptixRenderer.cpp:103-132)."""
    return np.ascontiguousarray(np.stack([np.asarray(f, np.float32)[cell_ids] for f in fields]))


def build_abr(bricks, scalars):
    bricks = np.ascontiguousarray(bricks)
    scalars = np.ascontiguousarray(scalars, dtype=np.float32)
    nf, nc = scalars.shape
    rp, lp, nr, nl = C.c_void_p(), C.c_void_p(), C.c_int64(), C.c_int64()
    lib().orc_build_abr(_p(bricks), C.c_int64(len(bricks)), _p(scalars), C.c_int64(nc), C.c_int(nf),
                        C.byref(rp), C.byref(nr), C.byref(lp), C.byref(nl))
    regions = np.frombuffer(C.string_at(rp, nr.value * 44), dtype=REGION_DT).copy()
    leaf = np.frombuffer(C.string_at(lp, nl.value * 4), dtype=np.int32).copy()
    lib().orc_free(rp)
    lib().orc_free(lp)
    return regions, leaf


def parse_abr(raw):
    """format written by oracle/ref_abr_main.cpp"""
    nr, nl = np.frombuffer(raw[:8], dtype="<i4")
    regions = np.frombuffer(raw[8:8 + 44 * nr], dtype=REGION_DT).copy()
    leaf = np.frombuffer(raw[8 + 44 * nr:8 + 44 * nr + 4 * nl], dtype="<i4").copy()
    return regions, leaf


class Scene:
    def __init__(self, bricks, scalars, regions, leaf):
        self.bricks = np.ascontiguousarray(bricks)
        self.scalars = np.ascontiguousarray(scalars, dtype=np.float32)
        self.regions = np.ascontiguousarray(regions)
        self.leaf = np.ascontiguousarray(leaf, dtype=np.int32)
        nf, nc = self.scalars.shape
        self.h = C.c_void_p(lib().orc_scene_create(_p(self.bricks), C.c_int64(len(self.bricks)), _p(self.scalars),
                                                   C.c_int64(nc), C.c_int(nf), _p(self.regions),
                                                   C.c_int64(len(self.regions)), _p(self.leaf), C.c_int64(len(self.leaf))))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_scene_destroy(self.h)
            self.h = None

    def sample(self, pts, channel=0, grad=False):
        pts = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
        n = len(pts)
        val = np.zeros(n, np.float32)
        ok = np.zeros(n, np.uint8)
        g = np.zeros((n, 3), np.float32) if grad else None
        st = OrcStats()
        lib().orc_sample_points(self.h, _p(pts), C.c_int64(n), C.c_int(channel), _p(val), _p(ok),
                                _p(g) if grad else None, C.byref(st))
        return (val, ok, g, st.as_dict()) if grad else (val, ok, st.as_dict())

    def render(self, frame, window=None, accum=None):
        W, H = frame.width, frame.height
        x0, y0, x1, y1 = window if window else (0, 0, W, H)
        if accum is None:
            accum = np.zeros((H, W, 4), np.float32)
        rgba = np.zeros((H, W), np.uint32)
        st = OrcStats()
        lib().orc_render(self.h, C.byref(frame), x0, y0, x1, y1, _p(accum), _p(rgba), C.byref(st))
        return accum, rgba, st.as_dict()

    def trace(self, seeds, num_timesteps, channels, steplen, wlo, whi):
        seeds = np.asarray(seeds, np.float32).reshape(-1, 3)
        tr = np.zeros((len(seeds), num_timesteps, 3), np.float32)
        tr[:, 0] = seeds
        ch = (C.c_int32 * 3)(*channels)
        lo = (C.c_float * 3)(*wlo)
        hi = (C.c_float * 3)(*whi)
        lib().orc_trace_streamlines(self.h, _p(tr), len(seeds), num_timesteps, ch, C.c_float(steplen), lo, hi)
        return tr


# ---- the reference's own binaries (oracle/_ref), when present ---------------------------
def have_ref():
    return all(os.path.exists(os.path.join(REF_DIR, b)) for b in ("exaBuilder", "ref_abr", "exaArtificial"))


def ref_build_bricks(cells, workdir, max_leaf_width=None):
    """run the reference's exaBuilder (builder/builder.cpp) on a cells array"""
    cf = os.path.join(workdir, "in.cells")
    bf = os.path.join(workdir, "out.bricks")
    np.ascontiguousarray(cells, dtype=np.int32).tofile(cf)
    cmd = [os.path.join(REF_DIR, "exaBuilder"), cf, "-o", bf]
    if max_leaf_width:
        cmd += ["--max-leaf-width", str(max_leaf_width)]
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
    return open(bf, "rb").read()


def have_ref_shader():
    return os.path.exists(os.path.join(REF_DIR, "ref_shader"))


def have_ref_shader_cd():
    return os.path.exists(os.path.join(REF_DIR, "ref_shader_cd"))


def ref_render(bricks_raw, fields, frame, workdir, accum=None, window=None, binary="ref_shader"):
    """one frame through the reference's OWN device source programs/exabrick.cu compiled as C++
    (oracle/ref_shader_main.cpp + ref_shim/shader_shim.h) -> (accum float32[H,W,4], rgba8 uint32[H,W]);
    binary = "ref_shader_cd": the build with -DANALYTIC_GRADIENTS=0"""
    bf = os.path.join(workdir, "s.bricks")
    open(bf, "wb").write(bricks_raw)
    sfs = []
    for i, f in enumerate(fields):
        p = os.path.join(workdir, "s%d.scalars" % i)
        np.ascontiguousarray(f, dtype=np.float32).tofile(p)
        sfs.append(p)
    ff, of = os.path.join(workdir, "frame.bin"), os.path.join(workdir, "out.bin")
    with open(ff, "wb") as fh:
        fh.write(bytes(frame))
        if frame.numTraces > 0 and frame.traces:
            n = frame.numTraces * frame.numTimesteps * 3
            fh.write(np.ctypeslib.as_array(frame.traces, shape=(n,)).astype(np.float32).tobytes())
        if frame.numMeshVerts > 0 and frame.numMeshTris > 0:
            fh.write(np.ctypeslib.as_array(frame.meshVerts, shape=(frame.numMeshVerts * 3,)).astype(np.float32).tobytes())
            fh.write(np.ctypeslib.as_array(frame.meshIndex, shape=(frame.numMeshTris * 3,)).astype(np.int32).tobytes())
    W, H = frame.width, frame.height
    if accum is not None:                       # progressive: the previous sum is read from out.bin
        with open(of, "wb") as fh:
            fh.write(np.ascontiguousarray(accum, np.float32).tobytes())
    elif os.path.exists(of):
        os.remove(of)
    env = dict(os.environ)
    if window:
        env["ORC_WINDOW"] = ",".join(str(int(x)) for x in window)
    subprocess.check_call([os.path.join(REF_DIR, binary), bf, ff, of] + sfs, stdout=subprocess.DEVNULL, env=env)
    raw = open(of, "rb").read()
    acc = np.frombuffer(raw[:W * H * 16], np.float32).reshape(H, W, 4).copy()
    rgba = np.frombuffer(raw[W * H * 16:W * H * 20], np.uint32).reshape(H, W).copy()
    return acc, rgba


def ref_sample_points(bricks_raw, fields, pts, channel, workdir):
    """tent-basis reconstruction (+ analytic gradient) at points by the reference's own samplePoint /
    samplePointWithDerivative behind its point query (programs/exabrick.cu:781-830, 883-942, compiled as C++)
    -> (value, valid uint8, grad[n,3])"""
    pts = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
    f = OrcFrame()
    f.width, f.height, f.numPrimaryChannels, f.dt = 1, 1, len(fields), 0.5
    f.xf_vx[0] = f.xf_vy[1] = f.xf_vz[2] = 1.0
    bf = os.path.join(workdir, "s.bricks")
    open(bf, "wb").write(bricks_raw)
    sfs = []
    for i, fl in enumerate(fields):
        p = os.path.join(workdir, "s%d.scalars" % i)
        np.ascontiguousarray(fl, dtype=np.float32).tofile(p)
        sfs.append(p)
    ff, of, pf = os.path.join(workdir, "frame.bin"), os.path.join(workdir, "pts.out"), os.path.join(workdir, "pts.in")
    open(ff, "wb").write(bytes(f))
    pts.tofile(pf)
    env = dict(os.environ, ORC_POINTS=pf, ORC_CHANNEL=str(int(channel)))
    subprocess.check_call([os.path.join(REF_DIR, "ref_shader"), bf, ff, of] + sfs, stdout=subprocess.DEVNULL, env=env)
    out = np.fromfile(of, np.float32).reshape(-1, 5)
    assert not (out[:, 1] == 2).any(), "samplePoint and samplePointWithDerivative disagree on the value"
    return out[:, 0].copy(), (out[:, 1] == 1).astype(np.uint8), out[:, 2:5].copy()


def ref_trace(bricks_raw, fields, seeds, num_timesteps, channels, steplen, wlo, whi, workdir):
    """RK4 advection by the reference's own computeTraces (programs/exabrick.cu:1531-1574, compiled as C++)"""
    seeds = np.asarray(seeds, np.float32).reshape(-1, 3)
    tr = np.zeros((len(seeds), num_timesteps, 3), np.float32)
    tr[:, 0] = seeds
    f = OrcFrame()
    f.width, f.height, f.numPrimaryChannels, f.dt = 1, 1, len(fields), 0.5
    f.xf_vx[0] = f.xf_vy[1] = f.xf_vz[2] = 1.0
    f.numTraces, f.numTimesteps, f.currentTimestep = len(seeds), num_timesteps, 0
    f.traces = tr.ctypes.data_as(C.POINTER(C.c_float))
    f.worldLo[:] = [float(x) for x in wlo]
    f.worldHi[:] = [float(x) for x in whi]
    os.environ["ORC_TRACE"] = "%r,%d,%d,%d" % (float(steplen), channels[0], channels[1], channels[2])
    try:
        bf = os.path.join(workdir, "s.bricks")
        open(bf, "wb").write(bricks_raw)
        sfs = []
        for i, fl in enumerate(fields):
            p = os.path.join(workdir, "s%d.scalars" % i)
            np.ascontiguousarray(fl, dtype=np.float32).tofile(p)
            sfs.append(p)
        ff, of = os.path.join(workdir, "frame.bin"), os.path.join(workdir, "traces.bin")
        with open(ff, "wb") as fh:
            fh.write(bytes(f))
            fh.write(tr.tobytes())
        subprocess.check_call([os.path.join(REF_DIR, "ref_shader"), bf, ff, of] + sfs, stdout=subprocess.DEVNULL)
        return np.fromfile(of, np.float32).reshape(tr.shape)
    finally:
        del os.environ["ORC_TRACE"]


def ref_build_abr(bricks_raw, fields, workdir):
    """run the reference's ExaBrickRegions::buildFrom (exa/Regions.cpp) via oracle/ref_abr_main.cpp"""
    bf = os.path.join(workdir, "x.bricks")
    open(bf, "wb").write(bricks_raw)
    sfs = []
    for i, f in enumerate(fields):
        p = os.path.join(workdir, "f%d.scalars" % i)
        np.ascontiguousarray(f, dtype=np.float32).tofile(p)
        sfs.append(p)
    of = os.path.join(workdir, "x.abr")
    subprocess.check_call([os.path.join(REF_DIR, "ref_abr"), bf, of] + sfs, stdout=subprocess.DEVNULL)
    return open(of, "rb").read()
