"""Python mirror of the reference's host classes over the C ABI (include/exabrick_b200.h).

Names follow the reference: ExaBricks (exa/ExaBricks.h), OptixRenderer's methods
(exa/OptixRenderer.h:35-95) on class Renderer, SimpleCamera maths
(glutViewer/Camera.cpp:94-120, glutViewer/OWLViewer.cpp:81-109, exa/viewer.cpp:226-238).
All compute goes through libexabrick_b200.so; nothing here falls back to numpy.
"""
import ctypes as C
import numpy as np
from . import capi
from .capi import lib, check

BRICK_DT = np.dtype([("lower", "<i4", 3), ("size", "<i4", 3), ("level", "<i4"), ("begin", "<u4")])
REGION_DT = np.dtype([("lo", "<f4", 3), ("hi", "<f4", 3), ("vlo", "<f4"), ("vhi", "<f4"),
                      ("leafListBegin", "<i4"), ("leafListSize", "<i4"), ("fcw", "<f4")])
KD_DT = np.dtype([("split_or_first_prim", "<u4"), ("axis_bits", "<u4"), ("max_level", "<i4")])
NUM_XF_VALUES = 128


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f3(v):
    return np.ascontiguousarray(v, dtype=np.float32).reshape(3)


class ExaBricks:
    """exa::ExaBricks (+ the exaBuilder executable as ExaBricks.build)."""

    def __init__(self, handle):
        self.h = C.c_void_p(handle)

    def __del__(self):
        if getattr(self, "h", None):
            lib.exa_bricks_destroy(self.h)
            self.h = None

    @staticmethod
    def build(cells, builder=capi.BUILDER_SAH_ALIKE, max_leaf_width=127):
        cells = np.ascontiguousarray(cells, dtype=np.int32).reshape(-1, 4)
        out = C.c_void_p()
        check(lib.exa_bricks_build(_p(cells), len(cells), builder, max_leaf_width, C.byref(out)))
        return ExaBricks(out.value)

    @staticmethod
    def build_gpu(cells, builder=capi.BUILDER_SAH_ALIKE, max_leaf_width=127, device=0):
        """exaBuilder on the GPU (exa_bricks_build_gpu); .timings = seconds per stage"""
        cells = np.ascontiguousarray(cells, dtype=np.int32).reshape(-1, 4)
        out, tm = C.c_void_p(), (C.c_double * 8)()
        check(lib.exa_bricks_build_gpu(_p(cells), len(cells), builder, max_leaf_width, device, C.byref(out), tm))
        b = ExaBricks(out.value)
        b.timings = {"upload_s": tm[0], "dedup_s": tm[1], "levels_s": tm[2], "write_s": tm[3], "total_s": tm[4], "levels": int(tm[5]), "nodes": int(tm[6])}
        return b

    @staticmethod
    def load(file_name):
        out = C.c_void_p()
        check(lib.exa_bricks_load(file_name.encode(), C.byref(out)))
        return ExaBricks(out.value)

    @staticmethod
    def from_bytes(raw):
        buf = np.frombuffer(raw, dtype=np.uint8)
        out = C.c_void_p()
        check(lib.exa_bricks_from_memory(_p(buf), len(buf), C.byref(out)))
        return ExaBricks(out.value)

    def save(self, bricks_file, kd_file=None):
        check(lib.exa_bricks_save(self.h, bricks_file.encode(), kd_file.encode() if kd_file else None))

    def to_bytes(self):
        """bytes of the .bricks file (builder/builder.cpp:895-902)"""
        hdr, ids = self.headers, self.cell_ids
        out = bytearray()
        for b in hdr:
            n = int(b["size"][0]) * int(b["size"][1]) * int(b["size"][2])
            out += np.concatenate([b["size"], b["lower"], [b["level"]]]).astype("<i4").tobytes()
            out += ids[int(b["begin"]):int(b["begin"]) + n].tobytes()
        return bytes(out)

    @property
    def num_bricks(self):
        return int(lib.exa_bricks_count(self.h))

    @property
    def total_num_cells(self):
        return int(lib.exa_bricks_num_cells(self.h))

    @property
    def headers(self):
        n = self.num_bricks
        if n == 0:
            return np.zeros(0, BRICK_DT)
        return np.frombuffer(C.string_at(lib.exa_bricks_headers(self.h), 32 * n), dtype=BRICK_DT).copy()

    @property
    def cell_ids(self):
        n = self.total_num_cells
        if n == 0:
            return np.zeros(0, np.int32)
        buf = (C.c_int32 * n).from_address(lib.exa_bricks_cell_ids(self.h))
        return np.frombuffer(buf, dtype=np.int32).copy()

    @property
    def kd_nodes(self):
        n = int(lib.exa_bricks_kd_count(self.h))
        if n == 0:
            return np.zeros(0, KD_DT)
        return np.frombuffer(C.string_at(lib.exa_bricks_kd_nodes(self.h), 12 * n), dtype=KD_DT).copy()

    def get_bounds(self):
        lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
        check(lib.exa_bricks_get_bounds(self.h, _p(lo), _p(hi)))
        return lo, hi


def load_scalar_field(file_name, compat=True):
    """ScalarField::load -> (values, valueRange)"""
    ptr, n, rng = C.c_void_p(), C.c_size_t(), np.zeros(2, np.float32)
    check(lib.exa_scalar_field_load(file_name.encode(), int(compat), C.byref(ptr), C.byref(n), _p(rng)))
    vals = np.frombuffer(C.string_at(ptr, 4 * n.value), dtype=np.float32).copy()
    lib.exa_free(ptr)
    return vals, rng


class ColorMapper:
    """exa::ColorMapper (exa/ColorMapper.h:29-83): piece-wise linear colour map over control points (t, r, g, b).
    Constructors as in the reference: colours (equally spaced), control points, a "t,(r,g,b)" string, PNG bytes;
    ColorMapper.embedded(i) / ColorMapper.names() are the maps of the viewer (exa/viewer.cpp:506-542)."""

    def __init__(self, colors=None, control_points=None, string=None, png=None):
        if string is not None:
            cp = self._from_c(lambda p, n: lib.exa_colormap_parse_string(string.encode(), p, n))
        elif png is not None:
            buf = np.frombuffer(bytes(png), np.uint8)
            cp = self._from_c(lambda p, n: lib.exa_colormap_from_png(_p(buf), len(buf), p, n))
        elif control_points is not None:
            cp = np.asarray(control_points, np.float32).reshape(-1, 4)
            cp = cp[np.argsort(cp[:, 0], kind="stable")]
        else:
            colors = np.asarray(colors, np.float32).reshape(-1, 3)
            t = np.arange(len(colors), dtype=np.float32) / np.float32(max(len(colors) - 1, 1))
            cp = np.concatenate([t[:, None], colors], axis=1)
        self.values_ = np.ascontiguousarray(cp, dtype=np.float32)

    @staticmethod
    def _from_c(call):
        ptr, n = C.c_void_p(), C.c_int()
        check(call(C.byref(ptr), C.byref(n)))
        try:
            return np.frombuffer(C.string_at(ptr, 16 * n.value), np.float32).reshape(-1, 4).copy()
        finally:
            lib.exa_free(ptr)

    @staticmethod
    def names():
        out = []
        for i in range(lib.exa_colormap_embedded_count()):
            name, ptr, n = C.c_char_p(), C.c_void_p(), C.c_int()
            check(lib.exa_colormap_embedded(i, C.byref(name), C.byref(ptr), C.byref(n)))
            lib.exa_free(ptr)
            out.append(name.value.decode())
        return out

    @staticmethod
    def embedded(index):
        name = C.c_char_p()
        cm = ColorMapper(control_points=ColorMapper._from_c(lambda p, n: lib.exa_colormap_embedded(int(index), C.byref(name), p, n)))
        cm.name = name.value.decode()
        return cm

    def __call__(self, t):
        out = np.zeros(3, np.float32)
        check(lib.exa_colormap_eval(_p(self.values_), len(self.values_), float(t), _p(out)))
        return out

    def table(self, n=NUM_XF_VALUES):
        """the 128-entry colour table the viewer feeds to updateXF (viewer.cpp:552-566)"""
        return np.stack([self(i / float(n - 1)) for i in range(n)])


def load_xf(file_name):
    """.xf file of the viewer (`--xf`, exa/viewer.cpp:139-145 writes it): 128 float32 opacities"""
    a = np.fromfile(file_name, np.float32)
    if len(a) != NUM_XF_VALUES:
        raise RuntimeError("mismatching xf size!?")
    return a


def save_xf(file_name, opacities):
    a = np.ascontiguousarray(opacities, np.float32)
    if a.shape != (NUM_XF_VALUES,):
        raise RuntimeError("mismatching xf size!?")
    a.tofile(file_name)


class Config:
    """exa::Config::parseConfigFile"""

    def __init__(self, file_name, scalar_compat=True):
        out = C.c_void_p()
        check(lib.exa_config_parse(file_name.encode(), int(scalar_compat), C.byref(out)))
        self.h = out
        self.fields = []
        for i in range(lib.exa_config_num_fields(self.h)):
            n, rng, name = C.c_size_t(), np.zeros(2, np.float32), C.c_char_p()
            ptr = lib.exa_config_field(self.h, i, C.byref(n), _p(rng), C.byref(name))
            vals = np.frombuffer(C.string_at(ptr, 4 * n.value), dtype=np.float32).copy()
            self.fields.append({"name": name.value.decode(), "values": vals, "valueRange": rng})
        self.voxelSpaceTransform = np.zeros(12, np.float32)
        check(lib.exa_config_voxel_space_transform(self.h, _p(self.voxelSpaceTransform)))

    def __del__(self):
        if getattr(self, "h", None):
            lib.exa_config_destroy(self.h)
            self.h = None

    def bricks_handle(self):
        return lib.exa_config_bricks(self.h)

    def get_bounds(self):
        lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
        check(lib.exa_config_get_bounds(self.h, _p(lo), _p(hi)))
        return lo, hi

    def triangles(self):
        """Config::surfaces (`triangles <file>`, Config.cpp:165-171) -> (vertices float32[n,3], indices int32[m,3]),
        concatenated"""
        v, i, nv, nt = C.c_void_p(), C.c_void_p(), C.c_size_t(), C.c_size_t()
        check(lib.exa_config_triangles(self.h, C.byref(v), C.byref(nv), C.byref(i), C.byref(nt)))
        verts = np.frombuffer(C.string_at(v, nv.value * 12), np.float32).reshape(-1, 3).copy() if nv.value else np.zeros((0, 3), np.float32)
        idx = np.frombuffer(C.string_at(i, nt.value * 12), np.int32).reshape(-1, 3).copy() if nt.value else np.zeros((0, 3), np.int32)
        return verts, idx


def simple_camera(pos, at, up, fovy_deg, width, height):
    """pos, dir00, dirDu, dirDv exactly as the viewer derives them in fp32
    (glutViewer/Camera.cpp:94-120, OWLViewer.cpp:81-109, viewer.cpp:226-238)."""
    f = np.float32
    pos, at, up = (np.asarray(v, f) for v in (pos, at, up))

    def norm(v):
        return (v * (f(1) / np.sqrt(np.dot(v, v), dtype=f))).astype(f)
    d = (at - pos).astype(f)
    focal = np.sqrt(np.dot(d, d), dtype=f)
    vz = -norm(d)
    vx = norm(np.cross(up, vz).astype(f))
    vy = norm(np.cross(vz, vx).astype(f))
    aspect = f(width) / f(height)
    h = f(2) * np.tan(f(fovy_deg) * f(np.pi) / f(360), dtype=f) * focal
    vertical = (h * vy).astype(f)
    horizontal = (h * aspect * vx).astype(f)
    lower_left = (-focal * vz - f(.5) * vertical - f(.5) * horizontal).astype(f)
    return pos, lower_left, (horizontal / f(width)).astype(f), (vertical / f(height)).astype(f)


def raw_to_cells(volume, threshold, max_level, refinement=2):
    """exaRawToCells (tools/fromVolume/raw2cells.cpp): volume[z, y, x] of uint8 / float32 / float64 ->
    (cells int32[n,4], scalars float32[n]) in the tool's output order"""
    vol = np.ascontiguousarray(volume)
    fmt = {np.dtype(np.uint8): 0, np.dtype(np.float32): 1, np.dtype(np.float64): 2}.get(vol.dtype)
    if fmt is None or vol.ndim != 3:
        raise RuntimeError("unknown format!?")
    nz, ny, nx = vol.shape
    c, s, n = C.c_void_p(), C.c_void_p(), C.c_size_t()
    check(lib.exa_raw_to_cells(_p(vol), fmt, nx, ny, nz, float(threshold), int(max_level), int(refinement), C.byref(c), C.byref(s), C.byref(n)))
    try:
        cells = np.frombuffer(C.string_at(c, n.value * 16), np.int32).reshape(-1, 4).copy()
        scal = np.frombuffer(C.string_at(s, n.value * 4), np.float32).copy()
    finally:
        lib.exa_free(c)
        lib.exa_free(s)
    return cells, scal


def load_triangle_meshes(file_name):
    """TriangleMesh::load (exa/TriangleMesh.cpp:21-68) -> (vertices float32[n,3], indices int32[m,3]), all meshes
    of the file concatenated"""
    v, i, nv, nt = C.c_void_p(), C.c_void_p(), C.c_size_t(), C.c_size_t()
    check(lib.exa_triangle_mesh_load(file_name.encode(), C.byref(v), C.byref(nv), C.byref(i), C.byref(nt)))
    try:
        verts = np.frombuffer(C.string_at(v, nv.value * 12), np.float32).reshape(-1, 3).copy()
        idx = np.frombuffer(C.string_at(i, nt.value * 12), np.int32).reshape(-1, 3).copy()
    finally:
        lib.exa_free(v)
        lib.exa_free(i)
    return verts, idx


def default_camera(world_lo, world_hi, width, height, fovy=70.0):
    """viewer default: from = center + (-.3,.7,1)*span, at = center, up = y (viewer.cpp:1283-1295)"""
    lo, hi = np.asarray(world_lo, np.float32), np.asarray(world_hi, np.float32)
    center = (lo + hi) * np.float32(0.5)
    span = hi - lo
    pos = center + np.array([-.3, .7, 1.], np.float32) * span
    return simple_camera(pos, center, [0, 1, 0], fovy, width, height)


class Renderer:
    """exa::OptixRenderer, headless.  fields: list of float32 arrays indexed by cell ID."""

    def __init__(self, bricks, fields, device=0, layout=None, part=None):
        """part = (rank, num_ranks): sort-last with the data partitioned (exa_renderer_create_partitioned): only the
        bricks of this rank's kd subtree are uploaded"""
        self.bricks = bricks
        if layout is not None:
            check(lib.exa_set_scalar_layout(int(layout)))
        fields = [np.ascontiguousarray(f, dtype=np.float32) for f in fields]
        ptrs = (C.c_void_p * len(fields))(*[f.ctypes.data for f in fields])
        counts = (C.c_size_t * len(fields))(*[len(f) for f in fields])
        handle = bricks.h if isinstance(bricks, ExaBricks) else C.c_void_p(bricks)
        out = C.c_void_p()
        if part is not None:
            check(lib.exa_renderer_create_partitioned(handle, ptrs, counts, len(fields), device, int(part[0]), int(part[1]), C.byref(out)))
        else:
            check(lib.exa_renderer_create(handle, ptrs, counts, len(fields), device, C.byref(out)))
        if layout is not None:
            lib.exa_set_scalar_layout(capi.LAYOUT_AUTO)
        self.h = out
        self.num_fields = len(fields)
        self.fb_size = (0, 0)
        self._trace_shape = self._pending_shape = (1000, 100)      # OptixRenderer.h:155-165 defaults

    def build_times(self):
        """device ms of the construction stages: ABR geometry, scalar gather, dual-cell records, value ranges"""
        ms = (C.c_double * 4)()
        check(lib.exa_renderer_get_build_times(self.h, ms))
        return {"abr_geometry_ms": ms[0], "gather_ms": ms[1], "dual_records_ms": ms[2], "value_ranges_ms": ms[3]}

    def residency(self):
        nb, nc = C.c_int(), C.c_uint64()
        check(lib.exa_renderer_get_residency(self.h, C.byref(nb), C.byref(nc)))
        return nb.value, int(nc.value)

    def __del__(self):
        self.close()

    def close(self):
        if getattr(self, "h", None):
            lib.exa_renderer_destroy(self.h)
            self.h = None

    def scalar_layout(self):
        """(layout, bytes): capi.LAYOUT_FLAT or capi.LAYOUT_DUAL and the size of the scalar data the kernels read"""
        lay, nbytes = C.c_int(), C.c_uint64()
        check(lib.exa_renderer_get_scalar_layout(self.h, C.byref(lay), C.byref(nbytes)))
        return lay.value, int(nbytes.value)

    def set_stream(self, cuda_stream):
        check(lib.exa_renderer_set_stream(self.h, C.c_void_p(cuda_stream)))

    # ---- OptixRenderer interface ----
    def setVoxelSpaceTransform(self, xfm12):
        m = np.ascontiguousarray(xfm12, dtype=np.float32).reshape(12)
        check(lib.exa_renderer_set_voxel_space_transform(self.h, _p(m)))

    def resizeFrameBuffer(self, fb_pointer, size):
        check(lib.exa_renderer_resize_framebuffer(self.h, C.c_void_p(fb_pointer) if fb_pointer else None, int(size[0]), int(size[1])))
        self.fb_size = (int(size[0]), int(size[1]))

    def updateCamera(self, pos, dir00, dirDu, dirDv):
        a, b, c, d = _f3(pos), _f3(dir00), _f3(dirDu), _f3(dirDv)
        check(lib.exa_renderer_update_camera(self.h, _p(a), _p(b), _p(c), _p(d)))

    def updateXF(self, chan, opacities, color_map, xf_domain, xf_opacity_scale=.1):
        op = np.ascontiguousarray(opacities, dtype=np.float32)
        cm = np.ascontiguousarray(color_map, dtype=np.float32)
        if cm.shape != (NUM_XF_VALUES, 3) or op.shape != (NUM_XF_VALUES,):
            raise RuntimeError("mismatching xf size!?")          # OptixRenderer.cpp:383
        check(lib.exa_renderer_update_xf(self.h, chan, _p(op), _p(cm), float(xf_domain[0]), float(xf_domain[1]), float(xf_opacity_scale)))

    def updateIsoValues(self, iso_values, channels, enabled):
        v = np.ascontiguousarray(iso_values, dtype=np.float32)
        c = np.ascontiguousarray(channels, dtype=np.int32)
        e = np.ascontiguousarray(enabled, dtype=np.int32)
        check(lib.exa_renderer_update_iso_values(self.h, _p(v), _p(c), _p(e)))

    def updateFrameID(self, frame_id):
        check(lib.exa_renderer_update_frame_id(self.h, int(frame_id)))

    def updateDt(self, dt):
        check(lib.exa_renderer_update_dt(self.h, float(dt)))

    def setSpaceSkipping(self, enable):
        check(lib.exa_renderer_set_space_skipping(self.h, int(bool(enable))))

    def setGradientShadingDVR(self, enable):
        check(lib.exa_renderer_set_gradient_shading_dvr(self.h, int(bool(enable))))

    def setGradientShadingISO(self, enable):
        check(lib.exa_renderer_set_gradient_shading_iso(self.h, int(bool(enable))))

    def setClipBox(self, enable, lo=None, hi=None):
        check(lib.exa_renderer_set_clip_box(self.h, int(bool(enable)), _p(_f3(lo)) if lo is not None else None, _p(_f3(hi)) if hi is not None else None))

    def updateContourPlanes(self, normals, offsets, channels, enabled):
        """OptixRenderer::updateContourPlanes (exa/OptixRenderer.cpp:504-531); arrays of MAX_CONTOUR_PLANES = 3"""
        n = np.ascontiguousarray(normals, dtype=np.float32).reshape(3, 3)
        o = np.ascontiguousarray(offsets, dtype=np.float32)
        c = np.ascontiguousarray(channels, dtype=np.int32)
        e = np.ascontiguousarray(enabled, dtype=np.int32)
        check(lib.exa_renderer_update_contour_planes(self.h, _p(n), _p(o), _p(c), _p(e)))

    def setTriangleMeshes(self, vertices=None, indices=None):
        """the `surfaces` of the OptixRenderer constructor (exa/OptixRenderer.cpp:556-608): world-space
        triangle meshes, concatenated; None removes them"""
        if vertices is None or indices is None or len(indices) == 0:
            check(lib.exa_renderer_set_triangle_meshes(self.h, None, 0, None, 0))
            return
        v = np.ascontiguousarray(vertices, dtype=np.float32).reshape(-1, 3)
        i = np.ascontiguousarray(indices, dtype=np.int32).reshape(-1, 3)
        check(lib.exa_renderer_set_triangle_meshes(self.h, _p(v), len(v), _p(i), len(i)))

    def setAO(self, enable, length=0.0):
        """frameState.ao (programs/FrameState.h:60-63; exa/viewer.cpp:953-954)"""
        check(lib.exa_renderer_set_ao(self.h, int(bool(enable)), float(length)))

    def setGradientMode(self, central_differences):
        """gradients by central differences (the reference's ANALYTIC_GRADIENTS=0 build) instead of the analytic ones"""
        check(lib.exa_renderer_set_gradient_mode(self.h, capi.GRADIENTS_CENTRAL_DIFFERENCES if central_differences else capi.GRADIENTS_ANALYTIC))

    def setClockScale(self, scale):
        """frameState.clockScale (viewer.cpp:956): > 0 = per-pixel cycle heat-map in the red channel"""
        check(lib.exa_renderer_set_clock_scale(self.h, float(scale)))

    def setErtThreshold(self, t):
        check(lib.exa_renderer_set_ert_threshold(self.h, float(t)))

    def setTracerEnabled(self, enable):
        check(lib.exa_renderer_set_tracer_enabled(self.h, int(bool(enable))))

    def setTracerParams(self, num_traces, num_timesteps, steplen, channels=(0, 1, 2), seed_lo=(.3, .3, .5), seed_hi=(.8, .8, .5)):
        ch = np.ascontiguousarray(channels, dtype=np.int32)
        check(lib.exa_renderer_set_tracer_params(self.h, num_traces, num_timesteps, float(steplen), _p(ch), _p(_f3(seed_lo)), _p(_f3(seed_hi))))
        self._pending_shape = (num_traces, num_timesteps)             # takes effect where the trace buffer is rebuilt

    def resetTracer(self):
        check(lib.exa_renderer_reset_tracer(self.h))
        self._trace_shape = self._pending_shape

    def setTracerSeeds(self, seeds):
        s = np.ascontiguousarray(seeds, dtype=np.float32).reshape(-1, 3)
        check(lib.exa_renderer_set_tracer_seeds(self.h, _p(s), len(s)))
        self._trace_shape = self._pending_shape = (len(s), self._pending_shape[1])

    def advanceTracer(self):
        rebuilt = C.c_int()
        check(lib.exa_renderer_advance_tracer(self.h, C.byref(rebuilt)))
        return bool(rebuilt.value)

    def traceAll(self):
        check(lib.exa_renderer_trace_all(self.h))

    def load_traces(self, traces, current_timestep):
        """trace vertices computed elsewhere: float32[numTraces, numTimesteps, 3] + the time step reached"""
        t = np.ascontiguousarray(traces, dtype=np.float32)
        check(lib.exa_renderer_set_traces(self.h, _p(t), t.shape[0], t.shape[1], int(current_timestep)))
        self._trace_shape = self._pending_shape = (t.shape[0], t.shape[1])

    def write_accum(self, accum):
        a = np.ascontiguousarray(accum, dtype=np.float32)
        check(lib.exa_renderer_write_accum(self.h, _p(a)))

    def readTraces(self):
        out = np.zeros(self._trace_shape + (3,), np.float32)
        check(lib.exa_renderer_read_traces(self.h, _p(out)))
        return out

    def render(self, sync=True):
        check(lib.exa_renderer_render(self.h) if sync else lib.exa_renderer_render_async(self.h))

    def sync(self):
        check(lib.exa_renderer_sync(self.h))

    # ---- headless extras ----
    def read_rgba8(self):
        out = np.zeros((self.fb_size[1], self.fb_size[0]), np.uint32)
        check(lib.exa_renderer_read_rgba8(self.h, _p(out)))
        return out

    def read_rgba8_async(self, pinned_ptr):
        """enqueue the copy of the frame just rendered into PINNED host memory (raw pointer);
        the next render overlaps it; sync() waits for both"""
        check(lib.exa_renderer_read_rgba8_async(self.h, C.c_void_p(pinned_ptr)))

    def read_accum(self):
        out = np.zeros((self.fb_size[1], self.fb_size[0], 4), np.float32)
        check(lib.exa_renderer_read_accum(self.h, _p(out)))
        return out

    def device_buffers(self):
        a, b = C.c_void_p(), C.c_void_p()
        check(lib.exa_renderer_device_buffers(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def render_stats(self):
        st = capi.ExaStats()
        check(lib.exa_renderer_render_stats(self.h, C.byref(st)))
        return st.as_dict()

    def get_abr(self):
        nr, nl = C.c_size_t(), C.c_size_t()
        check(lib.exa_renderer_get_abr(self.h, None, C.byref(nr), None, C.byref(nl)))
        regions = np.zeros(nr.value, REGION_DT)
        leaf = np.zeros(nl.value, np.int32)
        check(lib.exa_renderer_get_abr(self.h, _p(regions), C.byref(nr), _p(leaf), C.byref(nl)))
        return regions, leaf

    def kd_info(self):
        n, d = C.c_size_t(), C.c_int()
        check(lib.exa_renderer_get_kd_info(self.h, C.byref(n), C.byref(d)))
        return n.value, d.value

    def launch_params(self):
        lp, fs = capi.ExaLaunchParams(), capi.ExaFrameState()
        check(lib.exa_renderer_get_launch_params(self.h, C.byref(lp), C.byref(fs)))
        return bytes(lp.raw), bytes(fs.raw)

    def launch_count(self):
        return int(lib.exa_renderer_launch_count(self.h))

    def sample_points(self, xyz, channel=0, grad=False, out=None):
        """out = (values, valid[, grad]): arrays to fill (a caller in a loop reuses them; fresh arrays cost a page fault per
        4 KB on top of the query)"""
        xyz = np.ascontiguousarray(xyz, dtype=np.float32).reshape(-1, 3)
        n = len(xyz)
        if out is not None:
            val, ok, g = out[0], out[1], (out[2] if grad else None)
        else:
            val, ok = np.zeros(n, np.float32), np.zeros(n, np.uint8)
            g = np.zeros((n, 3), np.float32) if grad else None
        check(lib.exa_sample_points_host(self.h, _p(xyz), n, channel, _p(val), _p(ok), _p(g)))
        return (val, ok, g) if grad else (val, ok)

    def sample_points_noalloc(self, xyz, channel=0, grad=False, stats=False, out=None):
        """exa_sample_points_host_noalloc: at most capi.NOALLOC_MAX_POINTS points, device staging in the library's static
        scratch array; `out` = (values, valid[, grad]) arrays to fill (so that the call allocates nothing on the host either)"""
        xyz = np.ascontiguousarray(xyz, dtype=np.float32).reshape(-1, 3)
        n = len(xyz)
        if out is None:
            out = (np.zeros(n, np.float32), np.zeros(n, np.uint8)) + ((np.zeros((n, 3), np.float32),) if grad else ())
        val, ok, g = out[0], out[1], (out[2] if grad else None)
        st = capi.ExaStats() if stats else None
        check(lib.exa_sample_points_host_noalloc(self.h, _p(xyz), n, channel, _p(val), _p(ok), _p(g), C.byref(st) if stats else None))
        res = (val, ok, g) if grad else (val, ok)
        return res + (st.as_dict(),) if stats else res

    def sample_points_device(self, d_xyz, n, channel, d_values, d_valid, d_grad=None):
        check(lib.exa_sample_points_device(self.h, C.c_void_p(d_xyz), n, channel, C.c_void_p(d_values),
                                           C.c_void_p(d_valid) if d_valid else None, C.c_void_p(d_grad) if d_grad else None))

    def sample_points_stats(self, d_xyz, n, channel=0):
        st = capi.ExaStats()
        check(lib.exa_sample_points_stats(self.h, C.c_void_p(d_xyz), n, channel, C.byref(st)))
        return st.as_dict()

    # ---- multi-GPU ----
    def set_partition(self, mode, rank, num_ranks):
        check(lib.exa_renderer_set_partition(self.h, mode, rank, num_ranks))

    def render_partial(self, d_partial, d_surf=None):
        check(lib.exa_renderer_render_partial(self.h, C.c_void_p(d_partial), C.c_void_p(d_surf) if d_surf else None))

    def partition_order(self):
        """front-to-back order of the sort-last parts for the current camera"""
        order, n = (C.c_int * 16)(), C.c_int()
        check(lib.exa_renderer_partition_order(self.h, order, C.byref(n)))
        return list(order)[:max(n.value, 0)], n.value

    def composite_partials(self, d_partials, d_surfs, first, count, d_rgba8):
        """d_partials / d_surfs: device pointers of all parts' buffers, front-most first"""
        n = len(d_partials)
        pa = (C.c_void_p * n)(*d_partials)
        su = (C.c_void_p * n)(*d_surfs) if d_surfs else None
        check(lib.exa_renderer_composite_partials(self.h, pa, su, n, first, count, C.c_void_p(d_rgba8)))


class RendererGroup:
    """exa_group_*: N GPUs of one node driven by this one host thread (image tiles over a replicated scene, or sort-last
    over a partitioned one).  .renderers[i] are ordinary Renderer objects (not owning) for the per-device setters."""

    def __init__(self, bricks, fields, devices, mode=capi.PARTITION_IMAGE_TILES):
        fields = [np.ascontiguousarray(f, dtype=np.float32) for f in fields]
        ptrs = (C.c_void_p * len(fields))(*[f.ctypes.data for f in fields])
        counts = (C.c_size_t * len(fields))(*[len(f) for f in fields])
        dev = (C.c_int * len(devices))(*devices)
        out = C.c_void_p()
        check(lib.exa_group_create(bricks.h, ptrs, counts, len(fields), dev, len(devices), mode, C.byref(out)))
        self.h, self.bricks, self.size = out, bricks, (0, 0)
        self.renderers = []
        for i in range(lib.exa_group_size(self.h)):
            r = Renderer.__new__(Renderer)
            r.h, r.bricks, r.num_fields, r.fb_size = None, bricks, len(fields), (0, 0)
            r._trace_shape = r._pending_shape = (1000, 100)
            r.h = C.c_void_p(lib.exa_group_renderer(self.h, i))
            r.close = lambda: None                                    # owned by the group
            self.renderers.append(r)

    def close(self):
        if getattr(self, "h", None):
            for r in self.renderers:
                r.h = None
            lib.exa_group_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def rebalance(self, rounds=2):
        """exa_group_rebalance: image tiles dealt by measured cost; returns max / mean device load of the new deal"""
        ratio = C.c_double(1.0)
        check(lib.exa_group_rebalance(self.h, int(rounds), C.byref(ratio)))
        return ratio.value

    def apply(self, fd):
        """a tests/scenes.FrameDesc on every device (the frame buffers belong to the group)"""
        if self.size != (fd.width, fd.height):
            check(lib.exa_group_resize_framebuffer(self.h, fd.width, fd.height))
            self.size = (fd.width, fd.height)
        for r in self.renderers:
            r.fb_size = self.size
            fd.apply(r)

    def render(self, out=None):
        """one frame -> uint32[H, W] host array"""
        if out is None:
            out = np.zeros((self.size[1], self.size[0]), np.uint32)
        check(lib.exa_group_render(self.h, _p(out)))
        return out
