"""ctypes binding of libexabrick_b200.so (include/exabrick_b200.h).

The library is the product: importing this module fails loudly when it has not been built
(`python -m owlexabrick_b200.build`), and no function has a CPU fallback.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("EXA_LIB") or os.path.join(HERE, "libexabrick_b200.so")   # EXA_LIB: kernel-variant builds (tools/perf_quick.py)

if not os.path.exists(LIB_PATH):
    raise ImportError("%s is missing - build it with `python -m owlexabrick_b200.build` "
                      "(there is no CPU fallback)" % LIB_PATH)
lib = C.CDLL(LIB_PATH)

EXA_OK, EXA_ERR_INVALID, EXA_ERR_IO, EXA_ERR_CUDA, EXA_ERR_LIMIT = range(5)
BUILDER_SPATIAL_MEDIAN, BUILDER_SAH_ALIKE = 0, 1
PARTITION_NONE, PARTITION_SORT_LAST, PARTITION_IMAGE_TILES = 0, 1, 2
GRADIENTS_ANALYTIC, GRADIENTS_CENTRAL_DIFFERENCES = 0, 1
NOALLOC_MAX_POINTS = 65536          # EXA_NOALLOC_MAX_POINTS (include/exabrick_b200.h)
LAYOUT_AUTO, LAYOUT_FLAT, LAYOUT_DUAL = 0, 1, 2


class ExaError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("[exa error %d] %s" % (code, msg))
        self.code = code


class ExaStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("brickVisits", C.c_uint64), ("cornerLoads", C.c_uint64), ("segments", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


class ExaFrameState(C.Structure):
    _fields_ = [("raw", C.c_uint8 * 408)]


class ExaLaunchParams(C.Structure):
    _fields_ = [("raw", C.c_uint8 * 240)]


vp, sz, i32, f32 = C.c_void_p, C.c_size_t, C.c_int, C.c_float
_SIG = {
    "exa_last_error": (C.c_char_p, []),
    "exa_cuda_available": (i32, []),
    "exa_free": (None, [vp]),
    "exa_renderer_get_tile_cost_histogram": (i32, [vp, vp]),
    "exa_renderer_debug_counters": (i32, [vp, vp]),
    "exa_renderer_set_color_buffer": (i32, [vp, vp]),
    "exa_renderer_read_rgba8_async": (i32, [vp, vp]),
    "exa_device_malloc": (i32, [C.c_size_t, i32, C.POINTER(vp)]),
    "exa_device_free": (i32, [vp]),
    "exa_ipc_export": (i32, [vp, vp]),
    "exa_ipc_import": (i32, [vp, i32, C.POINTER(vp)]),
    "exa_ipc_close": (i32, [vp]),
    "exa_set_scalar_layout": (i32, [i32]),
    "exa_renderer_get_scalar_layout": (i32, [vp, C.POINTER(i32), C.POINTER(C.c_uint64)]),
    "exa_bricks_build": (i32, [vp, sz, i32, i32, C.POINTER(vp)]),
    "exa_bricks_build_gpu": (i32, [vp, sz, i32, i32, i32, C.POINTER(vp), vp]),
    "exa_bricks_load": (i32, [C.c_char_p, C.POINTER(vp)]),
    "exa_bricks_from_memory": (i32, [vp, sz, C.POINTER(vp)]),
    "exa_bricks_save": (i32, [vp, C.c_char_p, C.c_char_p]),
    "exa_bricks_destroy": (None, [vp]),
    "exa_bricks_count": (sz, [vp]),
    "exa_bricks_num_cells": (sz, [vp]),
    "exa_bricks_headers": (vp, [vp]),
    "exa_bricks_cell_ids": (vp, [vp]),
    "exa_bricks_get_bounds": (i32, [vp, vp, vp]),
    "exa_bricks_kd_count": (sz, [vp]),
    "exa_bricks_kd_nodes": (vp, [vp]),
    "exa_scalar_field_load": (i32, [C.c_char_p, i32, C.POINTER(vp), C.POINTER(sz), vp]),
    "exa_scalar_field_magnitude": (i32, [vp, vp, vp, sz, vp, vp]),
    "exa_colormap_eval": (i32, [vp, i32, f32, vp]),
    "exa_colormap_parse_string": (i32, [C.c_char_p, C.POINTER(vp), C.POINTER(i32)]),
    "exa_colormap_from_png": (i32, [vp, sz, C.POINTER(vp), C.POINTER(i32)]),
    "exa_colormap_embedded_count": (i32, []),
    "exa_colormap_embedded": (i32, [i32, C.POINTER(C.c_char_p), C.POINTER(vp), C.POINTER(i32)]),
    "exa_config_parse": (i32, [C.c_char_p, i32, C.POINTER(vp)]),
    "exa_config_destroy": (None, [vp]),
    "exa_config_bricks": (vp, [vp]),
    "exa_config_num_fields": (i32, [vp]),
    "exa_config_field": (vp, [vp, i32, C.POINTER(sz), vp, C.POINTER(C.c_char_p)]),
    "exa_config_voxel_space_transform": (i32, [vp, vp]),
    "exa_config_get_bounds": (i32, [vp, vp, vp]),
    "exa_renderer_create": (i32, [vp, vp, vp, i32, i32, C.POINTER(vp)]),
    "exa_renderer_destroy": (None, [vp]),
    "exa_renderer_set_stream": (i32, [vp, vp]),
    "exa_renderer_get_abr": (i32, [vp, vp, C.POINTER(sz), vp, C.POINTER(sz)]),
    "exa_renderer_get_kd_info": (i32, [vp, C.POINTER(sz), C.POINTER(i32)]),
    "exa_renderer_get_launch_params": (i32, [vp, vp, vp]),
    "exa_renderer_set_voxel_space_transform": (i32, [vp, vp]),
    "exa_renderer_resize_framebuffer": (i32, [vp, vp, i32, i32]),
    "exa_renderer_update_camera": (i32, [vp, vp, vp, vp, vp]),
    "exa_renderer_update_xf": (i32, [vp, i32, vp, vp, f32, f32, f32]),
    "exa_renderer_update_iso_values": (i32, [vp, vp, vp, vp]),
    "exa_renderer_update_frame_id": (i32, [vp, i32]),
    "exa_renderer_update_dt": (i32, [vp, f32]),
    "exa_renderer_set_space_skipping": (i32, [vp, i32]),
    "exa_renderer_set_gradient_shading_dvr": (i32, [vp, i32]),
    "exa_renderer_set_gradient_shading_iso": (i32, [vp, i32]),
    "exa_renderer_set_gradient_mode": (i32, [vp, i32]),
    "exa_renderer_set_clip_box": (i32, [vp, i32, vp, vp]),
    "exa_renderer_update_contour_planes": (i32, [vp, vp, vp, vp, vp]),
    "exa_renderer_set_ao": (i32, [vp, i32, f32]),
    "exa_renderer_set_clock_scale": (i32, [vp, C.c_float]),
    "exa_renderer_set_triangle_meshes": (i32, [vp, vp, sz, vp, sz]),
    "exa_scalar_field_expression": (i32, [vp, i32, sz, vp, i32, vp, vp]),
    "exa_raw_to_cells": (i32, [vp, i32, i32, i32, i32, f32, i32, i32, C.POINTER(vp), C.POINTER(vp), C.POINTER(sz)]),
    "exa_triangle_mesh_load": (i32, [C.c_char_p, C.POINTER(vp), C.POINTER(sz), C.POINTER(vp), C.POINTER(sz)]),
    "exa_config_triangles": (i32, [vp, C.POINTER(vp), C.POINTER(sz), C.POINTER(vp), C.POINTER(sz)]),
    "exa_renderer_set_ert_threshold": (i32, [vp, f32]),
    "exa_renderer_set_tracer_enabled": (i32, [vp, i32]),
    "exa_renderer_set_tracer_params": (i32, [vp, i32, i32, f32, vp, vp, vp]),
    "exa_renderer_reset_tracer": (i32, [vp]),
    "exa_renderer_set_tracer_seeds": (i32, [vp, vp, i32]),
    "exa_renderer_advance_tracer": (i32, [vp, C.POINTER(i32)]),
    "exa_renderer_trace_all": (i32, [vp]),
    "exa_renderer_read_traces": (i32, [vp, vp]),
    "exa_renderer_render": (i32, [vp]),
    "exa_renderer_render_async": (i32, [vp]),
    "exa_renderer_sync": (i32, [vp]),
    "exa_renderer_read_rgba8": (i32, [vp, vp]),
    "exa_renderer_read_accum": (i32, [vp, vp]),
    "exa_renderer_write_accum": (i32, [vp, vp]),
    "exa_renderer_set_traces": (i32, [vp, vp, i32, i32, i32]),
    "exa_renderer_device_buffers": (i32, [vp, C.POINTER(vp), C.POINTER(vp)]),
    "exa_sample_points_host": (i32, [vp, vp, sz, i32, vp, vp, vp]),
    "exa_sample_points_device": (i32, [vp, vp, sz, i32, vp, vp, vp]),
    "exa_renderer_render_stats": (i32, [vp, C.POINTER(ExaStats)]),
    "exa_sample_points_stats": (i32, [vp, vp, sz, i32, C.POINTER(ExaStats)]),
    "exa_renderer_launch_count": (C.c_uint64, [vp]),
    "exa_sample_points_host_noalloc": (i32, [vp, vp, sz, i32, vp, vp, vp, C.POINTER(ExaStats)]),
    "exa_renderer_set_partition": (i32, [vp, i32, i32, i32]),
    "exa_renderer_render_partial": (i32, [vp, vp, vp]),
    "exa_renderer_partition_order": (i32, [vp, vp, C.POINTER(i32)]),
    "exa_renderer_composite_partials": (i32, [vp, vp, vp, i32, sz, sz, vp]),
    "exa_renderer_create_partitioned": (i32, [vp, vp, vp, i32, i32, i32, i32, C.POINTER(vp)]),
    "exa_renderer_get_residency": (i32, [vp, C.POINTER(i32), C.POINTER(C.c_uint64)]),
    "exa_renderer_get_build_times": (i32, [vp, vp]),
    "exa_renderer_set_color_strips": (i32, [vp, vp, i32]),
    "exa_renderer_set_color_strips_rows": (i32, [vp, vp, vp, i32]),
    "exa_renderer_enable_tile_costs": (i32, [vp, i32]),
    "exa_renderer_get_tile_costs": (i32, [vp, vp, i32]),
    "exa_renderer_set_tile_list": (i32, [vp, vp, i32]),
    "exa_balance_tiles": (i32, [vp, i32, i32, i32, vp, C.POINTER(i32)]),
    "exa_group_create": (i32, [vp, vp, vp, i32, vp, i32, i32, C.POINTER(vp)]),
    "exa_group_destroy": (None, [vp]),
    "exa_group_size": (i32, [vp]),
    "exa_group_renderer": (vp, [vp, i32]),
    "exa_group_resize_framebuffer": (i32, [vp, i32, i32]),
    "exa_group_update_camera": (i32, [vp, vp, vp, vp, vp]),
    "exa_group_update_xf": (i32, [vp, i32, vp, vp, f32, f32, f32]),
    "exa_group_update_frame_id": (i32, [vp, i32]),
    "exa_group_set_voxel_space_transform": (i32, [vp, vp]),
    "exa_group_render": (i32, [vp, vp]),
    "exa_group_rebalance": (i32, [vp, i32, C.POINTER(C.c_double)]),
    "exa_peer_barrier": (i32, [vp, i32, i32, C.c_uint32, vp]),
    "exa_copy_async": (i32, [vp, vp, sz, vp]),
    "exa_host_register": (i32, [vp, sz]),
    "exa_host_unregister": (i32, [vp]),
    "exa_strip_rows": (i32, [i32, i32, i32, C.POINTER(i32), C.POINTER(i32)]),
}
EXPORTS = sorted(_SIG)
for _name, (_res, _args) in _SIG.items():
    _fn = getattr(lib, _name)          # AttributeError here = header/library mismatch
    _fn.restype = _res
    _fn.argtypes = _args


def last_error():
    return (lib.exa_last_error() or b"").decode("utf-8", "replace")


def check(rc):
    if rc != EXA_OK:
        raise ExaError(rc, last_error())
