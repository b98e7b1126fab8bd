"""Loader and prototypes of librt_b200.so (include/rt_b200.h, include/rt_b200_host.h).

The library is the product: if it is missing this module raises, it never falls back to
anything else (there is no CPU path in this package).
"""
import ctypes as C
import os

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librt_b200.so")

_lib = None


class RtError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"librt_b200 error {code}: {message}")
        self.code = code


def _declare(lib):
    P = C.POINTER
    vp = C.c_void_p
    lib.rt_abi_version.restype = C.c_int
    lib.rt_last_error.restype = C.c_char_p
    lib.rt_device_count.restype = C.c_int
    lib.rt_camera_default.argtypes = [C.c_int32, C.c_int32, P(abi.rt_camera)]
    lib.rt_camera_default.restype = None
    lib.rt_tile_rows_for_rank.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32, P(C.c_int32), C.c_int32]
    lib.rt_tile_rows_for_rank.restype = C.c_int32
    lib.rt_scene_create.argtypes = [P(abi.rt_scene_desc), C.c_int32, P(vp)]
    lib.rt_scene_update_materials.argtypes = [vp, P(abi.rt_material), C.c_int32]
    lib.rt_scene_set_stream.argtypes = [vp, vp]
    lib.rt_scene_destroy.argtypes = [vp]
    lib.rt_scene_destroy.restype = None
    lib.rt_trim.argtypes = [C.c_int32]
    lib.rt_frame_alloc.argtypes = [C.c_int32, C.c_size_t, P(vp)]
    lib.rt_frame_export.argtypes = [C.c_int32, vp, P(C.c_uint8)]
    lib.rt_frame_import.argtypes = [C.c_int32, P(C.c_uint8), P(vp)]
    lib.rt_frame_release.argtypes = [C.c_int32, vp, C.c_int32]
    lib.rt_frame_signal.argtypes = [C.c_int32, vp, C.c_size_t, C.c_int32, C.c_uint32, vp]
    lib.rt_frame_wait.argtypes = [C.c_int32, vp, C.c_size_t, C.c_int32, C.c_int32, C.c_uint32, C.c_uint32, vp]
    lib.rt_render.argtypes = [vp, P(abi.rt_camera), P(abi.rt_render_params), vp, vp, P(abi.rt_stats)]
    lib.rt_render_device.argtypes = [vp, P(abi.rt_camera), P(abi.rt_render_params), vp, vp, P(abi.rt_stats)]
    lib.rt_multi_create.argtypes = [P(abi.rt_scene_desc), P(C.c_int32), C.c_int32, P(vp)]
    lib.rt_multi_device_count.argtypes = [vp]
    lib.rt_multi_device_count.restype = C.c_int32
    lib.rt_multi_update_materials.argtypes = [vp, P(abi.rt_material), C.c_int32]
    lib.rt_multi_render.argtypes = [vp, P(abi.rt_camera), P(abi.rt_render_params), vp, vp, P(abi.rt_stats)]
    lib.rt_multi_render_device.argtypes = [vp, P(abi.rt_camera), P(abi.rt_render_params), vp, vp, P(abi.rt_stats)]
    lib.rt_multi_destroy.argtypes = [vp]
    lib.rt_multi_destroy.restype = None
    lib.rt_render_multi.argtypes = [P(abi.rt_scene_desc), P(C.c_int32), C.c_int32, P(abi.rt_camera), P(abi.rt_render_params),
                                    vp, vp, P(abi.rt_stats)]
    lib.rt_forest_build.argtypes = [vp, P(abi.rt_camera), P(abi.rt_render_params), P(vp), P(abi.rt_stats)]
    lib.rt_forest_eval.argtypes = [vp, P(C.c_float), vp, vp, P(abi.rt_stats)]
    lib.rt_forest_eval_filtered.argtypes = [vp, P(C.c_float), P(C.c_int32), C.c_int32, vp, vp, P(C.c_uint64),
                                           P(abi.rt_stats)]
    lib.rt_forest_get_stats.argtypes = [vp, P(abi.rt_forest_stats)]
    lib.rt_forest_size.argtypes = [vp, P(C.c_uint64)]
    lib.rt_forest_trees_with.argtypes = [vp, C.c_int32, P(C.c_uint64)]
    lib.rt_forest_trees_with_static.argtypes = [vp, C.c_int32, P(C.c_uint64)]
    lib.rt_forest_destroy.argtypes = [vp]
    lib.rt_forest_destroy.restype = None
    lib.rt_measure_fp32_peak.argtypes = [C.c_int32, P(C.c_double), P(C.c_int32), P(C.c_int32)]
    # debug hooks (include/rt_b200_debug.h)
    lib.rt_debug_set.argtypes = [C.c_char_p, C.c_int64]
    lib.rt_debug_get.argtypes = [C.c_char_p]
    lib.rt_debug_get.restype = C.c_int64
    lib.rt_debug_reset.restype = None
    lib.rt_debug_plan_items.argtypes = [C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_int32, P(C.c_int64)]
    lib.rt_debug_scene_spheres.argtypes = [P(abi.rt_scene_desc), C.c_int32, P(C.c_double), C.c_int32]
    lib.rt_debug_scene_scalars.argtypes = [P(abi.rt_scene_desc), P(C.c_double)]
    lib.rt_debug_scene_light_table.argtypes = [P(abi.rt_scene_desc), C.c_int32, P(C.c_double), C.c_int32, P(C.c_double)]
    lib.rt_debug_decode_item.argtypes = [C.c_uint32, C.c_int64, C.c_int32, P(C.c_int64), P(C.c_int64)]
    # host helpers
    lib.rt_builtin_scene.argtypes = [C.c_char_p, C.c_uint64, P(P(abi.rt_scene_desc))]
    lib.rt_synthetic_scene.argtypes = [C.c_int32, C.c_int32, C.c_uint64, C.c_float, C.c_float,
                                      P(P(abi.rt_scene_desc))]
    lib.rt_scene_desc_free.argtypes = [P(abi.rt_scene_desc)]
    lib.rt_scene_desc_free.restype = None
    lib.rt_scene_desc_save.argtypes = [P(abi.rt_scene_desc), C.c_char_p]
    lib.rt_scene_desc_load.argtypes = [C.c_char_p, P(P(abi.rt_scene_desc))]
    lib.rt_matrix_inverse.argtypes = [P(C.c_float), P(C.c_float)]
    lib.rt_matrix_compose.argtypes = [C.c_float] * 9 + [P(C.c_float)]
    lib.rt_matrix_compose.restype = None
    lib.rt_plane_axes.argtypes = [P(C.c_float)] * 3
    lib.rt_plane_axes.restype = None
    lib.rt_triangle_normal.argtypes = [P(C.c_float)] * 4
    lib.rt_triangle_normal.restype = None


# every symbol include/rt_b200.h and include/rt_b200_host.h declare
EXPORTS = [
    "rt_abi_version", "rt_last_error", "rt_device_count", "rt_camera_default", "rt_tile_rows_for_rank",
    "rt_scene_create", "rt_scene_update_materials", "rt_scene_set_stream", "rt_scene_destroy", "rt_trim", "rt_render", "rt_render_device",
    "rt_multi_create", "rt_multi_device_count", "rt_multi_update_materials", "rt_multi_render", "rt_multi_render_device",
    "rt_multi_destroy", "rt_render_multi",
    "rt_frame_alloc", "rt_frame_export", "rt_frame_import", "rt_frame_release", "rt_frame_signal", "rt_frame_wait",
    "rt_forest_build", "rt_forest_eval", "rt_forest_eval_filtered", "rt_forest_get_stats", "rt_forest_size",
    "rt_forest_trees_with", "rt_forest_trees_with_static", "rt_forest_destroy", "rt_measure_fp32_peak",
    "rt_builtin_scene", "rt_synthetic_scene", "rt_scene_desc_free", "rt_scene_desc_save", "rt_scene_desc_load",
    "rt_matrix_inverse", "rt_matrix_compose",
    "rt_plane_axes", "rt_triangle_normal",
]


# include/rt_b200_debug.h: test and measurement hooks, not part of the drop-in ABI
DEBUG_EXPORTS = ["rt_debug_set", "rt_debug_get", "rt_debug_reset", "rt_debug_plan_items", "rt_debug_decode_item",
                 "rt_debug_scene_spheres", "rt_debug_scene_scalars", "rt_debug_scene_light_table"]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C rust_tracer_b200/csrc). There is no fallback path.")
        handle = C.CDLL(LIB_PATH)
        _declare(handle)
        if handle.rt_abi_version() != 4:
            raise ImportError("librt_b200.so ABI version mismatch")
        _lib = handle
    return _lib


def check(rc):
    if rc < 0:
        raise RtError(rc, lib().rt_last_error().decode("utf-8", "replace"))
    return rc
