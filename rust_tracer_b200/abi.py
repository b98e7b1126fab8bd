"""ctypes mirror of include/rt_b200.h (struct layouts and enum values only).

Kept free of any library loading so that both the product binding
(rust_tracer_b200/_lib.py) and the test-only oracle binding (tests/orc.py) can share it.
"""
import ctypes as C

RT_OK = 0
RT_ERR_INVALID, RT_ERR_UNSUPPORTED, RT_ERR_CUDA, RT_ERR_NOMEM, RT_ERR_INTERNAL, RT_ERR_TIMEOUT = -1, -2, -3, -4, -5, -6

RT_SHAPE_SPHERE, RT_SHAPE_PLANE, RT_SHAPE_TRIANGLE, RT_SHAPE_CUBE = 0, 1, 2, 3
RT_COLOR_CONST, RT_COLOR_CHECKERBOARD = 0, 1
RT_MATERIAL_PHONG, RT_MATERIAL_TEXTURE_PHONG = 0, 1
RT_LIGHT_POINT, RT_LIGHT_AMBIENT = 0, 1
RT_FLAG_NONE, RT_FLAG_AUDIT_ALL_PAIRS, RT_FLAG_FOREST_SHADING, RT_FLAG_FRAME_ROWS = 0, 1, 2, 4
RT_FRAME_HANDLE_BYTES = 64
RT_FRAME_SIGNAL_WORDS = 256
RT_FRAME_WORD_RELEASED = 255


class rt_shape(C.Structure):
    _fields_ = [("kind", C.c_int32), ("material", C.c_int32),
                ("transform", C.c_float * 16), ("inv_transform", C.c_float * 16),
                ("p", C.c_float * 12)]


class rt_colorfn(C.Structure):
    _fields_ = [("kind", C.c_int32), ("rgb", C.c_float * 3)]


class rt_material(C.Structure):
    _fields_ = [("kind", C.c_int32), ("ambient", rt_colorfn), ("diffuse", rt_colorfn),
                ("specular", rt_colorfn), ("power", C.c_float), ("reflectivity", C.c_float),
                ("refraction_index", C.c_float)]


class rt_light(C.Structure):
    _fields_ = [("kind", C.c_int32), ("pos", C.c_float * 3), ("color", C.c_float * 3)]


class rt_scene_desc(C.Structure):
    _fields_ = [("shapes", C.POINTER(rt_shape)), ("n_shapes", C.c_int32),
                ("materials", C.POINTER(rt_material)), ("n_materials", C.c_int32),
                ("lights", C.POINTER(rt_light)), ("n_lights", C.c_int32),
                ("ambient", C.c_float * 3)]


class rt_camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("x_min", C.c_float), ("x_max", C.c_float),
                ("y_min", C.c_float), ("y_max", C.c_float), ("x_res", C.c_int32), ("y_res", C.c_int32)]


class rt_render_params(C.Structure):
    _fields_ = [("depth", C.c_int32), ("spp", C.c_int32), ("seed", C.c_uint64),
                ("rank", C.c_int32), ("world", C.c_int32), ("tile_rows", C.c_int32), ("flags", C.c_int32)]


class rt_stats(C.Structure):
    _fields_ = [("pixels", C.c_uint64), ("rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("nodes", C.c_uint64), ("sphere_tests", C.c_uint64), ("plane_tests", C.c_uint64),
                ("triangle_tests", C.c_uint64), ("cube_tests", C.c_uint64), ("exact_tests", C.c_uint64),
                ("algorithmic_flops", C.c_double), ("sweep_ms", C.c_float), ("shade_ms", C.c_float),
                ("total_ms", C.c_float), ("kernel_launches", C.c_uint32), ("filter_flops", C.c_double), ("tested_pairs", C.c_uint64),
                ("bound_evals_sphere", C.c_uint64), ("bound_evals_triangle", C.c_uint64), ("bound_evals_cube", C.c_uint64),
                ("hier_evals", C.c_uint64), ("exact_sphere_tests", C.c_uint64), ("exact_triangle_tests", C.c_uint64),
                ("exact_cube_tests", C.c_uint64), ("exact_flops", C.c_double), ("executed_flops", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class rt_forest_stats(C.Structure):
    _fields_ = [("num_trees", C.c_uint64), ("num_intersections", C.c_uint64),
                ("smallest_tree", C.c_uint64), ("largest_tree", C.c_uint64), ("median", C.c_uint64),
                ("p90", C.c_uint64), ("p95", C.c_uint64), ("p99", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def default_camera(x_res, y_res):
    """Camera::new (reference src/render.rs:166-176)."""
    cam = rt_camera()
    cam.origin[:] = [0.0, 0.0, -8.0]
    cam.x_min, cam.x_max, cam.y_min, cam.y_max = -3.0, 3.0, -3.0, 3.0
    cam.x_res, cam.y_res = x_res, y_res
    return cam
