"""numpy-facing wrapper over the C ABI (include/rt_b200.h).

Mirrors the reference's entry points for the render path:
  Scene.render            <- render::render                (reference src/render.rs:31)
  Scene.build_forest      <- render_tree::generate_ray_forest (src/render_tree.rs:147)
  Forest.eval             <- render_tree::render_forest       (src/render_tree.rs:121)
  Forest.eval_filtered    <- render_tree::render_forest_filter (src/render_tree.rs:129)
  Forest.stats/size/trees_with <- RayForest::{stats,size,trees_with} (src/render_tree.rs:62-93)
All compute happens in librt_b200.so on the GPU; nothing here computes pixels.
"""
import ctypes as C

import numpy as np

from . import abi
from ._lib import check, lib


class SceneDesc:
    """A flat scene description (rt_scene_desc) and whatever owns its arrays."""

    def __init__(self, desc_ptr=None, owner=None, builtin=False):
        self._ptr = desc_ptr
        self._owner = owner
        self._builtin = builtin

    @classmethod
    def builtin(cls, name, seed=0):
        """"default" (src/my_scene.rs), "s1k", "m10k" (SURVEY 8d)."""
        p = C.POINTER(abi.rt_scene_desc)()
        check(lib().rt_builtin_scene(name.encode(), C.c_uint64(seed), C.byref(p)))
        return cls(p, builtin=True)

    @classmethod
    def synthetic(cls, n_spheres, n_triangles, seed, r_lo, r_hi):
        p = C.POINTER(abi.rt_scene_desc)()
        check(lib().rt_synthetic_scene(n_spheres, n_triangles, C.c_uint64(seed), r_lo, r_hi, C.byref(p)))
        return cls(p, builtin=True)

    @classmethod
    def load(cls, path):
        """Read a scene description file (rt_scene_desc_load, include/rt_b200_host.h)."""
        p = C.POINTER(abi.rt_scene_desc)()
        check(lib().rt_scene_desc_load(str(path).encode(), C.byref(p)))
        return cls(p, builtin=True)

    def save(self, path):
        """Write this description as a scene file (rt_scene_desc_save)."""
        check(lib().rt_scene_desc_save(self._ptr, str(path).encode()))

    @classmethod
    def from_arrays(cls, shapes, materials, lights, ambient):
        """shapes/materials/lights: ctypes arrays of rt_shape / rt_material / rt_light."""
        d = abi.rt_scene_desc()
        d.shapes = C.cast(shapes, C.POINTER(abi.rt_shape))
        d.n_shapes = len(shapes)
        d.materials = C.cast(materials, C.POINTER(abi.rt_material))
        d.n_materials = len(materials)
        d.lights = C.cast(lights, C.POINTER(abi.rt_light))
        d.n_lights = len(lights)
        d.ambient[:] = list(ambient)
        return cls(C.pointer(d), owner=(shapes, materials, lights, d))

    @property
    def desc(self):
        return self._ptr.contents

    def materials_copy(self):
        d = self.desc
        arr = (abi.rt_material * d.n_materials)()
        C.memmove(arr, d.materials, C.sizeof(abi.rt_material) * d.n_materials)
        return arr

    def shape_kinds(self):
        d = self.desc
        return np.array([d.shapes[i].kind for i in range(d.n_shapes)], dtype=np.int32)

    def close(self):
        if self._builtin and self._ptr:
            lib().rt_scene_desc_free(self._ptr)
        self._ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def camera(x_res, y_res):
    """Camera::new (src/render.rs:166-176)."""
    cam = abi.rt_camera()
    lib().rt_camera_default(x_res, y_res, C.byref(cam))
    return cam


def tile_rows(height, rank=0, world=1, rows_per_tile=8):
    n = check(lib().rt_tile_rows_for_rank(height, rank, world, rows_per_tile, None, 0))
    rows = np.zeros(n, dtype=np.int32)
    check(lib().rt_tile_rows_for_rank(height, rank, world, rows_per_tile,
                                      rows.ctypes.data_as(C.POINTER(C.c_int32)), n))
    return rows


def _params(depth, spp=1, seed=0, rank=0, world=1, rows_per_tile=8, flags=0):
    p = abi.rt_render_params()
    p.depth, p.spp, p.seed = depth, spp, seed
    p.rank, p.world, p.tile_rows, p.flags = rank, world, rows_per_tile, flags
    return p


class Scene:
    """Device-resident scene tables (rt_scene)."""

    def __init__(self, desc, device=0):
        self._h = C.c_void_p()
        self.desc = desc
        check(lib().rt_scene_create(C.byref(desc.desc), device, C.byref(self._h)))
        self.device = device

    def set_stream(self, cuda_stream_handle):
        """Run on a caller-created CUDA stream (int handle); 0/None restores the private one."""
        check(lib().rt_scene_set_stream(self._h, C.c_void_p(cuda_stream_handle or None)))

    def update_materials(self, materials):
        check(lib().rt_scene_update_materials(self._h, materials, len(materials)))

    def render(self, width, height, depth, spp=1, seed=0, rank=0, world=1, rows_per_tile=8, flags=0,
               want_f32=True, want_u8=True, out_u8=None, cam=None):
        """Returns (rgb_f32 | None, rgb_u8 | None, stats dict); arrays are [rows, width, 3].
        out_u8: caller-owned uint8 array of that shape to fill (e.g. a view of pinned memory).
        cam: an abi.rt_camera to use instead of Camera::new(width, height)."""
        cam = cam if cam is not None else camera(width, height)
        p = _params(depth, spp, seed, rank, world, rows_per_tile, flags)
        n_rows = len(tile_rows(height, rank, world, rows_per_tile))
        f = np.zeros((n_rows, width, 3), dtype=np.float32) if want_f32 else None
        u = np.zeros((n_rows, width, 3), dtype=np.uint8) if want_u8 and out_u8 is None else None
        if out_u8 is not None:
            if out_u8.dtype != np.uint8 or out_u8.size != n_rows * width * 3 or not out_u8.flags.c_contiguous:
                raise ValueError("out_u8 must be a contiguous uint8 array of %d x %d x 3" % (n_rows, width))
            u, want_u8 = out_u8, True
        st = abi.rt_stats()
        check(lib().rt_render(self._h, C.byref(cam), C.byref(p),
                              f.ctypes.data if want_f32 else None, u.ctypes.data if want_u8 else None, C.byref(st)))
        return f, u, st.as_dict()

    def render_device(self, width, height, depth, d_rgb_f32=0, d_rgb_u8=0, spp=1, seed=0, rank=0, world=1,
                      rows_per_tile=8, flags=0):
        """Outputs are raw device pointers (ints), e.g. torch tensor .data_ptr()."""
        cam = camera(width, height)
        p = _params(depth, spp, seed, rank, world, rows_per_tile, flags)
        st = abi.rt_stats()
        check(lib().rt_render_device(self._h, C.byref(cam), C.byref(p), C.c_void_p(d_rgb_f32 or None),
                                     C.c_void_p(d_rgb_u8 or None), C.byref(st)))
        return st.as_dict()

    def build_forest(self, width, height, depth, flags=0):
        cam = camera(width, height)
        p = _params(depth, flags=flags)
        h = C.c_void_p()
        st = abi.rt_stats()
        check(lib().rt_forest_build(self._h, C.byref(cam), C.byref(p), C.byref(h), C.byref(st)))
        return Forest(h, self, width, height, st.as_dict())

    def close(self):
        if self._h:
            lib().rt_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MultiScene:
    """One scene resident on several GPUs, rendered by one call (rt_multi: the reference's single
    blocking render::render, src/render.rs:31, for a host with 1-8 GPUs)."""

    def __init__(self, desc, devices):
        self._h = C.c_void_p()
        self.desc = desc
        self.devices = list(devices)
        arr = (C.c_int32 * len(self.devices))(*self.devices)
        check(lib().rt_multi_create(C.byref(desc.desc), arr, len(self.devices), C.byref(self._h)))

    def update_materials(self, materials):
        check(lib().rt_multi_update_materials(self._h, materials, len(materials)))

    def render(self, width, height, depth, spp=1, seed=0, rows_per_tile=8, flags=0, want_f32=True, want_u8=True,
               out_u8=None, cam=None):
        """Whole frame into host arrays; returns (rgb_f32 | None, rgb_u8 | None, stats dict)."""
        cam = cam if cam is not None else camera(width, height)
        p = _params(depth, spp, seed, 0, 1, rows_per_tile, flags)
        f = np.zeros((height, width, 3), dtype=np.float32) if want_f32 else None
        u = np.zeros((height, width, 3), dtype=np.uint8) if want_u8 and out_u8 is None else None
        if out_u8 is not None:
            if out_u8.dtype != np.uint8 or out_u8.size != height * width * 3 or not out_u8.flags.c_contiguous:
                raise ValueError("out_u8 must be a contiguous uint8 array of %d x %d x 3" % (height, width))
            u, want_u8 = out_u8, True
        st = abi.rt_stats()
        check(lib().rt_multi_render(self._h, C.byref(cam), C.byref(p), f.ctypes.data if want_f32 else None,
                                    u.ctypes.data if want_u8 else None, C.byref(st)))
        return f, u, st.as_dict()

    def render_device(self, width, height, depth, d_rgb_f32=0, d_rgb_u8=0, spp=1, seed=0, rows_per_tile=8, flags=0):
        """Whole frame into device buffers on the first device (raw pointers), gathered by peer stores."""
        cam = camera(width, height)
        p = _params(depth, spp, seed, 0, 1, rows_per_tile, flags)
        st = abi.rt_stats()
        check(lib().rt_multi_render_device(self._h, C.byref(cam), C.byref(p), C.c_void_p(d_rgb_f32 or None),
                                           C.c_void_p(d_rgb_u8 or None), C.byref(st)))
        return st.as_dict()

    def close(self):
        if self._h:
            lib().rt_multi_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def render_multi(desc, devices, width, height, depth, spp=1, seed=0, rows_per_tile=8, out_u8=None):
    """rt_render_multi: scene description in, RGB8 image out, on the listed devices."""
    cam = camera(width, height)
    p = _params(depth, spp, seed, 0, 1, rows_per_tile, 0)
    u = out_u8 if out_u8 is not None else np.zeros((height, width, 3), dtype=np.uint8)
    arr = (C.c_int32 * len(devices))(*devices)
    st = abi.rt_stats()
    check(lib().rt_render_multi(C.byref(desc.desc), arr, len(devices), C.byref(cam), C.byref(p), None, u.ctypes.data, C.byref(st)))
    return u, st.as_dict()


def device_count():
    return check(lib().rt_device_count())


class Forest:
    """Device-resident ray forest (rt_forest)."""

    def __init__(self, handle, scene, width, height, build_stats):
        self._h, self.scene, self.width, self.height = handle, scene, width, height
        self.build_stats = build_stats

    @staticmethod
    def _amb(ambient):
        if ambient is None:
            return None
        return (C.c_float * 3)(*ambient)

    def eval(self, ambient=None):
        f = np.zeros((self.height, self.width, 3), dtype=np.float32)
        u = np.zeros((self.height, self.width, 3), dtype=np.uint8)
        st = abi.rt_stats()
        check(lib().rt_forest_eval(self._h, self._amb(ambient), f.ctypes.data, u.ctypes.data, C.byref(st)))
        return f, u, st.as_dict()

    def eval_filtered(self, mutated_ids, rgb_f32, rgb_u8, ambient=None):
        """Re-shades in place the pixels whose tree touches one of mutated_ids; returns their count."""
        ids = np.ascontiguousarray(mutated_ids, dtype=np.int32)
        n = C.c_uint64()
        st = abi.rt_stats()
        check(lib().rt_forest_eval_filtered(self._h, self._amb(ambient), ids.ctypes.data_as(C.POINTER(C.c_int32)),
                                            len(ids), rgb_f32.ctypes.data if rgb_f32 is not None else None,
                                            rgb_u8.ctypes.data if rgb_u8 is not None else None, C.byref(n),
                                            C.byref(st)))
        return n.value, st.as_dict()

    def stats(self):
        s = abi.rt_forest_stats()
        check(lib().rt_forest_get_stats(self._h, C.byref(s)))
        return s.as_dict()

    def size(self):
        n = C.c_uint64()
        check(lib().rt_forest_size(self._h, C.byref(n)))
        return n.value

    def trees_with(self, shape_id):
        n = C.c_uint64()
        check(lib().rt_forest_trees_with(self._h, shape_id, C.byref(n)))
        return n.value

    def trees_with_static(self, shape_id):
        """trees_with through the entry point that allocates nothing (static device scratch, at most
        RT_STATIC_TREES_MAX trees; not in the reference)."""
        n = C.c_uint64()
        check(lib().rt_forest_trees_with_static(self._h, shape_id, C.byref(n)))
        return n.value

    def close(self):
        if self._h:
            lib().rt_forest_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class tuned:
    """Test / tool hook (include/rt_b200_debug.h): override scheduling constants of the library for the
    duration of a `with` block, e.g. `with rt.tuned(table_cap=128): ...`.  Never changes a result."""

    def __init__(self, **knobs):
        self.knobs = knobs
        self.saved = {}

    def __enter__(self):
        for k, v in self.knobs.items():
            old = lib().rt_debug_get(k.encode())
            if old == -(1 << 63):
                raise KeyError(k)
            self.saved[k] = old
            check(lib().rt_debug_set(k.encode(), int(v)))
        return self

    def __exit__(self, *exc):
        for k, v in self.saved.items():
            lib().rt_debug_set(k.encode(), int(v))
        return False


def trim(device=-1):
    """Free the ray pools cached from destroyed scenes (rt_trim); -1 = every device."""
    check(lib().rt_trim(device))


def measure_fp32_peak(device=0):
    """(measured TFLOP/s, SM count, SM clock kHz) from the FFMA-saturation micro-kernel."""
    t, sms, khz = C.c_double(), C.c_int32(), C.c_int32()
    check(lib().rt_measure_fp32_peak(device, C.byref(t), C.byref(sms), C.byref(khz)))
    return t.value, sms.value, khz.value
