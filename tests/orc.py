"""ctypes binding of the CPU oracle (oracle/liborc.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module; the product package never does.
"""
import ctypes as C
import lzma
import os
import struct
import subprocess
import sys

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(_ROOT, "oracle", "liborc.so")


def _abi_mirror():
    """The ctypes mirror of include/rt_b200.h (struct layouts only, no library loading).  Inside the
    tests it is the product package's module; bench.py --impl reference must not import the product
    package at all, so there the same file is loaded on its own."""
    mod = sys.modules.get("rust_tracer_b200.abi")
    if mod is None and "rust_tracer_b200" not in sys.modules and os.environ.get("RT_ORC_STANDALONE"):
        import importlib.util
        spec = importlib.util.spec_from_file_location("rt_abi_mirror", os.path.join(_ROOT, "rust_tracer_b200", "abi.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    if mod is None:
        from rust_tracer_b200 import abi as mod
    return mod


abi = _abi_mirror()


def build(force=False):
    if force or not os.path.exists(_SO):
        subprocess.check_call(["make", "-C", os.path.join(_ROOT, "oracle"), "liborc.so"])
    return _SO


class orc_counters(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in
                ("rays", "shadow_rays", "nodes", "sphere_tests", "plane_tests", "triangle_tests", "cube_tests")]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.orc_forest_build.restype = C.c_void_p
        _lib.orc_forest_eval.restype = C.c_longlong
        _lib.orc_forest_trees_with.restype = C.c_uint64
        _lib.orc_forest_tree_size.restype = C.c_uint64
        _lib.orc_fresnel_reflection.restype = C.c_float
        _lib.orc_as_u8.restype = C.c_uint8
        _lib.orc_as_u8.argtypes = [C.c_float]
    return _lib


def fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def f32(*v):
    return np.asarray(v, dtype=np.float32).reshape(-1).copy()


class SceneArrays:
    """Owns the ctypes arrays behind an rt_scene_desc."""

    def __init__(self, shapes, materials, lights, ambient):
        self.shapes, self.materials, self.lights = shapes, materials, lights
        self.desc = abi.rt_scene_desc()
        self.desc.shapes = C.cast(shapes, C.POINTER(abi.rt_shape))
        self.desc.n_shapes = len(shapes)
        self.desc.materials = C.cast(materials, C.POINTER(abi.rt_material))
        self.desc.n_materials = len(materials)
        self.desc.lights = C.cast(lights, C.POINTER(abi.rt_light))
        self.desc.n_lights = len(lights)
        self.desc.ambient[:] = list(ambient)


def default_scene():
    """src/my_scene.rs restated by the oracle (independent of the product's builder)."""
    shapes = (abi.rt_shape * 6)()
    mats = (abi.rt_material * 6)()
    lights = (abi.rt_light * 3)()
    amb = (C.c_float * 3)()
    assert lib().orc_default_scene(shapes, mats, lights, amb) == 0
    return SceneArrays(shapes, mats, lights, list(amb))


_RTSCENE_HEADER = struct.Struct("<8sII3I3i3fIQ")  # rust_tracer_b200/csrc/host/scene_file.cpp: 64 bytes


def load_rtscene(path):
    """Independent reader of a scene description file (.rtscene, optionally .xz): header with record
    sizes and counts, then the rt_shape / rt_material / rt_light arrays as they are.  Lets the oracle
    render the benchmark scenes from committed fixtures without the product library."""
    opener = lzma.open if str(path).endswith(".xz") else open
    with opener(path, "rb") as fh:
        raw = fh.read()
    magic, version, _abi_v, sb, mb, lb, ns, nm, nl, ar, ag, ab, _zero, _checksum = _RTSCENE_HEADER.unpack_from(raw)
    if magic != b"RTSCENE\0" or version != 1 or (sb, mb, lb) != (C.sizeof(abi.rt_shape), C.sizeof(abi.rt_material), C.sizeof(abi.rt_light)):
        raise ValueError("%s is not a scene description file of this layout" % path)
    off = _RTSCENE_HEADER.size
    if len(raw) != off + sb * ns + mb * nm + lb * nl:
        raise ValueError("%s: truncated" % path)
    shapes = (abi.rt_shape * ns).from_buffer_copy(raw, off)
    off += sb * ns
    mats = (abi.rt_material * nm).from_buffer_copy(raw, off)
    off += mb * nm
    lights = (abi.rt_light * nl).from_buffer_copy(raw, off)
    return SceneArrays(shapes, mats, lights, (ar, ag, ab))


def render(desc, cam, depth, rows=None, mode=0, threads=None, spp=1, seed=0, want_f32=True):
    """Oracle render of the listed rows (all rows if None).  Returns (rgb_f32|None, rgb_u8, counters)."""
    if rows is None:
        rows = np.arange(cam.y_res, dtype=np.int32)
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    if threads is None:
        threads = os.cpu_count() or 1
    f = np.zeros((len(rows), cam.x_res, 3), dtype=np.float32) if want_f32 else None
    u = np.zeros((len(rows), cam.x_res, 3), dtype=np.uint8)
    cnt = orc_counters()
    rc = lib().orc_render(C.byref(desc), C.byref(cam), C.c_int(depth), C.c_int(spp), C.c_uint64(seed),
                          rows.ctypes.data_as(C.POINTER(C.c_int32)), C.c_int(len(rows)), C.c_int(mode),
                          C.c_int(threads), fptr(f) if want_f32 else None,
                          u.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(cnt))
    if rc != 0:
        raise RuntimeError("orc_render failed")
    return f, u, cnt


def intersect_batch(desc, o, d):
    o = np.ascontiguousarray(o, dtype=np.float32)
    d = np.ascontiguousarray(d, dtype=np.float32)
    n = o.shape[0]
    hit = np.zeros(n, dtype=np.int32)
    t = np.zeros(n, dtype=np.float32)
    rc = lib().orc_intersect_batch(C.byref(desc), fptr(o), fptr(d), C.c_int(n),
                                   hit.ctypes.data_as(C.POINTER(C.c_int32)), fptr(t))
    assert rc == 0
    return hit, t
