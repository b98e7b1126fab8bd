"""CPU: the C-ABI library loads, exports every symbol the headers declare, and its host
side (scene builders, matrix helpers) agrees with the oracle.  No compute call needs a GPU
here; the one that is made must fail loudly when there is none."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import orc
import rust_tracer_b200 as rt
from rust_tracer_b200 import _lib, abi
from conftest import HAS_GPU, ROOT


def _declared_functions(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text))


def test_library_exports_every_declared_symbol():
    declared = _declared_functions("rt_b200.h") | _declared_functions("rt_b200_host.h")
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    handle = rt.lib()
    for name in declared:
        assert hasattr(handle, name), name
    assert handle.rt_abi_version() == 4
    # the debug hooks have their own header and are not part of the drop-in ABI
    debug = _declared_functions("rt_b200_debug.h")
    assert debug == set(_lib.DEBUG_EXPORTS) and not (debug & declared)
    for name in debug:
        assert hasattr(handle, name), name


def test_struct_sizes_match_the_header():
    # computed from the C declarations: 2 ints + 44 floats; 1 + 3*4 + 3 words; ...
    assert C.sizeof(abi.rt_shape) == 184
    assert C.sizeof(abi.rt_colorfn) == 16
    assert C.sizeof(abi.rt_material) == 64
    assert C.sizeof(abi.rt_light) == 28
    assert C.sizeof(abi.rt_camera) == 36
    assert C.sizeof(abi.rt_render_params) == 32
    assert C.sizeof(abi.rt_stats) == 184


def _bytes(ptr, n, T):
    return C.string_at(ptr, n * C.sizeof(T))


def test_default_scene_is_bit_identical_to_the_oracles():
    da, ob = rt.SceneDesc.builtin("default"), orc.default_scene()  # keep the owners alive
    a, b = da.desc, ob.desc
    assert (a.n_shapes, a.n_materials, a.n_lights) == (6, 6, 3) == (b.n_shapes, b.n_materials, b.n_lights)
    assert _bytes(a.shapes, 6, abi.rt_shape) == _bytes(b.shapes, 6, abi.rt_shape)
    assert _bytes(a.materials, 6, abi.rt_material) == _bytes(b.materials, 6, abi.rt_material)
    assert _bytes(a.lights, 3, abi.rt_light) == _bytes(b.lights, 3, abi.rt_light)
    assert list(a.ambient) == list(b.ambient)


def test_synthetic_scenes_shape():
    for name, n_s, n_t in (("s1k", 1000, 0), ("m10k", 5000, 5000)):
        d = rt.SceneDesc.builtin(name)
        kinds = d.shape_kinds()
        assert kinds[0] == abi.RT_SHAPE_PLANE
        assert (kinds == abi.RT_SHAPE_SPHERE).sum() == n_s and (kinds == abi.RT_SHAPE_TRIANGLE).sum() == n_t
        assert d.desc.n_lights == 4 and d.desc.n_materials == 1 + n_s + n_t
        # deterministic
        e = rt.SceneDesc.builtin(name)
        assert _bytes(d.desc.shapes, d.desc.n_shapes, abi.rt_shape) == _bytes(e.desc.shapes, e.desc.n_shapes, abi.rt_shape)
        # everything stays above the floor (y = -2), SURVEY 8d
        for i in range(1, d.desc.n_shapes, 97):
            s = d.desc.shapes[i]
            if s.kind == abi.RT_SHAPE_SPHERE:
                assert s.transform[7] - 0.21 > -2.0
            else:
                assert min(s.p[1], s.p[4], s.p[7]) > -2.0


def test_matrix_inverse_matches_oracle_bitwise():
    rng = np.random.default_rng(3)
    L = rt.lib()
    for _ in range(200):
        v = rng.uniform(-3, 3, 9).astype(np.float32)
        m = np.zeros(16, np.float32)
        L.rt_matrix_compose(*[C.c_float(float(x)) for x in v[:6]], *[C.c_float(float(abs(x) + 0.1)) for x in v[6:]],
                            m.ctypes.data_as(C.POINTER(C.c_float)))
        a, b = np.zeros(16, np.float32), np.zeros(16, np.float32)
        assert L.rt_matrix_inverse(m.ctypes.data_as(C.POINTER(C.c_float)), a.ctypes.data_as(C.POINTER(C.c_float))) == 0
        assert orc.lib().orc_mat_inverse(orc.fptr(m), orc.fptr(b)) == 0
        assert a.tobytes() == b.tobytes()
    z = np.zeros(16, np.float32)
    assert L.rt_matrix_inverse(z.ctypes.data_as(C.POINTER(C.c_float)), a.ctypes.data_as(C.POINTER(C.c_float))) == abi.RT_ERR_INVALID


def test_plane_axes_and_triangle_normal_match_oracle():
    rng = np.random.default_rng(5)
    L, O = rt.lib(), orc.lib()
    P = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    for _ in range(100):
        n = rng.normal(size=3).astype(np.float32)
        u1, v1, u2, v2 = (np.zeros(3, np.float32) for _ in range(4))
        L.rt_plane_axes(P(n), P(u1), P(v1))
        O.orc_plane_axes(P(n), P(u2), P(v2))
        assert u1.tobytes() == u2.tobytes() and v1.tobytes() == v2.tobytes()
        tri = rng.uniform(-2, 2, (3, 3)).astype(np.float32)
        n1, n2 = np.zeros(3, np.float32), np.zeros(3, np.float32)
        L.rt_triangle_normal(P(tri[0].copy()), P(tri[1].copy()), P(tri[2].copy()), P(n1))
        O.orc_triangle_normal(P(tri[0].copy()), P(tri[1].copy()), P(tri[2].copy()), P(n2))
        assert n1.tobytes() == n2.tobytes()


def test_tile_rows_partition():
    for h, world, tr in ((1080, 1, 8), (1080, 2, 8), (2160, 8, 16), (37, 4, 5), (5, 8, 1)):
        seen = []
        for rank in range(world):
            rows = rt.tile_rows(h, rank, world, tr)
            assert all((v // tr) % world == rank for v in rows)
            assert list(rows) == sorted(rows)
            seen.extend(rows.tolist())
        assert sorted(seen) == list(range(h))  # every row exactly once
    with pytest.raises(rt.RtError):
        rt.tile_rows(10, 2, 2, 8)


def test_scene_validation_errors_do_not_need_a_gpu():
    d = rt.SceneDesc.builtin("default")
    mats = d.materials_copy()
    mats[3].diffuse.kind = 77  # a texture function the device does not have
    bad = rt.SceneDesc.from_arrays((abi.rt_shape * 0)(), mats, (abi.rt_light * 0)(), (0, 0, 0))
    h = C.c_void_p()
    rc = rt.lib().rt_scene_create(None, 0, C.byref(h))
    assert rc == abi.RT_ERR_INVALID and b"null" in rt.lib().rt_last_error()
    del bad


@pytest.mark.skipif(HAS_GPU, reason="only meaningful without a GPU")
def test_compute_fails_loudly_without_a_gpu():
    with pytest.raises(rt.RtError) as e:
        rt.Scene(rt.SceneDesc.builtin("default"))
    assert e.value.code == abi.RT_ERR_CUDA
    with pytest.raises(rt.RtError):
        rt.measure_fp32_peak(0)


def test_library_does_not_read_the_environment():
    """A drop-in must not change its behaviour because of what the caller's process exports: the
    library has no getenv at all; its scheduling constants are only reachable through
    rt_debug_set (include/rt_b200_debug.h)."""
    import subprocess
    out = subprocess.run(["nm", "-D", "--undefined-only", _lib.LIB_PATH], capture_output=True, text=True)
    assert out.returncode == 0
    assert "getenv" not in out.stdout
    L = rt.lib()
    assert L.rt_debug_get(b"item_max") == 16 and L.rt_debug_get(b"no_such_knob") == -(1 << 63)
    assert L.rt_debug_set(b"tri_bound_scale", 1) == abi.RT_ERR_INVALID  # correctness margins are not knobs
    assert L.rt_debug_set(b"item_max", 99) == abi.RT_ERR_INVALID
    with rt.tuned(item_max=3, table_cap=512):
        assert L.rt_debug_get(b"item_max") == 3 and L.rt_debug_get(b"table_cap") == 512
    assert L.rt_debug_get(b"item_max") == 16 and L.rt_debug_get(b"table_cap") == 0
    L.rt_debug_set(b"want_light", 9)
    L.rt_debug_reset()
    assert L.rt_debug_get(b"want_light") == 4


def test_debug_knobs_named_in_the_header_are_the_fields_of_tuning_h():
    """include/rt_b200_debug.h lists the knobs by name; each is a field of rt::Tuning (csrc/tuning.h) with the
    default the library reports, and the library knows no field the header does not name."""
    import os, re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "rt_b200_debug.h")).read()
    listed = re.findall(r'"([a-z_0-9]+)"', header[header.index("name: a field of rt::Tuning"):header.index("Process-wide")])
    tuning = open(os.path.join(root, "rust_tracer_b200", "csrc", "tuning.h")).read()
    fields = dict(re.findall(r"^\s+int (\w+) = ([^;]+);", tuning, flags=re.M))
    assert sorted(listed) == sorted(fields), (sorted(set(listed) ^ set(fields)))
    L = rt.lib()
    L.rt_debug_reset()
    for name, default in fields.items():
        assert L.rt_debug_get(name.encode()) == eval(default), name  # ("1 << 23" and the like)


def test_package_does_not_import_the_oracle():
    import subprocess, sys
    code = ("import sys; sys.path.insert(0, %r); import rust_tracer_b200 as rt; rt.lib(); "
            "import ctypes; bad=[m for m in sys.modules if m=='orc']; "
            "maps=open('/proc/self/maps').read(); assert 'liborc' not in maps and not bad; print('clean')" % ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert out.returncode == 0 and "clean" in out.stdout, out.stderr
