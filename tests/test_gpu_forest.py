"""GPU: the RayForest mode (reference src/render_tree.rs) through the C ABI vs the oracle."""
import ctypes as C

import numpy as np
import pytest

import orc
import rust_tracer_b200 as rt
from rust_tracer_b200 import abi

pytestmark = pytest.mark.gpu


def _orc_forest(desc, w, h, depth):
    cam = abi.default_camera(w, h)
    return orc.lib().orc_forest_build(C.byref(desc.desc), C.byref(cam), depth, 8), cam


def _orc_eval(handle, desc, w, h, ids=None, into=None):
    f = np.zeros((h, w, 3), np.float32) if into is None else into[0]
    u = np.zeros((h, w, 3), np.uint8) if into is None else into[1]
    arr = None if ids is None else np.ascontiguousarray(ids, np.int32)
    n = orc.lib().orc_forest_eval(C.c_void_p(handle), C.byref(desc.desc), None,
                                  None if arr is None else arr.ctypes.data_as(C.POINTER(C.c_int32)),
                                  0 if arr is None else len(arr), orc.fptr(f), u.ctypes.data_as(C.POINTER(C.c_uint8)))
    return f, u, n


def test_forest_render_matches_oracle_and_differs_from_basic():
    desc = rt.SceneDesc.builtin("default")
    w, h, depth = 200, 160, 8
    sc = rt.Scene(desc)
    forest = sc.build_forest(w, h, depth)
    f, u, _ = forest.eval()
    oh, _ = _orc_forest(desc, w, h, depth)
    of, ou, n = _orc_eval(oh, desc, w, h)
    assert n == w * h
    d = np.abs(u.astype(int) - ou.astype(int)).max(-1)
    assert (d <= 1).mean() >= 0.999 and np.array_equal(np.isnan(f), np.isnan(of))
    _, basic, _ = sc.render(w, h, depth, want_f32=False)
    assert (basic != u).any()  # the two modes do NOT give the same pixels
    # stats / size / trees_with (render_tree.rs:62-93)
    ost = np.zeros(8, np.uint64)
    orc.lib().orc_forest_stats(C.c_void_p(oh), ost.ctypes.data_as(C.POINTER(C.c_uint64)))
    st = forest.stats()
    assert list(st.values()) == list(ost)
    assert forest.size() == w * h
    for shape_id in range(0, 13):  # ids 0..11 alias cube triangles with top-level shapes
        assert forest.trees_with(shape_id) == orc.lib().orc_forest_trees_with(C.c_void_p(oh), shape_id)
    assert forest.build_stats["nodes"] == st["num_intersections"]
    orc.lib().orc_forest_free(C.c_void_p(oh))
    forest.close()
    sc.close()


def test_filtered_reshade_after_set_diffuse():
    """The GUI loop (src/gui.rs:163-235) and `bench --filter` (src/main.rs:170-205): edit the
    'blue' sphere's diffuse colour, then re-shade only the trees that hold shape 1."""
    desc = rt.SceneDesc.builtin("default")
    w, h, depth = 192, 128, 6
    sc = rt.Scene(desc)
    forest = sc.build_forest(w, h, depth)
    f, u, _ = forest.eval()
    mats = desc.materials_copy()
    mats[1].diffuse.rgb[:] = [1.0, 0.5, 0.0]
    sc.update_materials(mats)
    n, _ = forest.eval_filtered([1], f, u)
    assert n == forest.trees_with(1) and 0 < n < w * h

    edited = rt.SceneDesc.from_arrays((abi.rt_shape * 6)(*[desc.desc.shapes[i] for i in range(6)]), mats,
                                      (abi.rt_light * 3)(*[desc.desc.lights[i] for i in range(3)]), list(desc.desc.ambient))
    oh, _ = _orc_forest(desc, w, h, depth)
    of, ou, _ = _orc_eval(oh, desc, w, h)
    _, _, on = _orc_eval(oh, edited, w, h, ids=[1], into=(of, ou))
    assert on == n
    assert (u == ou).all()
    # an unfiltered eval with the new materials changes exactly the same pixels or more
    f2, u2, _ = forest.eval()
    o2f, o2u, _ = _orc_eval(oh, edited, w, h)
    assert (u2 == o2u).all()
    # empty mutated set: nothing is touched
    before = u.copy()
    n0, _ = forest.eval_filtered([], f, u)
    assert n0 == 0 and (u == before).all()
    orc.lib().orc_forest_free(C.c_void_p(oh))
    forest.close()
    sc.close()


def test_forest_explicit_ambient_argument():
    desc = rt.SceneDesc.builtin("default")
    sc = rt.Scene(desc)
    forest = sc.build_forest(96, 64, 4)
    _, a, _ = forest.eval()
    _, b, _ = forest.eval(ambient=(0.9, 0.1, 0.1))  # render_forest takes ambient explicitly (render_tree.rs:121)
    assert (a != b).any()
    forest.close()
    sc.close()
