"""CPU: the C++ oracle against a SECOND restatement of the reference, written independently from
the Rust sources in scalar numpy.float32 Python (tests/pyref.py).  The reference itself cannot
be run here (no Rust toolchain) and has no test or golden image for trace_ray, camera, shadows,
Fresnel, refraction, checkerboard or quantisation, so the oracle's fidelity rests on
known-answer tests plus line-by-line reading; two independent readings that agree BIT FOR BIT on
every f32 of a frame -- and on the ray and node counts -- are what can be had beyond that."""
import os

import numpy as np
import pytest

import orc
import pyref
import rust_tracer_b200 as rt
from rust_tracer_b200 import abi
from golden.make_golden import CASES

HERE = os.path.dirname(os.path.abspath(__file__))


def _same(desc, w, h, depth):
    cam = abi.default_camera(w, h)
    f, u8, cnt = pyref.render(desc, cam, depth)
    of, ou, oc = orc.render(desc, cam, depth)
    assert (f.view(np.uint32) == of.view(np.uint32)).all() or np.array_equal(f, of, equal_nan=True)
    assert (u8 == ou).all()
    assert cnt["rays"] == oc.rays and cnt["nodes"] == oc.nodes
    return f, u8


def test_default_scene_bit_for_bit():
    """src/my_scene.rs: spheres (one under T.Rz.S), both planes with the checkerboard, the cube,
    three coloured point lights; depth 6 reaches refraction inside the glass sphere and the cube."""
    d = rt.SceneDesc.builtin("default")
    f, _ = _same(d.desc, 64, 48, 6)
    assert np.isfinite(f).all() and f.max() > 0.5


def test_default_scene_depth_semantics():
    d = rt.SceneDesc.builtin("default")
    f0, _ = _same(d.desc, 16, 12, 0)
    assert (f0 == 0).all()           # depth 0 renders black (src/render.rs:43-45)
    f1, _ = _same(d.desc, 16, 12, 1)
    assert f1.max() > 0              # depth 1: no reflected / refracted energy, lights only


@pytest.mark.parametrize("seed", [5, 11])
def test_synthetic_scene_bit_for_bit(seed):
    """Top-level triangles, anisotropic spheres (every 10th), mirror and glass materials, 4 lights."""
    d = rt.SceneDesc.synthetic(30, 30, seed, 0.1, 0.4)
    _same(d.desc, 40, 30, 4)


def test_golden_fixture_is_what_the_second_restatement_renders():
    """The committed golden image (an oracle output) equals the independent restatement too."""
    name = "default_basic_96x72_d8"
    mk, w, h, depth, mode, spp, seed = CASES[name]
    assert mode == 0 and spp == 1
    g = np.load(os.path.join(HERE, "golden", name + ".npz"))
    owner = mk()  # keeps the arrays behind the description alive
    f, u8, cnt = pyref.render(owner.desc, abi.default_camera(w, h), depth)
    assert np.array_equal(f, g["rgb_f32"], equal_nan=True) and (u8 == g["rgb_u8"]).all()
    assert [cnt["rays"], cnt["nodes"]] == [int(g["counters"][0]), int(g["counters"][2])]


def test_scalar_pieces_match_the_oracle():
    """Schlick's Fresnel term and Color::as_u8 on edge values, through both restatements."""
    L = orc.lib()
    for x in (-1.0, -0.0, 0.0, 0.5, 0.999, 1.0, 1.0039216, 2.0, 300.0, float("nan"), float("inf")):
        assert pyref.as_u8(np.float32(x)) == L.orc_as_u8(x)


def test_rayforest_mode_bit_for_bit():
    """render_tree.rs: generate_ray_forest + render_forest (its own weights: the reflected term takes
    eye_dir as the light direction, the refracted term is not multiplied by the diffuse colour), and
    the per-tree shape-id sets behind trees_with / render_forest_filter, cube aliasing included."""
    import ctypes as C
    d = rt.SceneDesc.builtin("default")
    w, h, depth = 48, 36, 6
    cam = abi.default_camera(w, h)
    f, u8, cnt, ids = pyref.render_forest(d.desc, cam, depth)
    of, ou, oc = orc.render(d.desc, cam, depth, mode=1)
    assert np.array_equal(f, of, equal_nan=True) and (u8 == ou).all()
    assert cnt["rays"] == oc.rays and cnt["nodes"] == oc.nodes
    bf, _, _ = orc.render(d.desc, cam, depth, mode=0)
    assert not np.array_equal(f, bf)  # the two renderers of the reference do differ
    L = orc.lib()
    handle = L.orc_forest_build(C.byref(d.desc), C.byref(cam), depth, 4)
    try:
        for shape_id in range(0, 13):  # ids 6..11 can only be inner triangles of the cube
            mine = sum(1 for s in ids.values() if shape_id in s)
            assert mine == L.orc_forest_trees_with(C.c_void_p(handle), shape_id), shape_id
        assert sum(1 for s in ids.values() if 7 in s) > 0  # the aliasing quirk is exercised
    finally:
        L.orc_forest_free(C.c_void_p(handle))


def test_forest_golden_fixture_is_what_the_second_restatement_renders():
    name = "default_forest_96x72_d8"
    mk, w, h, depth, mode, spp, seed = CASES[name]
    assert mode == 1 and spp == 1
    g = np.load(os.path.join(HERE, "golden", name + ".npz"))
    owner = mk()
    f, u8, cnt, _ = pyref.render_forest(owner.desc, abi.default_camera(w, h), depth)
    assert np.array_equal(f, g["rgb_f32"], equal_nan=True) and (u8 == g["rgb_u8"]).all()
