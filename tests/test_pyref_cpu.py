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


def test_default_scene_description_bit_for_bit(default_desc):
    """The host builder's flat description of src/my_scene.rs (transforms through Matrix::translate /
    rotate / scale / mat_mul, inverses through the reference's Gauss-Jordan, plane texture axes, every
    material and light constant) against the values pyref derives from the Rust source on its own."""
    shapes, lights, ambient = pyref.default_scene_values()
    d = default_desc.desc
    assert (d.n_shapes, d.n_materials, d.n_lights) == (len(shapes), len(shapes), len(lights))
    bits = lambda xs: [np.float32(x).view(np.uint32) for x in xs]
    ident = pyref.m_flat(pyref.m_identity())
    for i, (kind, transform, p12, mat) in enumerate(shapes):
        s = d.shapes[i]
        assert s.kind == kind and s.material == i
        t = pyref.m_flat(transform) if transform is not None else ident
        inv = pyref.m_flat(pyref.m_inverse(transform)) if transform is not None else ident
        assert bits(s.transform) == bits(t), i
        assert bits(s.inv_transform) == bits(inv), i
        if p12 is not None:
            assert bits(s.p) == bits(p12), i
        m = d.materials[i]
        mk, amb, dif, spec, power, refl, ridx = mat
        assert m.kind == mk
        for got, want in ((m.ambient, amb), (m.diffuse, dif), (m.specular, spec)):
            assert got.kind == want[0] and bits(got.rgb) == bits(want[1]), i
        assert bits([m.power, m.reflectivity, m.refraction_index]) == bits([power, refl, ridx])
    for i, (pos, col) in enumerate(lights):
        assert d.lights[i].kind == abi.RT_LIGHT_POINT
        assert bits(d.lights[i].pos) == bits(pos) and bits(d.lights[i].color) == bits(col)
    assert bits(d.ambient) == bits(ambient)


def test_matrix_inverse_bit_for_bit_on_random_transforms():
    """Gauss-Jordan of src/math/matrix.rs:88-142 (pivot swap, EPSILON-skipped rows, final division) in
    pyref against the host library's rt_matrix_inverse, on T.Rz.Rx.S compositions and on matrices that
    need the row swap."""
    import ctypes as C
    L = rt.lib()
    rng = np.random.default_rng(17)
    cases = []
    for _ in range(40):
        t, a, sc = rng.uniform(-3, 3, 3), rng.uniform(-180, 180, 2), rng.uniform(0.2, 2.5, 3)
        cases.append(pyref.m_mul(pyref.m_mul(pyref.m_mul(pyref.m_translate(*t), pyref.m_rotate_z(a[0])), pyref.m_rotate_x(a[1])),
                                 pyref.m_scale(*sc)))
    cases.append(pyref._rows((0, 1, 0, 2), (1, 0, 0, -1), (0, 0, 2, 0.5), (0, 0, 0, 1)))  # zero pivot: rows swap
    cases.append(pyref._rows((0, 0, 1, 0), (0, 1, 0, 0), (1, 0, 0, 0), (0, 0, 0, 1)))
    for m in cases:
        flat = (C.c_float * 16)(*[float(x) for x in pyref.m_flat(m)])
        out = (C.c_float * 16)()
        assert L.rt_matrix_inverse(flat, out) == 0
        want = pyref.m_flat(pyref.m_inverse(m))
        assert [np.float32(x).view(np.uint32) for x in out] == [np.float32(x).view(np.uint32) for x in want]
    singular = (C.c_float * 16)(*([0.0] * 16))
    assert L.rt_matrix_inverse(singular, out) == abi.RT_ERR_INVALID
    with pytest.raises(ValueError):
        pyref.m_inverse(pyref._rows(*([(0, 0, 0, 0)] * 4)))


def test_the_references_own_bench_scene():
    """`render_128x128` (src/render.rs:233-249), the one render workload the reference defines besides
    my_scene.rs: one sphere under scale(1, 2.25, 1), Phong(WHITE, RED, WHITE, 60, reflectivity 1, index 0),
    no lights, ambient BLACK (Scene::new), 128x128, depth 5.  Nothing emits light, so the frame is black;
    what the bench measures is the rays: primary + reflection rays off the sphere."""
    import ctypes as C
    shapes, mats, lights = (abi.rt_shape * 1)(), (abi.rt_material * 1)(), (abi.rt_light * 0)()
    t = pyref.m_scale(1.0, 2.25, 1.0)
    shapes[0].kind, shapes[0].material = abi.RT_SHAPE_SPHERE, 0
    shapes[0].transform[:] = [float(x) for x in pyref.m_flat(t)]
    shapes[0].inv_transform[:] = [float(x) for x in pyref.m_flat(pyref.m_inverse(t))]
    m = mats[0]
    m.kind = abi.RT_MATERIAL_PHONG
    for fn, rgb in ((m.ambient, (1, 1, 1)), (m.diffuse, (1, 0, 0)), (m.specular, (1, 1, 1))):
        fn.kind = abi.RT_COLOR_CONST
        fn.rgb[:] = rgb
    m.power, m.reflectivity, m.refraction_index = 60.0, 1.0, 0.0
    desc = rt.SceneDesc.from_arrays(shapes, mats, lights, (0.0, 0.0, 0.0))
    cam = abi.default_camera(128, 128)
    of, ou, oc = orc.render(desc.desc, cam, 5)
    assert (of == 0).all() and (ou == 0).all()
    hits = int(oc.nodes)
    assert oc.shadow_rays == 0 and oc.rays == 128 * 128 + hits  # every node reflects once (reflectivity 1), none refracts
    assert 0 < hits < 128 * 128
    small = abi.default_camera(32, 32)  # the second restatement on a thumbnail of the same scene
    f, u8, cnt = pyref.render(desc.desc, small, 5)
    sf, su, sc = orc.render(desc.desc, small, 5)
    assert np.array_equal(f, sf) and cnt["rays"] == sc.rays and cnt["nodes"] == sc.nodes
