"""Pins the oracle against every known-answer unit test the reference holds for the render
path.  Each test names the reference test it restates (file:line in
/root/reference/src).  CPU only."""
import ctypes as C

import numpy as np
import pytest

import orc
from rust_tracer_b200 import abi

EPS = np.float32(np.finfo(np.float32).eps)
L = orc.lib()
fp = orc.fptr


def mat(fn, *a):
    out = np.zeros(16, np.float32)
    fn(*a, fp(out))
    return out.reshape(4, 4)


def scale(x, y, z):
    return mat(L.orc_mat_scale, C.c_float(x), C.c_float(y), C.c_float(z))


def translate(x, y, z):
    return mat(L.orc_mat_translate, C.c_float(x), C.c_float(y), C.c_float(z))


def rotate(axis, deg):
    return mat(L.orc_mat_rotate, C.c_int(axis), C.c_float(deg))


def mul(a, b):
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    return mat(L.orc_mat_mul, fp(a), fp(b))


def inverse(a):
    a = np.ascontiguousarray(a, np.float32)
    out = np.zeros(16, np.float32)
    rc = L.orc_mat_inverse(fp(a), fp(out))
    return rc, out.reshape(4, 4)


def pt_mul(m, p):
    m, p = np.ascontiguousarray(m, np.float32), orc.f32(*p)
    out = np.zeros(3, np.float32)
    L.orc_mat_pt_mul(fp(m), fp(p), fp(out))
    return out


def vec_mul(m, v):
    m, v = np.ascontiguousarray(m, np.float32), orc.f32(*v)
    out = np.zeros(3, np.float32)
    L.orc_mat_vec3_mul(fp(m), fp(v), fp(out))
    return out


# ---- src/math/matrix.rs ------------------------------------------------------------------
def test_matrix_creation_and_transformations():  # matrix.rs:304 creation, :431 transformations
    ident = mat(L.orc_mat_identity)
    assert (ident == np.eye(4, dtype=np.float32)).all()
    assert (scale(2, 3, 4) == np.diag([2, 3, 4, 1]).astype(np.float32)).all()
    t = np.eye(4, dtype=np.float32)
    t[:3, 3] = [2, 3, 4]
    assert (translate(2, 3, 4) == t).all()  # translation lives in column 3


def test_matrix_transpose():  # matrix.rs:354
    tr = lambda a: mat(L.orc_mat_transpose, fp(np.ascontiguousarray(a, np.float32)))
    assert (tr(scale(2, 3, 4)) == scale(2, 3, 4)).all()
    t = translate(1, 2, 3)
    assert not (tr(t) == t).all() and (tr(tr(t)) == t).all()
    for axis in range(3):
        assert np.abs(tr(rotate(axis, 90.)) - rotate(axis, 270.)).max() < EPS


def test_matrix_mul():  # matrix.rs:380
    s = scale(2, 2, 2)
    assert np.abs(mul(s, mat(L.orc_mat_identity)) - s).max() < EPS
    assert np.abs(mul(s, s) - scale(4, 4, 4)).max() < EPS


def test_matrix_inverse():  # matrix.rs:395
    ident = np.eye(4, dtype=np.float32)
    rc, inv = inverse(scale(2, 3, 4))
    assert rc == 0 and np.abs(mul(scale(2, 3, 4), inv) - ident).max() < 2 * EPS
    rc, inv = inverse(ident)
    assert rc == 0 and np.abs(inv - ident).max() < 2 * EPS
    t = translate(2, 2, -4)
    rc, inv = inverse(t)
    assert rc == 0 and np.abs(mul(inv, t) - ident).max() < 2 * EPS
    sh = ident.copy()
    sh[1, 3] = 4
    rc, inv = inverse(sh)
    assert rc == 0 and np.abs(mul(inv, sh) - ident).max() < 2 * EPS
    r = rotate(0, 82.)
    rc, inv = inverse(r)
    assert rc == 0 and np.abs(mul(r, inv) - ident).max() < 2 * EPS


def test_matrix_inverse_singular_panics():  # matrix.rs:106 panic!("Singular Matrix")
    rc, _ = inverse(np.zeros((4, 4), np.float32))
    assert rc == 1


def test_pt_and_vec_transforms():  # matrix.rs:459-532, point.rs:161-230, vector3.rs:230-300
    assert (pt_mul(scale(2, 3, 4), (1, 1, 1)) == [2, 3, 4]).all()
    assert (pt_mul(translate(2, 3, 4), (1, 1, 1)) == [3, 4, 5]).all()
    assert (vec_mul(translate(2, 3, 4), (1, 1, 1)) == [1, 1, 1]).all()  # vectors ignore translation
    assert np.abs(pt_mul(rotate(0, 90.), (1, 1, 1)) - [1, -1, 1]).max() < EPS
    assert np.abs(pt_mul(rotate(1, 90.), (1, 1, 1)) - [1, 1, -1]).max() < EPS
    assert np.abs(pt_mul(rotate(2, 90.), (1, 1, 1)) - [-1, 1, 1]).max() < EPS


def test_ray_transform():  # ray.rs:198-226: origin takes the translation, direction does not
    d = _one_shape_desc(abi.RT_SHAPE_SPHERE, translate(0, 0, 0))
    # exercised through Sphere::intersect below; here the primitive ops:
    assert (pt_mul(translate(1, 2, 3), (0, 0, 0)) == [1, 2, 3]).all()
    assert (vec_mul(translate(1, 2, 3), (0, 0, 1)) == [0, 0, 1]).all()
    del d


# ---- src/scene/color.rs:192 ----------------------------------------------------------------
def test_color_ops():
    a, b = orc.f32(0.2, 0.2, 0.2), orc.f32(0.1, 0.2, 0.3)
    m, sm, ad = (np.zeros(3, np.float32) for _ in range(3))
    L.orc_color_ops(fp(a), fp(b), C.c_float(0.1), fp(m), fp(sm), fp(ad))
    assert np.abs(m - [0.02, 0.04, 0.06]).max() < EPS
    assert np.abs(sm - [0.01, 0.02, 0.03]).max() < EPS
    assert np.abs(ad - [0.3, 0.4, 0.5]).max() < EPS


def test_as_u8_saturates_truncates_nan_to_zero():  # color.rs:43-46 with Rust >= 1.45 casts
    f = lambda x: L.orc_as_u8(C.c_float(x))
    assert f(0.0) == 0 and f(1.0) == 255 and f(2.5) == 255 and f(-0.3) == 0
    assert f(0.5) == 127 and f(float("nan")) == 0 and f(float("inf")) == 255
    assert f(0.999) == 254


# ---- shapes ----------------------------------------------------------------------------------
def _material(kind=abi.RT_MATERIAL_PHONG, amb=(1, 1, 1), dif=(1, 1, 1), spec=(1, 1, 1), power=60., refl=1., idx=0.):
    m = abi.rt_material()
    m.kind = kind
    for f, c in ((m.ambient, amb), (m.diffuse, dif), (m.specular, spec)):
        f.kind = abi.RT_COLOR_CONST
        f.rgb[:] = list(c)
    m.power, m.reflectivity, m.refraction_index = power, refl, idx
    return m


def _one_shape_desc(kind, transform=None, p=None, material=None):
    shapes = (abi.rt_shape * 1)()
    s = shapes[0]
    s.kind, s.material = kind, 0
    t = np.eye(4, dtype=np.float32) if transform is None else np.ascontiguousarray(transform, np.float32)
    rc, inv = inverse(t)
    assert rc == 0
    s.transform[:] = list(t.reshape(-1))
    s.inv_transform[:] = list(inv.reshape(-1))
    if p is not None:
        s.p[:] = list(p)
    mats = (abi.rt_material * 1)(material or _material())
    lights = (abi.rt_light * 0)()
    return orc.SceneArrays(shapes, mats, lights, (0, 0, 0))


def _intersect(sa, o, d):
    rec = np.zeros(15, np.float32)
    hit = L.orc_intersect(C.byref(sa.desc), fp(orc.f32(*o)), fp(orc.f32(*d)), fp(rec))
    return hit, rec


def test_sphere_intersection_no_transform():  # sphere.rs:178
    sa = _one_shape_desc(abi.RT_SHAPE_SPHERE)
    hit, rec = _intersect(sa, (0, 0, 2), (0, 0, -1))
    assert hit == 1 and rec[1] == 1.0
    hit, _ = _intersect(sa, (0, 0, 2), (0, 1, 0))
    assert hit == 0
    hit, rec = _intersect(sa, (0, 1, 2), (0, 0, -1))  # tangent ray
    assert hit == 1 and rec[1] == 2.0


def test_sphere_intersection_transform():  # sphere.rs:198
    sa = _one_shape_desc(abi.RT_SHAPE_SPHERE, mul(translate(0, 2, -2), scale(2, 2, 2)))
    hit, rec = _intersect(sa, (0, 0, 2), (0, 0, -1))
    assert hit == 1 and rec[1] == 4.0
    hit, _ = _intersect(sa, (0, 0, 2), (0, 1, 0))
    assert hit == 0
    hit, rec = _intersect(sa, (0, 2, 2), (0, 0, -1))
    assert hit == 1 and rec[1] == 2.0


def _tri_p(v0, v1, v2):
    n = np.zeros(3, np.float32)
    L.orc_triangle_normal(fp(orc.f32(*v0)), fp(orc.f32(*v1)), fp(orc.f32(*v2)), fp(n))
    return list(v0) + list(v1) + list(v2) + list(n), n


def test_triangle_creation():  # triangle.rs:130: CCW -> +Z, CW -> -Z
    _, n = _tri_p((0, 0, 0), (1, 0, 0), (0, 1, 0))
    assert np.abs(n - [0, 0, 1]).max() < EPS
    _, n = _tri_p((1, 0, 0), (0, 0, 0), (0, 1, 0))
    assert np.abs(n - [0, 0, -1]).max() < EPS


def test_triangle_intersection():  # triangle.rs:152
    p, _ = _tri_p((2, -2, 0), (-2, -2, 0), (-2, 2, 0))
    sa = _one_shape_desc(abi.RT_SHAPE_TRIANGLE, p=p)
    hit, rec = _intersect(sa, (0, 0, -4), (0, 0, 1))
    assert hit == 1
    assert rec[1] == 4.0                                   # t
    assert (rec[3:6] == [0, 0, 0]).all()                   # point
    assert np.abs(rec[9:12] - [0, 0, -1]).max() < EPS      # normal
    assert np.abs(rec[6:9] - [0, 0, -1]).max() < EPS       # eye_dir
    assert rec[12] == 1.0                                  # entering


def test_triangle_behind_ray_not_intersection():  # triangle.rs:176
    p, _ = _tri_p((2, -2, 0), (-2, -2, 0), (-2, 2, 0))
    sa = _one_shape_desc(abi.RT_SHAPE_TRIANGLE, p=p)
    hit, _ = _intersect(sa, (0, 0, -4), (0, 0, -1))
    assert hit == 0


def test_triangle_shading():  # triangle.rs:193: 0.5-grey Phong lit head-on reflects WHITE
    p, _ = _tri_p((2, -1, 0), (-1, -1, 0), (-1, 2, 0))
    m = _material(amb=(.5, .5, .5), dif=(.5, .5, .5), spec=(.5, .5, .5), power=60., refl=0., idx=0.)
    sa = _one_shape_desc(abi.RT_SHAPE_TRIANGLE, p=p, material=m)
    hit, rec = _intersect(sa, (0, 0, -4), (0, 0, 1))
    assert hit == 1
    point, eye, normal = rec[3:6], rec[6:9], rec[9:12]
    ldir = np.zeros(3, np.float32)
    L.orc_vec3_norm(fp(orc.f32(*(np.float32([0, 0, -4]) - point))), fp(ldir))
    out = np.zeros(3, np.float32)
    L.orc_reflected_energy(C.byref(m), fp(orc.f32(1, 1, 1)), fp(ldir), fp(eye.copy()), fp(normal.copy()),
                           C.c_float(rec[13]), C.c_float(rec[14]), fp(out))
    assert np.abs(out - [1, 1, 1]).max() < EPS


def test_plane_texture_coords():  # plane.rs:122: u axis is perpendicular to the normal
    u, v = np.zeros(3, np.float32), np.zeros(3, np.float32)
    n = orc.f32(0, 1, 0)
    L.orc_plane_axes(fp(n), fp(u), fp(v))
    assert float(np.dot(n, u)) == 0.0


def test_plane_negative_t_is_a_hit():  # plane.rs:59-84 has no t >= 0 rejection (hard part 2)
    u, v = np.zeros(3, np.float32), np.zeros(3, np.float32)
    L.orc_plane_axes(fp(orc.f32(0, 1, 0)), fp(u), fp(v))
    sa = _one_shape_desc(abi.RT_SHAPE_PLANE, p=[0, -2, 0, 0, 1, 0] + list(u) + list(v))
    hit, rec = _intersect(sa, (0, -3, 0), (0, -1, 0))  # below the floor, heading down
    assert hit == 1 and rec[1] == -1.0 and rec[12] == 0.0  # t < 0, entering = false
    hit, _ = _intersect(sa, (0, 0, 0), (0, 1, 0))  # one-sided: from behind the normal there is no hit
    assert hit == 0


def test_cube_keeps_inner_triangle_id():  # cube.rs:93-101 never overwrites the triangle's id
    sa = _one_shape_desc(abi.RT_SHAPE_CUBE)
    hit, rec = _intersect(sa, (0.1, 0.2, -4), (0, 0, 1))
    assert hit == 1 and rec[0] in (0.0, 1.0) and abs(rec[1] - 3.5) < 1e-6  # front face = triangles 0, 1


# ---- src/render_tree.rs:265 ray_tree_size -----------------------------------------------------
def test_ray_tree_size():
    sa = orc.default_scene()
    cam = abi.default_camera(32, 32)
    for depth, hi in ((1, 1), (2, 3), (3, 7)):
        f = L.orc_forest_build(C.byref(sa.desc), C.byref(cam), depth, 1)
        sizes = [L.orc_forest_tree_size(C.c_void_p(f), u, v) for u in range(32) for v in range(32)]
        assert 1 <= min(sizes) and max(sizes) <= hi  # a binary tree of depth d has at most 2^d - 1 nodes
        L.orc_forest_free(C.c_void_p(f))


# ---- workload shape measured by the survey -----------------------
@pytest.mark.parametrize("res,depth,rays_px,nodes_px", [(256, 8, 10.11, 2.48), (128, 5, 8.11, 2.00), (128, 3, 6.54, 1.63)])
def test_default_scene_workload_shape(res, depth, rays_px, nodes_px):
    sa = orc.default_scene()
    _, _, c = orc.render(sa.desc, abi.default_camera(res, res), depth, want_f32=False)
    px = res * res
    assert abs(c.rays / px - rays_px) < 0.006
    assert abs(c.nodes / px - nodes_px) < 0.006
    assert (c.sphere_tests + c.plane_tests + 12 * c.cube_tests) == 17 * c.rays
