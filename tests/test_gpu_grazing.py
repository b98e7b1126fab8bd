"""GPU: adversarial test of the triangle bound ("filter margins").

The reference's triangle test (src/scene/triangle.rs:51-80, Moller-Trumbore in f32) divides by
det = e1 . (d x e2), which goes to zero as the ray approaches the triangle's plane; it only
refuses |det| < f32::EPSILON.  For a ray within ~1e-5 rad of the plane the computed barycentric
coordinates can be off by a large fraction of the triangle, so the reference may report a hit for
a ray whose line passes OUTSIDE the triangle.  The sweep's filter hands a (ray, triangle) pair to
the exact test only if the line passes the triangle's bounding sphere (1.5 x the circumradius
about the centroid + margins): a reference hit outside that sphere would be a silently wrong
pixel.  These scenes aim whole bundles of such rays at thousands of triangles -- planes through
the camera (every primary ray towards the triangle grazes it), zoomed windows so that
neighbouring rays differ by 1e-6..1e-4 rad, hit points out at the vertices (u or v near 1),
triangle sizes 0.02..2, cameras inside the scene and 1e2..1e4 scene radii away -- and compare the
filtered sweep with the all-pairs audit, which runs every pair through the exact test the way the
reference does.  Zero differing pixels and equal node counts are required.
"""
import ctypes as C

import numpy as np
import pytest

import rust_tracer_b200 as rt
from rust_tracer_b200 import abi

pytestmark = pytest.mark.gpu


def _triangle(L, v0, v1, v2, material):
    s = abi.rt_shape()
    s.kind, s.material = abi.RT_SHAPE_TRIANGLE, material
    ident = [1.0, 0, 0, 0, 0, 1.0, 0, 0, 0, 0, 1.0, 0, 0, 0, 0, 1.0]
    s.transform[:] = ident
    s.inv_transform[:] = ident
    a, b, c = (np.asarray(v, dtype=np.float32) for v in (v0, v1, v2))
    n = np.zeros(3, np.float32)
    P = lambda x: x.ctypes.data_as(C.POINTER(C.c_float))
    L.rt_triangle_normal(P(a), P(b), P(c), P(n))
    s.p[:] = [float(x) for x in np.concatenate([a, b, c, n])]
    return s


def _grazing_scene(rng, origin, target, half_width, n_tri, size_lo, size_hi, tilt):
    """n_tri triangles whose planes pass within tilt[0]..tilt[1] rad of `origin`, scattered over the square of
    half-width `half_width` around `target` (a point of the z = 0 image plane) and over a range of
    depths behind it; plus 3 point lights and a few spheres so that shadow rays graze as well."""
    L = rt.lib()
    origin = np.asarray(origin, dtype=np.float64)
    target = np.asarray(target, dtype=np.float64)
    dist0 = np.linalg.norm(target - origin)
    shapes = []
    for j in range(n_tri):
        # a point the camera sees through the window, 1x..2.5x as far as the image plane
        aim = target + np.array([rng.uniform(-1, 1) * half_width, rng.uniform(-1, 1) * half_width, 0.0])
        a = (aim - origin) / np.linalg.norm(aim - origin)          # in-plane direction 1: the line of sight
        g = origin + a * dist0 * rng.uniform(1.0, 2.5)               # where the triangle sits
        r = rng.normal(size=3)
        b = np.cross(a, r)
        b /= np.linalg.norm(b)                                       # in-plane direction 2, perpendicular to the line of sight
        n = np.cross(a, b)
        # tilt the plane about b: the camera is off it by |g - o| sin eps; log-uniform over tilt[0]..tilt[1]
        eps = float(np.exp(rng.uniform(np.log(tilt[0]), np.log(tilt[1])))) * (1.0 if rng.uniform() < 0.5 else -1.0)
        a2 = a * np.cos(eps) + n * np.sin(eps)
        s = float(np.exp(rng.uniform(np.log(size_lo), np.log(size_hi))))
        # three vertices in the (tilted) plane, random shape, including slivers
        pts = [g + s * (rng.uniform(-1, 1) * a2 + rng.uniform(-1, 1) * b) for _ in range(3)]
        shapes.append(_triangle(L, pts[0], pts[1], pts[2], j % 3))
    base = rt.SceneDesc.builtin("default")
    mats = (abi.rt_material * 3)(base.desc.materials[0], base.desc.materials[1], base.desc.materials[2])
    lights = (abi.rt_light * 3)(*[base.desc.lights[i] for i in range(3)])
    for i in range(3):  # lights near the triangles' planes too: their shadow rays graze other triangles
        lights[i].kind = abi.RT_LIGHT_POINT
        lights[i].pos[:] = [float(x) for x in origin + (target - origin) * rng.uniform(0.2, 3.0) + rng.normal(size=3) * half_width]
        lights[i].color[:] = [0.4, 0.4, 0.4]
    arr = (abi.rt_shape * len(shapes))(*shapes)
    return rt.SceneDesc.from_arrays(arr, mats, lights, (0.1, 0.1, 0.1)), base


def _camera(w, h, origin, centre, half_width):
    cam = abi.default_camera(w, h)
    cam.origin[:] = [float(x) for x in origin]
    cam.x_min, cam.x_max = float(centre[0] - half_width), float(centre[0] + half_width)
    cam.y_min, cam.y_max = float(centre[1] - half_width), float(centre[1] + half_width)
    return cam


CASES = [
    # origin,               window centre, half width, triangles, sizes,        tilt (rad): log-uniform over the range
    ((0.0, 0.0, -8.0),      (0.0, 0.0),    3.0,        3000,      (0.05, 2.0),  (1e-6, 3e-2)),   # the reference's camera, whole view
    ((0.0, 0.0, -8.0),      (0.5, -0.3),   0.05,       3000,      (0.02, 0.5),  (1e-7, 1e-3)),   # zoomed 60x: 2.4e-5 rad between rays
    ((0.0, 0.0, -8.0),      (-1.0, 1.0),   0.002,      2000,      (0.02, 0.2),  (1e-8, 1e-4)),   # zoomed 1500x: 1e-6 rad between rays
    ((1.0, 0.5, -2.0),      (0.2, 0.1),    0.3,        3000,      (0.05, 1.0),  (1e-6, 1e-2)),   # a camera close to the triangles
    ((0.0, 0.0, -8.0e2),    (0.0, 0.0),    3.0,        2000,      (0.1, 2.0),   (1e-7, 3e-2)),   # origins 1e2 scene radii away ...
    ((0.0, 0.0, -8.0e3),    (0.0, 0.0),    3.0,        2000,      (0.1, 2.0),   (1e-7, 3e-2)),   # ... 1e3 ...
    ((3.0e4, 1.0e4, -8.0e4), (0.0, 0.0),   3.0,        2000,      (0.1, 2.0),   (1e-10, 1e-6)),  # ... 1e4: the "heavy" class
]


@pytest.mark.parametrize("case", range(len(CASES)))
def test_grazing_triangles_filter_equals_audit(case):
    origin, centre, half_width, n_tri, (size_lo, size_hi), tilt = CASES[case]
    rng = np.random.default_rng(1000 + case)
    desc, _keep = _grazing_scene(rng, origin, (centre[0], centre[1], 0.0), half_width, n_tri, size_lo, size_hi, tilt)
    w, h, depth = 512, 384, 3
    cam = _camera(w, h, origin, centre, half_width)
    sc = rt.Scene(desc)
    try:
        f1, u1, s1 = sc.render(w, h, depth, cam=cam)
        f2, u2, s2 = sc.render(w, h, depth, cam=cam, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS)
        with rt.tuned(origin_tables=0):  # the same rays through the general (7 FMA) filter
            sg = rt.Scene(desc)
            try:
                f3, u3, s3 = sg.render(w, h, depth, cam=cam)
            finally:
                sg.close()
    finally:
        sc.close()
    pairs = s2["rays"] * n_tri
    bad = int((~np.isclose(f1, f2, rtol=0, atol=0, equal_nan=True)).any(axis=-1).sum())
    bad_g = int((~np.isclose(f3, f2, rtol=0, atol=0, equal_nan=True)).any(axis=-1).sum())
    hit = float((s2["nodes"]) / max(1, s2["rays"] - s2["shadow_rays"]))
    print(f"case {case}: {s2['rays']} rays x {n_tri} triangles = {pairs:.3g} pairs, {hit:.2f} of tree rays hit; "
          f"pixels differing from the audit: origin-table filter {bad}, general filter {bad_g}")
    assert s2["nodes"] >= 100, "the scene no longer puts grazing triangles in front of the rays"
    assert bad == 0 and bad_g == 0
    assert s1["nodes"] == s2["nodes"] == s3["nodes"] and s1["rays"] == s2["rays"] == s3["rays"]
    assert pairs >= 1e8
