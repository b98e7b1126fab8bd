"""A SECOND, independent restatement of the reference's per-pixel path, in scalar Python over
numpy.float32 -- TEST INFRASTRUCTURE, written from the Rust sources alone (not from oracle/), so
that a transcription slip in the C++ oracle shows up as a bit difference between the two
(tests/test_pyref_cpu.py).  Every function cites the reference lines it follows; every
arithmetic step is one IEEE f32 operation in the reference's own order (numpy.float32 scalars
round after each operation, like Rust's f32; `powf`, `atan2f`, `acosf` are glibc's, called
through ctypes, which is what Rust's f32 methods reach on linux-gnu).

Far too slow for anything but thumbnails (a few hundred pixels): that is all it is for.
Geometry comes in as the flat rt_scene_desc (matrices and their inverses are the host's, whose
Gauss-Jordan has its own known-answer tests, tests/test_oracle_kat.py)."""
import ctypes as C
import ctypes.util

import numpy as np

F = np.float32
ZERO, ONE, TWO, HALF = F(0.0), F(1.0), F(2.0), F(0.5)
EPS = F(np.finfo(np.float32).eps)  # std::f32::EPSILON
PI = F(np.pi)                      # std::f32::consts::PI
OFFSET = F(0.0002)

_libm = C.CDLL(ctypes.util.find_library("m") or "libm.so.6")
for _name, _n in (("powf", 2), ("atan2f", 2), ("acosf", 1)):
    _fn = getattr(_libm, _name)
    _fn.restype = C.c_float
    _fn.argtypes = [C.c_float] * _n


def powf(a, b):
    return F(_libm.powf(float(a), float(b)))


def atan2f(a, b):
    return F(_libm.atan2f(float(a), float(b)))


def acosf(a):
    return F(_libm.acosf(float(a)))


# ---- src/math/vector3.rs, point.rs (tuples of three f32) -------------------------------------------

def vadd(a, b):  # vector3.rs:56-62 / point.rs add
    return (a[0] + b[0], a[1] + b[1], a[2] + b[2])


def vsub(a, b):  # vector3.rs:65-71 / point.rs sub
    return (a[0] - b[0], a[1] - b[1], a[2] - b[2])


def vneg(a):  # vector3.rs:30-36
    return (-a[0], -a[1], -a[2])


def vscale(a, s):  # scalar_mul: component * a (vector3.rs:39-45)
    return (a[0] * s, a[1] * s, a[2] * s)


def vdiv(a, d):  # vector3.rs:48-54
    return (a[0] / d, a[1] / d, a[2] / d)


def dot(a, b):  # vector3.rs:85-87
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]


def len2(a):  # vector3.rs:74-76
    return a[0] * a[0] + a[1] * a[1] + a[2] * a[2]


def norm(a):  # vector3.rs:90-93
    return vdiv(a, np.sqrt(len2(a)))


def cross(a, b):  # vector3.rs:96-102
    return (a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])


def reflect(a, about):  # vector3.rs:112-114: 2. * (self.dot(about)) * about - self
    return vsub(vscale(about, TWO * dot(a, about)), a)


# ---- src/math/matrix.rs (m = 16 f32, row-major) -----------------------------------------------------

def vec3_mul(m, v):  # matrix.rs:229-235
    return (v[0] * m[0] + v[1] * m[1] + v[2] * m[2],
            v[0] * m[4] + v[1] * m[5] + v[2] * m[6],
            v[0] * m[8] + v[1] * m[9] + v[2] * m[10])


def pt_mul(m, p):  # matrix.rs:237-252
    return (p[0] * m[0] + p[1] * m[1] + p[2] * m[2] + m[3],
            p[0] * m[4] + p[1] * m[5] + p[2] * m[6] + m[7],
            p[0] * m[8] + p[1] * m[9] + p[2] * m[10] + m[11])


def transpose(m):  # matrix.rs:77-86
    return [m[c * 4 + r] for r in range(4) for c in range(4)]


def ray_mul(m, ray):  # matrix.rs:254-256
    return (pt_mul(m, ray[0]), vec3_mul(m, ray[1]))


def ray_at(ray, t):  # `t * ray` = origin + direction * t (ray.rs:50-52, 63-77)
    return vadd(ray[0], vscale(ray[1], t))


# ---- src/math/matrix.rs: construction (4x4 lists of f32) and the Gauss-Jordan inverse -----------------

for _name in ("sinf", "cosf"):
    _fn = getattr(_libm, _name)
    _fn.restype = C.c_float
    _fn.argtypes = [C.c_float]


def _rows(*rows):
    return [[F(x) for x in r] for r in rows]


def m_identity():
    return _rows((1, 0, 0, 0), (0, 1, 0, 0), (0, 0, 1, 0), (0, 0, 0, 1))


def m_scale(x, y, z):  # matrix.rs:144-153
    return _rows((x, 0, 0, 0), (0, y, 0, 0), (0, 0, z, 0), (0, 0, 0, 1))


def m_translate(x, y, z):  # matrix.rs:155-164
    return _rows((1, 0, 0, x), (0, 1, 0, y), (0, 0, 1, z), (0, 0, 0, 1))


def _rads(angle):  # angle / 180.0 * pi
    return F(angle) / F(180.0) * PI


def m_rotate_x(angle):  # matrix.rs:166-178
    r = _rads(angle)
    c, s = F(_libm.cosf(float(r))), F(_libm.sinf(float(r)))
    return _rows((1, 0, 0, 0), (0, c, -s, 0), (0, s, c, 0), (0, 0, 0, 1))


def m_rotate_z(angle):  # matrix.rs:194-206
    r = _rads(angle)
    c, s = F(_libm.cosf(float(r))), F(_libm.sinf(float(r)))
    return _rows((c, -s, 0, 0), (s, c, 0, 0), (0, 0, 1, 0), (0, 0, 0, 1))


def m_mul(a, b):  # matrix.rs:57-75: sum starts at 0., += s[row][i] * a[i][col]
    out = []
    for row in range(4):
        r = []
        for col in range(4):
            acc = ZERO
            for i in range(4):
                acc = acc + a[row][i] * b[i][col]
            r.append(acc)
        out.append(r)
    return out


def m_inverse(m):  # matrix.rs:88-142 (panics "Singular Matrix" -> ValueError here)
    m = [list(r) for r in m]
    inv = m_identity()
    for col in range(4):
        if abs(m[col][col]) < EPS:
            big = col
            for row in range(4):
                if abs(m[row][col]) > abs(m[big][col]):
                    big = row
            if big == col:
                raise ValueError("Singular Matrix")
            for j in range(4):
                m[big][j], m[col][j] = m[col][j], m[big][j]
                inv[big][j], inv[col][j] = inv[col][j], inv[big][j]
        for row in range(4):
            if row != col:
                coeff = m[row][col] / m[col][col]
                if abs(coeff) >= EPS:
                    for j in range(4):
                        m[row][j] = m[row][j] - coeff * m[col][j]
                        inv[row][j] = inv[row][j] - coeff * inv[col][j]
                    m[row][col] = ZERO
    for row in range(4):
        for col in range(4):
            inv[row][col] = inv[row][col] / m[row][row]
    return inv


def m_flat(m):
    return [x for r in m for x in r]


def plane_axes(normal):  # Plane::new, plane.rs:22-30
    w = (ZERO, ONE, ZERO) if np.sqrt(len2(cross(normal, (ONE, ZERO, ZERO)))) <= EPS else (ONE, ZERO, ZERO)
    u = norm(cross(normal, w))
    return u, norm(cross(normal, u))


def default_scene_values():
    """src/my_scene.rs:45-120 read a third time: [(kind, transform | None, p12 | None, material)], lights, ambient.
    material = (kind, ambient fn, diffuse fn, specular fn, power, reflectivity, refraction index),
    fn = (0, rgb) constant | (1, rgb) checkerboard."""
    W, K = (ONE, ONE, ONE), (ZERO, ZERO, ZERO)
    dim_white = (F(0.1) * ONE, F(0.1) * ONE, F(0.1) * ONE)  # fn dim_white: 0.1 * WHITE
    const = lambda c: (0, tuple(F(x) for x in c))
    phong = lambda a, d, sp, pw, rf, ri: (0, const(a), const(d), const(sp), F(pw), F(rf), F(ri))
    tex = (1, (0, dim_white), (1, W), (0, dim_white), F(60.0), F(0.0), F(0.0))
    t1 = m_mul(m_mul(m_translate(-1.0, 0.0, 0.0), m_rotate_z(75.0)), m_scale(1.0, 0.25, 1.0))
    t2 = m_translate(1.0, -1.0, 0.0)
    t3 = m_mul(m_translate(0.0, -0.5, -3.0), m_scale(0.6, 0.6, 0.6))
    t6 = m_mul(m_translate(-1.0, -1.0, -4.0), m_rotate_x(-45.0))

    def plane(origin, normal):
        o, n = tuple(F(x) for x in origin), tuple(F(x) for x in normal)
        u, v = plane_axes(n)
        return list(o) + list(n) + list(u) + list(v)

    shapes = [
        (0, t1, None, phong((0.1, 0.1, 0.1), (1, 0, 0), W, 60.0, 0.5, 0.0)),
        (0, t2, None, phong(K, (0, 0, 1), (0, 0, 0.1), 600.0, 0.4, 0.0)),
        (0, t3, None, phong(K, W, W, 60.0, 0.7, 1.333)),
        (1, None, plane((0, -2, 2), (0, 0, -1)), tex),
        (1, None, plane((0, -2, 0), (0, 1, 0)), tex),
        (3, t6, None, phong(K, W, W, 60.0, 0.0, 1.333)),
    ]
    lights = [((4, 4, 0), (1, 0, 0)), ((-1, 2, -4), (0, 1, 0)), ((0, 8, -4), (0, 0, 1))]
    return shapes, lights, (F(0.1), F(0.1), F(0.1))


# ---- scene loading ----------------------------------------------------------------------------------

def _f(seq):
    return [F(x) for x in seq]


class SceneData:
    def __init__(self, desc):
        self.shapes = []
        for i in range(desc.n_shapes):
            s = desc.shapes[i]
            p = _f(s.p)
            self.shapes.append(dict(kind=s.kind, id=i, material=s.material, transform=_f(s.transform), inv=_f(s.inv_transform),
                                    a=tuple(p[0:3]), b=tuple(p[3:6]), c=tuple(p[6:9]), d=tuple(p[9:12])))
        self.materials = []
        for i in range(desc.n_materials):
            m = desc.materials[i]
            fn = lambda cf: (cf.kind, tuple(_f(cf.rgb)))
            self.materials.append(dict(kind=m.kind, ambient=fn(m.ambient), diffuse=fn(m.diffuse), specular=fn(m.specular),
                                       power=F(m.power), reflectivity=F(m.reflectivity), refraction_index=F(m.refraction_index)))
        self.lights = [dict(kind=desc.lights[i].kind, pos=tuple(_f(desc.lights[i].pos)), color=tuple(_f(desc.lights[i].color)))
                       for i in range(desc.n_lights)]
        self.ambient = tuple(_f(desc.ambient))
        self.cube_tris = _cube_triangles()


# ---- src/scene/triangle.rs ---------------------------------------------------------------------------

def triangle_normal(v0, v1, v2):  # Triangle::new, triangle.rs:25-29 (v0v2 = v2 - v1, as written there)
    return norm(cross(vsub(v1, v0), vsub(v2, v1)))


def intersect_triangle(v0, v1, v2, normal, ray, sid, material):  # triangle.rs:51-94
    o, d = ray
    v0v1 = vsub(v1, v0)
    v0v2 = vsub(v2, v0)
    pvec = cross(d, v0v2)
    det = dot(v0v1, pvec)
    if abs(det) < EPS:
        return None
    inv_det = ONE / det
    tvec = vsub(o, v0)
    u = dot(tvec, pvec) * inv_det
    if u < ZERO or u > ONE:
        return None
    qvec = cross(tvec, v0v1)
    v = dot(d, qvec) * inv_det
    if v < ZERO or u + v > ONE:
        return None
    t = dot(v0v2, qvec) * inv_det
    if t < ZERO:
        return None
    return dict(id=sid, t=t, material=material, point=ray_at(ray, t), eye_dir=vneg(norm(d)), normal=normal,
                entering=bool(det > ZERO), tex=(u, v))


# ---- src/scene/cube.rs -------------------------------------------------------------------------------

def _cube_triangles():  # Cube::new, cube.rs:21-71: vertices, triangles, and the ORDER they are added in
    h = HALF
    v = [(h, h, -h), (h, -h, -h), (-h, -h, -h), (-h, h, -h), (h, h, h), (-h, h, h), (-h, -h, h), (h, -h, h)]
    tf1, tf2 = (1, 2, 3), (0, 1, 3)
    tk1, tk2 = (7, 5, 4), (5, 7, 6)
    tr1, tr2 = (0, 4, 7), (7, 1, 0)
    tl1, tl2 = (5, 3, 6), (6, 3, 2)
    tb1, tb2 = (1, 7, 6), (6, 2, 1)
    tt1, tt2 = (5, 4, 0), (0, 3, 5)
    order = [tf1, tf2, tk1, tk2, tr1, tr2, tl1, tl2, tt1, tt2, tb1, tb2]
    return [(v[a], v[b], v[c], triangle_normal(v[a], v[b], v[c])) for a, b, c in order]


def intersect_cube(scene, s, ray):  # cube.rs:89-102
    tr = ray_mul(s["inv"], ray)
    best = None
    for k, (a, b, c, n) in enumerate(scene.cube_tris):  # the inner Scene::intersect; ids are 0..11 (add_shape)
        i = intersect_triangle(a, b, c, n, tr, k, s["material"])
        if i is not None and (best is None or i["t"] < best["t"]):
            best = i
    if best is None:
        return None
    best["point"] = ray_at(ray, best["t"])
    best["eye_dir"] = vneg(norm(ray[1]))
    best["normal"] = norm(vec3_mul(transpose(s["inv"]), best["normal"]))
    return best  # id stays the inner triangle's (the aliasing quirk: cube.rs never resets it)


# ---- src/scene/sphere.rs -----------------------------------------------------------------------------

def solve_quadratic(a, b, c):  # sphere.rs:126-145
    discr = b * b - F(4.0) * a * c
    if discr < ZERO:
        return None
    if abs(discr) < EPS:
        x = F(-0.5) * b / a
        return (x, x)
    q = F(-0.5) * (b + np.sqrt(discr)) if b > ZERO else F(-0.5) * (b - np.sqrt(discr))
    return (q / a, c / q)


def intersect_sphere(s, ray):  # sphere.rs:57-98
    tr = ray_mul(s["inv"], ray)
    l = vsub(tr[0], (ZERO, ZERO, ZERO))
    a = len2(tr[1])
    b = TWO * dot(tr[1], l)
    c = len2(l) - ONE
    sol = solve_quadratic(a, b, c)
    if sol is None:
        return None
    t0, t1 = sol
    if t0 > t1:
        t0, t1 = t1, t0
    if t0 < ZERO and t1 < ZERO:
        return None
    t = t1 if t0 < ZERO else t0
    entering = bool(t0 > ZERO)
    point = ray_at(ray, t)
    normal = norm(vec3_mul(transpose(s["inv"]), ray_at(tr, t)))
    if not entering:
        normal = vneg(normal)
    u = (ONE + atan2f(normal[2], normal[0]) / PI) * HALF  # get_texture_coord, sphere.rs:40-45
    v = acosf(normal[1]) / PI
    return dict(id=s["id"], t=t, material=s["material"], point=point, eye_dir=vneg(norm(ray[1])), normal=normal,
                entering=entering, tex=(u, v))


# ---- src/scene/plane.rs ------------------------------------------------------------------------------

def intersect_plane(s, ray):  # plane.rs:59-84; a = origin, b = normal, c = u axis, d = v axis
    tr = ray_mul(s["inv"], ray)
    denom = -dot(s["b"], tr[1])
    if not (denom > EPS):
        return None
    dirv = vsub(s["a"], tr[0])
    t = -dot(dirv, s["b"]) / denom
    point = ray_at(ray, t)
    return dict(id=s["id"], t=t, material=s["material"], point=point, eye_dir=vneg(norm(ray[1])),
                normal=vec3_mul(s["transform"], s["b"]), entering=bool(t >= ZERO), tex=(dot(s["c"], point), dot(s["d"], point)))


# ---- src/scene/mod.rs --------------------------------------------------------------------------------

def intersect_shape(scene, s, ray):
    if s["kind"] == 0:
        return intersect_sphere(s, ray)
    if s["kind"] == 1:
        return intersect_plane(s, ray)
    if s["kind"] == 2:
        return intersect_triangle(s["a"], s["b"], s["c"], s["d"], ray, s["id"], s["material"])
    return intersect_cube(scene, s, ray)


def scene_intersect(scene, ray, counters=None):  # Scene::intersect, mod.rs:98-116: first shape wins ties, `<` only
    if counters is not None:
        counters["rays"] += 1
    nearest = None
    for s in scene.shapes:
        i = intersect_shape(scene, s, ray)
        if i is None:
            continue
        if nearest is None or i["t"] < nearest["t"]:
            nearest = i
    return nearest


def light_energy(scene, light, p, counters):  # PointLight / AmbientLight::get_energy, mod.rs:189-206, 236-238
    if light["kind"] == 1:
        return (ZERO, ZERO, ZERO), light["color"]
    d = norm(vsub(light["pos"], p))
    i = scene_intersect(scene, (p, d), counters)
    if i is not None and len2(vsub(i["point"], p)) < len2(vsub(light["pos"], p)):
        return d, (ZERO, ZERO, ZERO)
    return d, light["color"]


# ---- src/scene/material.rs, color.rs, my_scene.rs ------------------------------------------------------

def cmul(a, b):
    return (a[0] * b[0], a[1] * b[1], a[2] * b[2])


def cscale(s, c):  # f32 * Color
    return (s * c[0], s * c[1], s * c[2])


def cadd(a, b):
    return (a[0] + b[0], a[1] + b[1], a[2] + b[2])


def _as_i32(x):  # `as i32`: truncation towards zero, saturating, NaN -> 0
    if x != x:
        return 0
    return int(max(min(float(np.trunc(x)), 2147483647.0), -2147483648.0))


def checkerboard(base, tx):  # my_scene.rs:26-43 (base is WHITE there)
    u, v = _as_i32(abs(tx[0])), _as_i32(abs(tx[1]))
    same_sign = (tx[0] < ZERO and tx[1] < ZERO) or (tx[0] > ZERO and tx[1] > ZERO)
    white = (u % 2 == v % 2) if same_sign else (u % 2 != v % 2)
    return base if white else cscale(HALF, base)


def colorfn(cf, tx):
    kind, rgb = cf
    return checkerboard(rgb, tx) if kind == 1 else rgb


def lambert(light_dir, normal, light, surface):  # material.rs:193-195
    return cmul(cscale(dot(light_dir, normal), light), surface)


def phong(power, eye_dir, light_dir, normal, light, surface):  # material.rs:197-213
    h = norm(vadd(norm(eye_dir), norm(light_dir)))
    m_dot_h = dot(normal, h)
    if m_dot_h < ZERO:
        return (ZERO, ZERO, ZERO)
    return cmul(cscale(powf(m_dot_h, power), light), surface)


def reflected_energy(m, incoming, light_dir, i):  # Phong / TexturePhong::get_reflected_energy, material.rs:77-93, 162-185
    diffuse = lambert(light_dir, i["normal"], incoming, colorfn(m["diffuse"], i["tex"]))
    specular = phong(m["power"], i["eye_dir"], light_dir, i["normal"], incoming, colorfn(m["specular"], i["tex"]))
    return cadd(diffuse, specular)


# ---- src/render.rs -----------------------------------------------------------------------------------

def fresnel_reflection(light_dir, normal, n1, n2):  # render.rs:129-134
    m_dot_r = dot(light_dir, normal)
    r0 = ((n1 - n2) / (n1 + n2)) * ((n1 - n2) / (n1 + n2))
    x = ONE - m_dot_r
    x2 = x * x
    return r0 + (ONE - r0) * (x * (x2 * x2))  # powi(5)


def reflect_ray(ray, i):  # render.rs:105-110
    d = vneg(norm(reflect(ray[1], i["normal"])))
    return (vadd(i["point"], vscale(d, OFFSET)), d)


def refract_ray(ray, i, n1, n2):  # render.rs:112-125
    ratio = n1 / n2
    m_dot_r = -dot(ray[1], i["normal"])
    cos_theta_sqrd = ONE - ratio * ratio * (ONE - m_dot_r * m_dot_r)
    if not (cos_theta_sqrd > ZERO):
        return None
    cos_theta = np.sqrt(cos_theta_sqrd)
    d = vadd(vscale(ray[1], ratio), vscale(i["normal"], ratio * m_dot_r - cos_theta))
    return (vadd(i["point"], vscale(d, OFFSET)), d)


def trace_ray(scene, ray, depth, counters):  # render.rs:40-103
    black = (ZERO, ZERO, ZERO)
    if depth == 0:
        return black
    i = scene_intersect(scene, ray, counters)
    if i is None:
        return black
    counters["nodes"] += 1
    m = scene.materials[i["material"]]
    n1, n2 = (ONE, m["refraction_index"]) if i["entering"] else (m["refraction_index"], ONE)
    ambient = cmul(colorfn(m["ambient"], i["tex"]), scene.ambient)
    p = vadd(i["point"], vscale(i["normal"], OFFSET))  # get_light_energy, render.rs:142-153
    lights = black
    for light in scene.lights:
        ldir, lenergy = light_energy(scene, light, p, counters)
        fres = fresnel_reflection(ldir, i["normal"], n1, n2)
        lights = cadd(lights, cscale(fres, reflected_energy(m, lenergy, ldir, i)))
    reflected = black
    if m["reflectivity"] > EPS:
        rr = reflect_ray(ray, i)
        energy = trace_ray(scene, rr, depth - 1, counters)
        fres = fresnel_reflection(rr[1], i["normal"], n1, n2)
        reflected = cscale(fres, reflected_energy(m, energy, rr[1], i))
    refracted = black
    if m["refraction_index"] > EPS:
        rr = refract_ray(ray, i, n1, n2)
        inner = black
        if rr is not None:
            fres = ONE - fresnel_reflection(rr[1], vneg(i["normal"]), n1, n2)  # fresnel_refraction, render.rs:138-140
            inner = cscale(fres, trace_ray(scene, rr, depth - 1, counters))
        refracted = cmul(colorfn(m["diffuse"], i["tex"]), inner)
    return cadd(cadd(cadd(ambient, lights), reflected), refracted)


# ---- src/render_tree.rs ------------------------------------------------------------------------------

def build_ray_tree(scene, ray, depth, shapes, counters):  # render_tree.rs:166-212
    if depth == 0:
        return None
    i = scene_intersect(scene, ray, counters)
    if i is None:
        return None
    counters["nodes"] += 1
    shapes.add(i["id"])  # a cube contributes the id of its inner triangle (cube.rs:93-101)
    m = scene.materials[i["material"]]
    n1, n2 = (ONE, m["refraction_index"]) if i["entering"] else (m["refraction_index"], ONE)
    p = vadd(i["point"], vscale(i["normal"], OFFSET))
    lights = [light_energy(scene, light, p, counters) for light in scene.lights]
    reflected = build_ray_tree(scene, reflect_ray(ray, i), depth - 1, shapes, counters) if m["reflectivity"] > EPS else None
    refracted = None
    if m["refraction_index"] > EPS:
        rr = refract_ray(ray, i, n1, n2)
        if rr is not None:
            refracted = build_ray_tree(scene, rr, depth - 1, shapes, counters)
    return (i, lights, reflected, refracted)


def render_ray_tree(scene, tree, ambient):  # render_tree.rs:214-255 -> (colour, direction)
    black = (ZERO, ZERO, ZERO)
    if tree is None:
        return black, (ZERO, ZERO, ZERO)
    i, lights, reflected, refracted = tree
    m = scene.materials[i["material"]]
    n1, n2 = (ONE, m["refraction_index"]) if i["entering"] else (m["refraction_index"], ONE)
    lsum = black
    for ldir, lenergy in lights:
        lsum = cadd(lsum, cscale(fresnel_reflection(ldir, i["normal"], n1, n2), reflected_energy(m, lenergy, ldir, i)))
    energy, d = render_ray_tree(scene, reflected, ambient)
    refl = cscale(fresnel_reflection(d, i["normal"], n1, n2), reflected_energy(m, energy, i["eye_dir"], i))  # eye_dir, as written
    energy, d = render_ray_tree(scene, refracted, ambient)
    refr = cscale(ONE - fresnel_reflection(d, vneg(i["normal"]), n1, n2), energy)
    amb = cmul(colorfn(m["ambient"], i["tex"]), ambient)
    return cadd(cadd(cadd(amb, lsum), refl), refr), vneg(i["eye_dir"])


def render_forest(desc, cam, depth):
    """generate_ray_forest + render_forest (render_tree.rs:121-127, 147-164):
    (rgb_f32, rgb_u8, counters, per-pixel sets of shape ids)."""
    scene = SceneData(desc)
    f = np.zeros((cam.y_res, cam.x_res, 3), dtype=np.float32)
    u8 = np.zeros((cam.y_res, cam.x_res, 3), dtype=np.uint8)
    counters = dict(rays=0, nodes=0)
    ids = {}
    with np.errstate(all="ignore"):
        for v in range(cam.y_res):
            for u in range(cam.x_res):
                shapes = set()
                tree = build_ray_tree(scene, camera_ray(cam, u, v), depth, shapes, counters)
                c, _ = render_ray_tree(scene, tree, scene.ambient)
                f[v, u] = c
                u8[v, u] = [as_u8(x) for x in c]
                ids[(u, v)] = shapes
    return f, u8, counters, ids


def camera_ray(cam, u, v):  # Camera::get_ray, render.rs:178-185
    x_delta = (F(cam.x_max) - F(cam.x_min)) / F(cam.x_res)
    y_delta = (F(cam.y_max) - F(cam.y_min)) / F(cam.y_res)
    x = F(cam.x_min) + F(u) * x_delta
    y = F(cam.y_max) - F(v) * y_delta
    origin = tuple(_f(cam.origin))
    return (origin, norm(vsub((x, y, ZERO), origin)))


def as_u8(x):  # Color::as_u8, color.rs:43-46: `(255. * c) as u8` saturates, NaN -> 0
    y = F(255.0) * x
    if y != y:
        return 0
    return int(max(min(float(np.trunc(y)), 255.0), 0.0))


def render(desc, cam, depth):
    """Returns (rgb_f32 [rows, cols, 3], rgb_u8, counters) for the whole frame."""
    scene = SceneData(desc)
    f = np.zeros((cam.y_res, cam.x_res, 3), dtype=np.float32)
    u8 = np.zeros((cam.y_res, cam.x_res, 3), dtype=np.uint8)
    counters = dict(rays=0, nodes=0)
    with np.errstate(all="ignore"):
        for v in range(cam.y_res):
            for u in range(cam.x_res):
                c = trace_ray(scene, camera_ray(cam, u, v), depth, counters)
                f[v, u] = c
                u8[v, u] = [as_u8(x) for x in c]
    return f, u8, counters
