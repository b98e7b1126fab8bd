// shade.cuh -- full hit record of the winning shape, materials, Fresnel, secondary rays.
// Everything here runs once per ray-tree node (not in the all-pairs sweep) and keeps the
// reference's operand order; only powf / atan2f / acosf come from the CUDA math library
// (the reference uses glibc's: a few ulp apart, smooth, far below 1/255).
#pragma once
#include "scene_dev.cuh"

namespace rt {

// Intersection (reference src/scene/intersection.rs:8-17)
struct Hit {
    int id;       // shape id; for a cube the inner triangle 0..11 (cube.rs:93-101 never overwrites it)
    int shape;    // top-level shape id
    int kind;     // KIND_* of the top-level shape
    int material;
    float t;
    V3 point, eye, normal;
    bool entering;
    float tu, tv;
};

// sphere.rs:40-45
RT_DEV void sphere_tex(V3 n, float &u, float &v) {
    u = xmul(xadd(1.f, xdiv(atan2f(n.z, n.x), PI_F)), 0.5f);
    v = xdiv(acosf(n.y), PI_F);
}

// Rebuild the Intersection the winning shape's intersect() returns.  The key only names
// the winner; t and everything else are recomputed with the exact chain.  A sphere's texture coordinates
// (sphere.rs:40-45: atan2f + acosf, ~250 instructions) are left to hit_tex(): nothing but a textured
// material reads them, and the spheres of the big scenes have none.
RT_DEV bool hit_record(const DevScene &sc, unsigned long long key, V3 o, V3 d, Hit &h) {
    if (key == KEY_MISS) return false;
    int shape = key_shape(key);
    int4 meta = sc.shape_meta[shape];
    const float4 *g = sc.geom + 3 * shape;
    h.shape = shape;
    h.kind = meta.x;
    h.id = shape;
    h.material = meta.y;
    h.eye = vneg(vnorm(d));
    float t = 0.f;
    if (meta.x == KIND_SPHERE) { // sphere.rs:57-98
        M34 inv = load_m34(g);
        bool entering = false;
        exact_sphere(inv, o, d, t, entering);
        V3 to = m_pt(inv, o), td = m_vec(inv, d);
        V3 n_obj = ray_at(to, td, t);
        V3 n = vnorm(m_vec_transposed(inv, n_obj));
        if (!entering) n = vneg(n);
        h.normal = n;
        h.entering = entering;
        h.tu = 0.f; h.tv = 0.f; // hit_tex()
    } else if (meta.x == KIND_PLANE) { // plane.rs:59-84
        const float4 *pd = sc.plane_data + 7 * meta.z;
        V3 origin = v3(pd[0].x, pd[0].y, pd[0].z), normal = v3(pd[1].x, pd[1].y, pd[1].z);
        exact_plane(load_m34(g), origin, normal, o, d, t);
        h.entering = t >= 0.f;
        M34 fwd;
        fwd.r0 = pd[4]; fwd.r1 = pd[5]; fwd.r2 = pd[6];
        h.normal = m_vec(fwd, normal);
        V3 p = ray_at(o, d, t);
        h.tu = vdot(v3(pd[2].x, pd[2].y, pd[2].z), p);
        h.tv = vdot(v3(pd[3].x, pd[3].y, pd[3].z), p);
    } else if (meta.x == KIND_TRIANGLE) { // triangle.rs:51-94
        Tri tr = load_tri(g);
        float u = 0.f, v = 0.f, det = 0.f;
        exact_triangle(tr, o, d, t, u, v, det);
        h.normal = tr.n;
        h.entering = det > 0.f;
        h.tu = u; h.tv = v;
    } else { // cube.rs:89-102
        M34 inv = load_m34(g);
        float u = 0.f, v = 0.f, det = 0.f;
        int which = 0;
        exact_cube(inv, sc.cube_tris, o, d, t, which, u, v, det);
        Tri tr = load_tri(sc.cube_tris + 3 * which);
        h.id = which;
        h.normal = vnorm(m_vec_transposed(inv, tr.n));
        h.entering = det > 0.f;
        h.tu = u; h.tv = v;
    }
    h.t = t;
    h.point = ray_at(o, d, t);
    return true;
}

// The part of the Intersection hit_record leaves out: call before evaluating a colour function that reads
// the texture coordinate (m.ambient / diffuse / specular of kind CHECKERBOARD).
RT_DEV void hit_tex(Hit &h) {
    if (h.kind == KIND_SPHERE) sphere_tex(h.normal, h.tu, h.tv);
}
RT_DEV bool material_reads_tex(const DevMaterial &m) { return (m.ambient.kind | m.diffuse.kind | m.specular.kind) != 0; }

// ---- materials (material.rs) ------------------------------------------------------------

// my_scene.rs:26-43 on a base colour.  `as i32` saturates and maps NaN to 0, which is
// exactly cvt.rzi.s32.f32.
RT_DEV V3 eval_colorfn(const DevColorFn &f, float tu, float tv) {
    V3 base = v3(f.r, f.g, f.b);
    if (f.kind == 0) return base;
    int u = __float2int_rz(fabsf(tu));
    int v = __float2int_rz(fabsf(tv));
    bool same_sign = (tu < 0.f && tv < 0.f) || (tu > 0.f && tv > 0.f);
    bool same_parity = (u % 2) == (v % 2);
    bool full = same_sign ? same_parity : !same_parity;
    return full ? base : v3(xmul(0.5f, base.x), xmul(0.5f, base.y), xmul(0.5f, base.z));
}

RT_DEV V3 cmul(V3 a, V3 b) { return v3(xmul(a.x, b.x), xmul(a.y, b.y), xmul(a.z, b.z)); }
RT_DEV V3 cscale(float s, V3 c) { return v3(xmul(s, c.x), xmul(s, c.y), xmul(s, c.z)); }

// Phong / TexturePhong::get_reflected_energy (material.rs:77-93,165-186,193-213)
RT_DEV V3 reflected_energy(const DevMaterial &m, V3 incoming, V3 light_dir, const Hit &h) {
    V3 kd = eval_colorfn(m.diffuse, h.tu, h.tv);
    V3 ks = eval_colorfn(m.specular, h.tu, h.tv);
    V3 diffuse = cmul(cscale(vdot(light_dir, h.normal), incoming), kd); // lambert, unclamped
    V3 hv = vnorm(vadd(vnorm(h.eye), vnorm(light_dir)));
    float m_dot_h = vdot(h.normal, hv);
    V3 specular = v3(0.f, 0.f, 0.f);
    if (!(m_dot_h < 0.f)) specular = cmul(cscale(powf(m_dot_h, m.power), incoming), ks);
    return vadd(diffuse, specular);
}

// render.rs:129-134; powi(5) = x * ((x*x)*(x*x))
RT_DEV float fresnel_reflection(V3 dir, V3 n, float n1, float n2) {
    float m_dot_r = vdot(dir, n);
    float q = xdiv(xsub(n1, n2), xadd(n1, n2));
    float r0 = xmul(q, q);
    float x = xsub(1.f, m_dot_r);
    float x2 = xmul(x, x);
    float p5 = xmul(x, xmul(x2, x2));
    return xadd(r0, xmul(xsub(1.f, r0), p5));
}

// render.rs:105-110
RT_DEV void reflect_ray(V3 d, const Hit &h, V3 &ro, V3 &rd) {
    float s = xmul(2.f, vdot(d, h.normal)); // vector3.rs:113-115
    V3 refl = vsub(vscale(h.normal, s), d);
    rd = vneg(vnorm(refl));
    ro = vadd(h.point, vscale(rd, 0.0002f));
}

// render.rs:112-125
RT_DEV bool refract_ray(V3 d, const Hit &h, float n1, float n2, V3 &ro, V3 &rd) {
    float ratio = xdiv(n1, n2);
    float m_dot_r = -vdot(d, h.normal);
    float cos2 = xsub(1.f, xmul(xmul(ratio, ratio), xsub(1.f, xmul(m_dot_r, m_dot_r))));
    if (!(cos2 > 0.f)) return false;
    float cos_theta = xsqrt(cos2);
    rd = vadd(vscale(d, ratio), vscale(h.normal, xsub(xmul(ratio, m_dot_r), cos_theta)));
    ro = vadd(h.point, vscale(rd, 0.0002f));
    return true;
}

RT_DEV void refraction_indices(const DevMaterial &m, bool entering, float &n1, float &n2) { // render.rs:51-55
    if (entering) { n1 = 1.f; n2 = m.refraction_index; }
    else { n1 = m.refraction_index; n2 = 1.f; }
}

// Color::as_u8 (color.rs:43-46): (255*c) as u8 -> truncate, saturate, NaN -> 0
RT_DEV unsigned char as_u8(float c) {
    unsigned int v = __float2uint_rz(xmul(255.f, c)); // saturating, NaN -> 0
    return (unsigned char)(v > 255u ? 255u : v);
}

} // namespace rt
