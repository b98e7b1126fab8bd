// exact.cuh -- reference-order f32 arithmetic for the decisions that must match the
// reference bit for bit: every + - * / sqrt is one IEEE
// round-to-nearest operation, written with the __f*_rn intrinsics so that nvcc can never
// contract a*b+c into an FMA, whatever -fmad says.  No rsqrt, no reciprocal tricks.
//
// Cites are into the reference tree (/root/reference/src).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rt {

#define RT_DEV __device__ __forceinline__

constexpr float F32_EPS = 1.1920929e-07f; // std::f32::EPSILON
constexpr float PI_F = 3.14159274101257324f; // std::f32::consts::PI

RT_DEV float xmul(float a, float b) { return __fmul_rn(a, b); }
RT_DEV float xadd(float a, float b) { return __fadd_rn(a, b); }
RT_DEV float xsub(float a, float b) { return __fsub_rn(a, b); }
RT_DEV float xdiv(float a, float b) { return __fdiv_rn(a, b); }
RT_DEV float xsqrt(float a) { return __fsqrt_rn(a); }

struct V3 {
    float x, y, z;
};
RT_DEV V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_DEV V3 vneg(V3 a) { return v3(-a.x, -a.y, -a.z); }
RT_DEV V3 vadd(V3 a, V3 b) { return v3(xadd(a.x, b.x), xadd(a.y, b.y), xadd(a.z, b.z)); }
RT_DEV V3 vsub(V3 a, V3 b) { return v3(xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)); }
RT_DEV V3 vscale(V3 a, float s) { return v3(xmul(a.x, s), xmul(a.y, s), xmul(a.z, s)); } // vector3.rs:39-45
RT_DEV V3 vdivs(V3 a, float d) { return v3(xdiv(a.x, d), xdiv(a.y, d), xdiv(a.z, d)); }  // vector3.rs:48-54
RT_DEV float vdot(V3 a, V3 b) {                                                          // vector3.rs:86-88
    return xadd(xadd(xmul(a.x, b.x), xmul(a.y, b.y)), xmul(a.z, b.z));
}
RT_DEV float vlen2(V3 a) { return vdot(a, a); }           // vector3.rs:75-77
RT_DEV V3 vnorm(V3 a) { return vdivs(a, xsqrt(vlen2(a))); } // vector3.rs:80-94
RT_DEV V3 vcross(V3 a, V3 b) {                            // vector3.rs:97-103
    return v3(xsub(xmul(a.y, b.z), xmul(a.z, b.y)), xsub(xmul(a.z, b.x), xmul(a.x, b.z)),
              xsub(xmul(a.x, b.y), xmul(a.y, b.x)));
}
// ray.rs:50-52  origin + direction * a
RT_DEV V3 ray_at(V3 o, V3 d, float t) { return vadd(o, vscale(d, t)); }

// A 3x4 affine matrix = rows 0..2 of the reference's row-major 4x4 (matrix.rs:5-9).
struct M34 {
    float4 r0, r1, r2;
};
// matrix.rs:237-252  ((p.x*m0 + p.y*m1) + p.z*m2) + m3
RT_DEV V3 m_pt(const M34 &m, V3 p) {
    return v3(xadd(xadd(xadd(xmul(p.x, m.r0.x), xmul(p.y, m.r0.y)), xmul(p.z, m.r0.z)), m.r0.w),
              xadd(xadd(xadd(xmul(p.x, m.r1.x), xmul(p.y, m.r1.y)), xmul(p.z, m.r1.z)), m.r1.w),
              xadd(xadd(xadd(xmul(p.x, m.r2.x), xmul(p.y, m.r2.y)), xmul(p.z, m.r2.z)), m.r2.w));
}
// matrix.rs:229-235
RT_DEV V3 m_vec(const M34 &m, V3 v) {
    return v3(xadd(xadd(xmul(v.x, m.r0.x), xmul(v.y, m.r0.y)), xmul(v.z, m.r0.z)),
              xadd(xadd(xmul(v.x, m.r1.x), xmul(v.y, m.r1.y)), xmul(v.z, m.r1.z)),
              xadd(xadd(xmul(v.x, m.r2.x), xmul(v.y, m.r2.y)), xmul(v.z, m.r2.z)));
}
// transpose().vec3_mul(v) (matrix.rs:77-86,229-235): row r of the transpose = column r
RT_DEV V3 m_vec_transposed(const M34 &m, V3 v) {
    return v3(xadd(xadd(xmul(v.x, m.r0.x), xmul(v.y, m.r1.x)), xmul(v.z, m.r2.x)),
              xadd(xadd(xmul(v.x, m.r0.y), xmul(v.y, m.r1.y)), xmul(v.z, m.r2.y)),
              xadd(xadd(xmul(v.x, m.r0.z), xmul(v.y, m.r1.z)), xmul(v.z, m.r2.z)));
}

// ---- hit keys -------------------------------------------------------------------------
// Scene::intersect keeps the smallest t and replaces only on strict `<` (scene/mod.rs:
// 98-116), so the winner is the lexicographic minimum of (t, shape id).  A 64-bit key
// (orderable(t) << 32 | id) turns that into one integer min.  -0.0 is folded onto +0.0
// first (they compare equal in the reference).
constexpr unsigned long long KEY_MISS = 0xFFFFFFFFFFFFFFFFull;

RT_DEV uint32_t f32_orderable(float t) {
    uint32_t b = __float_as_uint(t + 0.0f); // -0.0 + 0.0 = +0.0
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
RT_DEV float f32_from_orderable(uint32_t u) {
    uint32_t b = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
    return __uint_as_float(b);
}
RT_DEV unsigned long long make_key(float t, int shape) {
    return ((unsigned long long)f32_orderable(t) << 32) | (unsigned long long)(uint32_t)shape;
}
RT_DEV float key_t(unsigned long long k) { return f32_from_orderable((uint32_t)(k >> 32)); }
RT_DEV int key_shape(unsigned long long k) { return (int)(uint32_t)(k & 0xFFFFFFFFull); }
// A ray with a NaN component "hits" the first sphere / triangle / cube with t = NaN and
// nothing can replace it (every comparison with NaN is false: sphere.rs:61-76,126-145,
// triangle.rs:56-80, scene/mod.rs:105-111).  Such hits carry this key prefix.
constexpr uint32_t KEY_NAN_PREFIX = 0xFFFFFFFEu;

// ---- exact t of one shape (the decision part of each intersect()) --------------------

// sphere.rs:57-76 + solve_quadratic :126-145.  Returns hit; t and t0 (for `entering`).
RT_DEV bool exact_sphere(const M34 &inv, V3 o, V3 d, float &t, bool &entering) {
    V3 to = m_pt(inv, o);
    V3 td = m_vec(inv, d);
    float a = vlen2(td);
    float b = xmul(2.f, vdot(td, to));
    float c = xsub(vlen2(to), 1.f);
    float discr = xsub(xmul(b, b), xmul(xmul(4.f, a), c));
    float t0, t1;
    if (discr < 0.f) return false;
    if (fabsf(discr) < F32_EPS) {
        float x = xdiv(xmul(-0.5f, b), a);
        t0 = x; t1 = x;
    } else {
        float sq = xsqrt(discr);
        float q = (b > 0.f) ? xmul(-0.5f, xadd(b, sq)) : xmul(-0.5f, xsub(b, sq));
        t0 = xdiv(q, a);
        t1 = xdiv(c, q);
    }
    if (t0 > t1) { float tmp = t1; t1 = t0; t0 = tmp; }
    if (t0 < 0.f && t1 < 0.f) return false;
    t = (t0 < 0.f) ? t1 : t0;
    entering = t0 > 0.f;
    return true;
}

struct Tri {
    V3 v0, v1, v2, n;
};
// triangle.rs:51-80.  u, v are the texture coordinates; det decides `entering`.
RT_DEV bool exact_triangle(const Tri &tr, V3 o, V3 d, float &t, float &u, float &v, float &det) {
    V3 v0v1 = vsub(tr.v1, tr.v0);
    V3 v0v2 = vsub(tr.v2, tr.v0);
    V3 pvec = vcross(d, v0v2);
    det = vdot(v0v1, pvec);
    if (fabsf(det) < F32_EPS) return false;
    float inv_det = xdiv(1.0f, det);
    V3 tvec = vsub(o, tr.v0);
    u = xmul(vdot(tvec, pvec), inv_det);
    if (u < 0.f || u > 1.f) return false;
    V3 qvec = vcross(tvec, v0v1);
    v = xmul(vdot(d, qvec), inv_det);
    if (v < 0.f || xadd(u, v) > 1.f) return false;
    t = xmul(vdot(v0v2, qvec), inv_det);
    if (t < 0.f) return false;
    return true;
}

RT_DEV Tri load_tri(const float4 *g) {
    float4 a = g[0], b = g[1], c = g[2];
    Tri t;
    t.v0 = v3(a.x, a.y, a.z);
    t.v1 = v3(a.w, b.x, b.y);
    t.v2 = v3(b.z, b.w, c.x);
    t.n = v3(c.y, c.z, c.w);
    return t;
}
RT_DEV M34 load_m34(const float4 *g) {
    M34 m;
    m.r0 = g[0]; m.r1 = g[1]; m.r2 = g[2];
    return m;
}

// cube.rs:89-102 + nested Scene::intersect over the 12 triangles (scene/mod.rs:98-116):
// first hit, then strict `<`.  `which` = inner triangle index (the id the reference keeps).
RT_DEV bool exact_cube(const M34 &inv, const float4 *cube_tris, V3 o, V3 d, float &t, int &which, float &u,
                       float &v, float &det) {
    V3 to = m_pt(inv, o);
    V3 td = m_vec(inv, d);
    bool have = false;
    for (int k = 0; k < 12; ++k) {
        Tri tr = load_tri(cube_tris + 3 * k);
        float tk, uk, vk, dk;
        if (exact_triangle(tr, to, td, tk, uk, vk, dk)) {
            if (!have || tk < t) { t = tk; which = k; u = uk; v = vk; det = dk; have = true; }
        }
    }
    return have;
}

// plane.rs:59-64.  p[0..3] = origin, normal, u axis, v axis.
RT_DEV bool exact_plane(const M34 &inv, V3 origin, V3 normal, V3 o, V3 d, float &t) {
    V3 to = m_pt(inv, o);
    V3 td = m_vec(inv, d);
    float denom = -vdot(normal, td);
    if (!(denom > F32_EPS)) return false;
    V3 dir = vsub(origin, to);
    t = xdiv(-vdot(dir, normal), denom);
    return true;
}

} // namespace rt
