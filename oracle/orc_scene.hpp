// ORACLE -- TEST INFRASTRUCTURE ONLY (see orc_math.hpp).
// Restates src/scene/{mod,sphere,plane,triangle,cube,intersection,material}.rs.
#pragma once
#include <vector>

#include "../include/rt_b200.h"
#include "orc_math.hpp"

namespace orc {

struct Counters {
    uint64_t rays = 0, shadow_rays = 0, nodes = 0;
    uint64_t sphere_tests = 0, plane_tests = 0, triangle_tests = 0, cube_tests = 0;
    void add(const Counters &o) {
        rays += o.rays; shadow_rays += o.shadow_rays; nodes += o.nodes;
        sphere_tests += o.sphere_tests; plane_tests += o.plane_tests;
        triangle_tests += o.triangle_tests; cube_tests += o.cube_tests;
    }
};

// src/scene/intersection.rs:8-17 (material is an index instead of Rc<RefCell<dyn Material>>)
struct Intersection {
    int32_t id;
    float t;
    int32_t material;
    Point3 point;
    Vector3 eye_dir;
    Vector3 normal;
    bool entering;
    float tex_u, tex_v;
};

// ---- materials: src/scene/material.rs -------------------------------------------------

// src/my_scene.rs:26-43 generalised to a base colour (WHITE there): base or 0.5*base.
// Rust `f32 as i32`: truncate toward zero, saturate, NaN -> 0.
inline int32_t f32_as_i32(float x) {
    if (!(x == x)) return 0;
    if (x >= 2147483648.0f) return INT32_MAX;
    if (x <= -2147483648.0f) return INT32_MIN;
    return (int32_t)x;
}

inline Color checkerboard(const rt_colorfn &f, float tx0, float tx1) {
    int32_t u = f32_as_i32(std::fabs(tx0));
    int32_t v = f32_as_i32(std::fabs(tx1));
    Color base(f.rgb[0], f.rgb[1], f.rgb[2]);
    bool same_sign = (tx0 < 0.f && tx1 < 0.f) || (tx0 > 0.f && tx1 > 0.f);
    if (same_sign) {
        if (u % 2 == v % 2) return base;
        return 0.5f * base;
    } else {
        if (u % 2 != v % 2) return base;
        return 0.5f * base;
    }
}

inline Color eval_colorfn(const rt_colorfn &f, float tx0, float tx1) {
    if (f.kind == RT_COLOR_CHECKERBOARD) return checkerboard(f, tx0, tx1);
    return Color(f.rgb[0], f.rgb[1], f.rgb[2]);
}

// material.rs:193-195  `light_dir.dot(normal) * light * surface`  (unclamped)
inline Color lambert(const Vector3 &light_dir, const Vector3 &normal, const Color &light, const Color &surface) {
    return (light_dir.dot(normal) * light) * surface;
}

// material.rs:197-213
inline Color phong(float power, const Vector3 &eye_dir, const Vector3 &light_dir, const Vector3 &normal,
                   const Color &light, const Color &surface) {
    Vector3 h = (eye_dir.norm() + light_dir.norm()).norm();
    float m_dot_h = normal.dot(h);
    if (m_dot_h < 0.f) return BLACK;
    return (std::pow(m_dot_h, power) * light) * surface; // powf
}

// Phong / TexturePhong::get_reflected_energy (material.rs:77-93,165-186)
inline Color get_reflected_energy(const rt_material &m, const Color &incoming, const Vector3 &light_dir,
                                  const Intersection &i) {
    Color kd = eval_colorfn(m.diffuse, i.tex_u, i.tex_v);
    Color ks = eval_colorfn(m.specular, i.tex_u, i.tex_v);
    Color diffuse = lambert(light_dir, i.normal, incoming, kd);
    Color specular = phong(m.power, i.eye_dir, light_dir, i.normal, incoming, ks);
    return diffuse + specular;
}

// ---- shapes ---------------------------------------------------------------------------

// src/scene/triangle.rs:7-94
struct Triangle {
    int32_t id = 0;
    int32_t material = 0;
    Point3 v0, v1, v2;
    Vector3 normal;

    // Triangle::new (:17-39): normal = ((v1 - v0) x (v2 - v1)).norm()
    static Triangle make(const Point3 &a, const Point3 &b, const Point3 &c, int32_t material) {
        Triangle t;
        t.v0 = a; t.v1 = b; t.v2 = c;
        t.material = material;
        Vector3 e01 = b - a;
        Vector3 e12 = c - b;
        t.normal = e01.cross(e12).norm();
        return t;
    }

    // :51-94
    bool intersect(const Ray &ray, Intersection &out) const {
        Vector3 v0v1 = v1 - v0;
        Vector3 v0v2 = v2 - v0;
        Vector3 pvec = ray.direction.cross(v0v2);
        float det = v0v1.dot(pvec);
        if (std::fabs(det) < EPS) return false;
        float inv_det = 1.0f / det;
        Vector3 tvec = ray.origin - v0;
        float u = tvec.dot(pvec) * inv_det;
        if (u < 0.f || u > 1.f) return false;
        Vector3 qvec = tvec.cross(v0v1);
        float v = ray.direction.dot(qvec) * inv_det;
        if (v < 0.f || u + v > 1.f) return false;
        float t = v0v2.dot(qvec) * inv_det;
        if (t < 0.f) return false;
        out.id = id;
        out.t = t;
        out.material = material;
        out.point = ray.at(t);
        out.eye_dir = -(ray.direction.norm());
        out.normal = normal; // `if t < 0. {-n} else {n}` is dead code after the return above
        out.entering = det > 0.f;
        out.tex_u = u;
        out.tex_v = v;
        return true;
    }
};

// Scene::intersect over a triangle list (src/scene/mod.rs:98-116): first hit, then strict <.
inline bool nearest_triangle(const std::vector<Triangle> &tris, const Ray &ray, Intersection &best) {
    bool have = false;
    float nearest = 0.f;
    Intersection cur;
    for (const Triangle &t : tris) {
        if (t.intersect(ray, cur)) {
            if (!have) { nearest = cur.t; best = cur; have = true; }
            else if (cur.t < nearest) { nearest = cur.t; best = cur; }
        }
    }
    return have;
}

struct Shape {
    int32_t kind = RT_SHAPE_SPHERE;
    int32_t id = 0;
    int32_t material = 0;
    Matrix transform = Matrix::identity();
    Matrix inv_transform = Matrix::identity();
    // plane
    Point3 origin;
    Vector3 normal, u_axis, v_axis;
    // triangle
    Triangle tri;
    // cube: nested scene of 12 triangles with ids 0..11 (cube.rs:21-77)
    std::vector<Triangle> cube_tris;
};

// Plane::new (src/scene/plane.rs:22-42): texture axes
inline void plane_axes(const Vector3 &normal, Vector3 &u, Vector3 &v) {
    Vector3 w = (normal.cross(Vector3(1.f, 0.f, 0.f)).len() <= EPS) ? Vector3(0.f, 1.f, 0.f)
                                                                      : Vector3(1.f, 0.f, 0.f);
    u = normal.cross(w).norm();
    v = normal.cross(u).norm();
}

// Cube::new (src/scene/cube.rs:21-77): the 12 triangles in add order
inline std::vector<Triangle> cube_triangles(int32_t material) {
    Point3 v0(0.5f, 0.5f, -0.5f), v1(0.5f, -0.5f, -0.5f), v2(-0.5f, -0.5f, -0.5f), v3(-0.5f, 0.5f, -0.5f);
    Point3 v4(0.5f, 0.5f, 0.5f), v5(-0.5f, 0.5f, 0.5f), v6(-0.5f, -0.5f, 0.5f), v7(0.5f, -0.5f, 0.5f);
    std::vector<Triangle> t;
    t.push_back(Triangle::make(v1, v2, v3, material)); // tf1
    t.push_back(Triangle::make(v0, v1, v3, material)); // tf2
    t.push_back(Triangle::make(v7, v5, v4, material)); // tk1
    t.push_back(Triangle::make(v5, v7, v6, material)); // tk2
    t.push_back(Triangle::make(v0, v4, v7, material)); // tr1
    t.push_back(Triangle::make(v7, v1, v0, material)); // tr2
    t.push_back(Triangle::make(v5, v3, v6, material)); // tl1
    t.push_back(Triangle::make(v6, v3, v2, material)); // tl2
    t.push_back(Triangle::make(v5, v4, v0, material)); // tt1
    t.push_back(Triangle::make(v0, v3, v5, material)); // tt2
    t.push_back(Triangle::make(v1, v7, v6, material)); // tb1
    t.push_back(Triangle::make(v6, v2, v1, material)); // tb2
    for (size_t k = 0; k < t.size(); ++k) t[k].id = (int32_t)k; // add_shape: id = index
    return t;
}

// src/scene/sphere.rs:126-145
inline bool solve_quadratic(float a, float b, float c, float &x0, float &x1) {
    float discr = b * b - 4.f * a * c;
    if (discr < 0.f) return false;
    if (std::fabs(discr) < EPS) {
        float x = -0.5f * b / a;
        x0 = x; x1 = x;
        return true;
    }
    float q = (b > 0.f) ? -0.5f * (b + std::sqrt(discr)) : -0.5f * (b - std::sqrt(discr));
    x0 = q / a;
    x1 = c / q;
    return true;
}

// src/scene/sphere.rs:40-45
inline void sphere_tex(const Vector3 &n, float &u, float &v) {
    u = (1.f + std::atan2(n.z, n.x) / PI_F) * 0.5f;
    v = std::acos(n.y) / PI_F;
}

// src/scene/sphere.rs:57-98
inline bool sphere_intersect(const Shape &s, const Ray &ray, Intersection &out) {
    Ray tr = s.inv_transform.ray_mul(ray);
    Vector3 l = tr.origin - Point3(0.f, 0.f, 0.f);
    float a = tr.direction.len2();
    float b = 2.f * tr.direction.dot(l);
    float c = l.len2() - 1.f;
    float t0, t1;
    if (!solve_quadratic(a, b, c, t0, t1)) return false;
    if (t0 > t1) { float tmp = t1; t1 = t0; t0 = tmp; }
    if (t0 < 0.f && t1 < 0.f) return false;
    float t = (t0 < 0.f) ? t1 : t0;
    bool entering = t0 > 0.f;
    Point3 point = ray.at(t);
    Point3 n_obj = tr.at(t);
    Vector3 normal = s.inv_transform.transpose().vec3_mul(to_vec(n_obj)).norm();
    if (!entering) normal = -normal;
    out.id = s.id;
    out.t = t;
    out.material = s.material;
    out.point = point;
    out.eye_dir = -(ray.direction.norm());
    out.normal = normal;
    out.entering = entering;
    sphere_tex(normal, out.tex_u, out.tex_v);
    return true;
}

// src/scene/plane.rs:59-84 (negative t IS returned as a hit)
inline bool plane_intersect(const Shape &s, const Ray &ray, Intersection &out) {
    Ray tr = s.inv_transform.ray_mul(ray);
    float denom = -(s.normal.dot(tr.direction));
    if (!(denom > EPS)) return false;
    Vector3 dir = s.origin - tr.origin;
    float t = -(dir.dot(s.normal)) / denom;
    Point3 point = ray.at(t);
    out.id = s.id;
    out.t = t;
    out.material = s.material;
    out.entering = t >= 0.f;
    out.point = point;
    out.eye_dir = -(ray.direction.norm());
    out.normal = s.transform.vec3_mul(s.normal); // forward matrix, not re-normalised
    out.tex_u = s.u_axis.dot(to_vec(point));
    out.tex_v = s.v_axis.dot(to_vec(point));
    return true;
}

// src/scene/cube.rs:89-102 (id stays the inner triangle's 0..11)
inline bool cube_intersect(const Shape &s, const Ray &ray, Intersection &out) {
    Ray tr = s.inv_transform.ray_mul(ray);
    if (!nearest_triangle(s.cube_tris, tr, out)) return false;
    out.point = ray.at(out.t);
    out.eye_dir = -(ray.direction.norm());
    out.normal = s.inv_transform.transpose().vec3_mul(out.normal).norm();
    return true;
}

struct Light {
    int32_t kind;
    Point3 pos;
    Color color;
};

struct Scene {
    std::vector<Shape> shapes;
    std::vector<rt_material> materials;
    std::vector<Light> lights;
    Color ambient;
    int n_spheres = 0, n_planes = 0, n_triangles = 0, n_cubes = 0;

    bool shape_intersect(const Shape &s, const Ray &ray, Intersection &out) const {
        switch (s.kind) {
        case RT_SHAPE_SPHERE: return sphere_intersect(s, ray, out);
        case RT_SHAPE_PLANE: return plane_intersect(s, ray, out);
        case RT_SHAPE_TRIANGLE: {
            bool h = s.tri.intersect(ray, out);
            return h;
        }
        default: return cube_intersect(s, ray, out);
        }
    }

    // Scene::intersect (src/scene/mod.rs:98-116)
    bool intersect(const Ray &ray, Intersection &best, Counters &cnt) const {
        cnt.rays += 1;
        cnt.sphere_tests += n_spheres;
        cnt.plane_tests += n_planes;
        cnt.triangle_tests += n_triangles;
        cnt.cube_tests += n_cubes;
        bool have = false;
        float nearest = 0.f;
        Intersection cur;
        for (const Shape &s : shapes) {
            if (shape_intersect(s, ray, cur)) {
                if (!have) { nearest = cur.t; best = cur; have = true; }
                else if (cur.t < nearest) { nearest = cur.t; best = cur; }
            }
        }
        return have;
    }
};

// Build the oracle's scene from the flat ABI description.
inline bool scene_from_desc(const rt_scene_desc &d, Scene &sc) {
    sc = Scene();
    sc.ambient = Color(d.ambient[0], d.ambient[1], d.ambient[2]);
    sc.materials.assign(d.materials, d.materials + d.n_materials);
    for (int i = 0; i < d.n_lights; ++i) {
        Light l;
        l.kind = d.lights[i].kind;
        l.pos = Point3(d.lights[i].pos[0], d.lights[i].pos[1], d.lights[i].pos[2]);
        l.color = Color(d.lights[i].color[0], d.lights[i].color[1], d.lights[i].color[2]);
        sc.lights.push_back(l);
    }
    for (int i = 0; i < d.n_shapes; ++i) {
        const rt_shape &in = d.shapes[i];
        if (in.material < 0 || in.material >= d.n_materials) return false;
        Shape s;
        s.kind = in.kind;
        s.id = i; // add_shape (src/scene/mod.rs:40-44)
        s.material = in.material;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) {
                s.transform.mat[r][c] = in.transform[r * 4 + c];
                s.inv_transform.mat[r][c] = in.inv_transform[r * 4 + c];
            }
        const float *p = in.p;
        switch (in.kind) {
        case RT_SHAPE_SPHERE: sc.n_spheres++; break;
        case RT_SHAPE_PLANE:
            s.origin = Point3(p[0], p[1], p[2]);
            s.normal = Vector3(p[3], p[4], p[5]);
            s.u_axis = Vector3(p[6], p[7], p[8]);
            s.v_axis = Vector3(p[9], p[10], p[11]);
            sc.n_planes++;
            break;
        case RT_SHAPE_TRIANGLE:
            s.tri.v0 = Point3(p[0], p[1], p[2]);
            s.tri.v1 = Point3(p[3], p[4], p[5]);
            s.tri.v2 = Point3(p[6], p[7], p[8]);
            s.tri.normal = Vector3(p[9], p[10], p[11]);
            s.tri.id = i;
            s.tri.material = in.material;
            sc.n_triangles++;
            break;
        case RT_SHAPE_CUBE:
            s.cube_tris = cube_triangles(in.material);
            sc.n_cubes++;
            break;
        default: return false;
        }
        sc.shapes.push_back(s);
    }
    return true;
}

} // namespace orc
