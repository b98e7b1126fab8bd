// builtin_scenes.cpp -- the reference's compiled-in scene (src/my_scene.rs) and the two
// synthetic benchmark scenes, built through the host scene API and flattened for the ABI.
#include <cstring>
#include <new>
#include <string>

#include "../../../include/rt_b200_host.h"
#include "rt_scene.hpp"

using namespace rust_tracer;
using namespace rust_tracer::colors;

namespace {

// src/my_scene.rs:45-120
void create_scene(Scene &scene) {
    const Color DIM_WHITE(0.1f, 0.1f, 0.1f), DIM_BLUE(0.f, 0.f, 0.1f);
    const ColorFun dim_white = ColorFun::constant(0.1f * WHITE); // fn dim_white (:11-13)
    const ColorFun checkerboard = ColorFun::checkerboard(WHITE);  // fn checkerboard (:26-43)

    auto phong = std::make_shared<Phong>(DIM_WHITE, RED, WHITE, 60.f, 0.5f, 0.f);
    Sphere sph(phong);
    sph.set_transform(Matrix::translate(-1.0f, 0.f, 0.f) * Matrix::rotate_z(75.f) * Matrix::scale(1.0f, 0.25f, 1.0f));
    scene.add_shape(std::move(sph));

    phong = std::make_shared<Phong>(BLACK, BLUE, DIM_BLUE, 600.f, 0.4f, 0.f);
    Sphere sph2 = Sphere::new_with_name("blue", phong);
    sph2.set_transform(Matrix::translate(1.f, -1.f, 0.f));
    scene.add_shape(std::move(sph2));

    phong = std::make_shared<Phong>(BLACK, WHITE, WHITE, 60.f, 0.7f, 1.333f);
    Sphere sph4(phong);
    sph4.set_transform(Matrix::translate(0.f, -0.5f, -3.f) * Matrix::scale(0.6f, 0.6f, 0.6f));
    scene.add_shape(std::move(sph4));

    auto plane_material = std::make_shared<TexturePhong>(dim_white, checkerboard, dim_white, 60.f, 0.f, 0.f);
    scene.add_shape(Plane(Point3(0.f, -2.f, 2.f), Vector3(0.f, 0.f, -1.f), plane_material));
    plane_material = std::make_shared<TexturePhong>(dim_white, checkerboard, dim_white, 60.f, 0.f, 0.f);
    scene.add_shape(Plane(Point3(0.f, -2.f, 0.f), Vector3(0.f, 1.f, 0.f), plane_material));

    auto cube_material = std::make_shared<Phong>(BLACK, WHITE, WHITE, 60.f, 0.f, 1.333f);
    Cube cube(cube_material);
    cube.set_transform(Matrix::translate(-1.f, -1.0f, -4.f) * Matrix::rotate_x(-45.0f));
    scene.add_shape(std::move(cube));

    scene.add_light(PointLight(Point3(4.f, 4.0f, 0.f), Color(1.f, 0.f, 0.f)));
    scene.add_light(PointLight(Point3(-1.f, 2.0f, -4.f), Color(0.f, 1.f, 0.f)));
    scene.add_light(PointLight(Point3(0.f, 8.0f, -4.f), Color(0.f, 0.f, 1.f)));
    scene.set_ambient(Color(0.1f, 0.1f, 0.1f));
}

struct SplitMix64 {
    uint64_t s;
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    float uniform() { return (float)(next() >> 40) * (1.0f / 16777216.0f); }
    float range(float lo, float hi) { return lo + uniform() * (hi - lo); }
};

// material mix by index: 0-4 opaque, 5-7 mirror, 8-9 glass
MaterialRef synthetic_material(int i, SplitMix64 &rng) {
    Color col(rng.range(0.2f, 1.f), rng.range(0.2f, 1.f), rng.range(0.2f, 1.f));
    int k = i % 10;
    if (k <= 4) return std::make_shared<Phong>(0.1f * col, col, WHITE, 60.f, 0.f, 0.f);
    if (k <= 7) return std::make_shared<Phong>(BLACK, col, WHITE, 60.f, 0.5f, 0.f);
    return std::make_shared<Phong>(BLACK, WHITE, WHITE, 60.f, 0.7f, 1.333f);
}

void synthetic_scene(Scene &scene, int n_spheres, int n_triangles, uint64_t seed, float r_lo, float r_hi) {
    SplitMix64 rng{seed};
    const ColorFun dim_white = ColorFun::constant(0.1f * WHITE);
    // the floor of the default scene (plane4 of src/my_scene.rs:95-107), added first
    scene.add_shape(Plane(Point3(0.f, -2.f, 0.f), Vector3(0.f, 1.f, 0.f),
                          std::make_shared<TexturePhong>(dim_white, ColorFun::checkerboard(WHITE), dim_white, 60.f, 0.f, 0.f)));
    for (int i = 0; i < n_spheres; ++i) {
        float cx = rng.range(-3.5f, 3.5f), cy = rng.range(-1.8f, 3.0f), cz = rng.range(-2.f, 6.f);
        float r = rng.range(r_lo, r_hi);
        MaterialRef m = synthetic_material(i, rng);
        Sphere s(m);
        if (i % 10 == 0) { // every 10th: a rotated ellipsoid, to exercise the affine path
            float theta = rng.range(0.f, 180.f);
            s.set_transform(Matrix::translate(cx, cy, cz) * Matrix::rotate_z(theta) * Matrix::scale(r, r * 0.5f, r));
        } else {
            s.set_transform(Matrix::translate(cx, cy, cz) * Matrix::scale(r, r, r));
        }
        scene.add_shape(std::move(s));
    }
    for (int i = 0; i < n_triangles; ++i) {
        Point3 v0(rng.range(-3.5f, 3.5f), rng.range(-1.8f, 3.0f), rng.range(-2.f, 6.f));
        Vector3 e1(rng.range(-0.15f, 0.15f), rng.range(-0.15f, 0.15f), rng.range(-0.15f, 0.15f));
        Vector3 e2(rng.range(-0.15f, 0.15f), rng.range(-0.15f, 0.15f), rng.range(-0.15f, 0.15f));
        MaterialRef m = synthetic_material(i, rng);
        scene.add_shape(Triangle(v0, v0 + e1, v0 + e2, m));
    }
    scene.add_light(PointLight(Point3(4.f, 4.f, -2.f), 0.6f * WHITE));
    scene.add_light(PointLight(Point3(-4.f, 4.f, -2.f), 0.4f * WHITE));
    scene.add_light(PointLight(Point3(0.f, 8.f, -4.f), 0.3f * WHITE));
    scene.add_light(PointLight(Point3(0.f, 2.f, -7.f), 0.2f * WHITE));
    scene.set_ambient(Color(0.1f, 0.1f, 0.1f));
}

int export_scene(const Scene &scene, rt_scene_desc **out) {
    OwnedDesc *o = new (std::nothrow) OwnedDesc();
    if (!o) return RT_ERR_NOMEM;
    scene.flatten(o->flat);
    o->desc = o->flat.desc;
    *out = &o->desc;
    return RT_OK;
}

} // namespace

extern "C" {

int rt_synthetic_scene(int32_t n_spheres, int32_t n_triangles, uint64_t seed, float r_lo, float r_hi,
                       rt_scene_desc **out) {
    if (!out || n_spheres < 0 || n_triangles < 0 || !(r_lo > 0.f) || !(r_hi >= r_lo)) return RT_ERR_INVALID;
    try {
        Scene scene;
        synthetic_scene(scene, n_spheres, n_triangles, seed, r_lo, r_hi);
        return export_scene(scene, out);
    } catch (const std::exception &) {
        return RT_ERR_INVALID;
    }
}

int rt_builtin_scene(const char *name, uint64_t seed, rt_scene_desc **out) {
    if (!name || !out) return RT_ERR_INVALID;
    std::string n(name);
    try {
        if (n == "default") {
            Scene scene;
            create_scene(scene);
            return export_scene(scene, out);
        }
        if (n == "s1k") return rt_synthetic_scene(1000, 0, seed ? seed : 1000, 0.05f, 0.20f, out);
        if (n == "m10k") return rt_synthetic_scene(5000, 5000, seed ? seed : 10000, 0.03f, 0.10f, out);
    } catch (const std::exception &) {
        return RT_ERR_INVALID;
    }
    return RT_ERR_INVALID;
}

void rt_scene_desc_free(rt_scene_desc *desc) { delete reinterpret_cast<OwnedDesc *>(desc); }

int rt_matrix_inverse(const float *m16, float *out16) {
    Matrix m;
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) m.mat[r][c] = m16[r * 4 + c];
    try {
        m.inverse().store(out16);
    } catch (const std::runtime_error &) {
        return RT_ERR_INVALID;
    }
    return RT_OK;
}

void rt_matrix_compose(float tx, float ty, float tz, float rx, float ry, float rz, float sx, float sy, float sz,
                       float *out16) {
    Matrix m = Matrix::translate(tx, ty, tz) * Matrix::rotate_z(rz) * Matrix::rotate_y(ry) * Matrix::rotate_x(rx) *
               Matrix::scale(sx, sy, sz);
    m.store(out16);
}

void rt_plane_axes(const float *n, float *u3, float *v3) {
    Plane p(Point3(0.f, 0.f, 0.f), Vector3(n[0], n[1], n[2]), nullptr);
    u3[0] = p.u().x; u3[1] = p.u().y; u3[2] = p.u().z;
    v3[0] = p.v().x; v3[1] = p.v().y; v3[2] = p.v().z;
}

void rt_triangle_normal(const float *v0, const float *v1, const float *v2, float *n3) {
    Triangle t(Point3(v0[0], v0[1], v0[2]), Point3(v1[0], v1[1], v1[2]), Point3(v2[0], v2[1], v2[2]), nullptr);
    n3[0] = t.normal().x; n3[1] = t.normal().y; n3[2] = t.normal().z;
}

} // extern "C"
