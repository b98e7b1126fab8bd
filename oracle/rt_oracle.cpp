// ORACLE -- TEST INFRASTRUCTURE ONLY (see orc_math.hpp).
// C entry points of liborc.so, bound with ctypes by tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs.  Build: oracle/Makefile
// (g++ -O2 -ffp-contract=off, mirroring opt-level = 2 of .cargo/config:4-6).
#include <atomic>
#include <cstring>
#include <thread>

#include "orc_render.hpp"

using namespace orc;

namespace {

void mat_out(const Matrix &m, float *out) {
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) out[r * 4 + c] = m.mat[r][c];
}
Matrix mat_in(const float *in) {
    Matrix m;
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) m.mat[r][c] = in[r * 4 + c];
    return m;
}
rt_colorfn cf(int kind, float r, float g, float b) {
    rt_colorfn f;
    f.kind = kind;
    f.rgb[0] = r; f.rgb[1] = g; f.rgb[2] = b;
    return f;
}
rt_material phong_mat(Color a, Color d, Color s, float power, float refl, float idx) {
    rt_material m;
    m.kind = RT_MATERIAL_PHONG;
    m.ambient = cf(RT_COLOR_CONST, a.r, a.g, a.b);
    m.diffuse = cf(RT_COLOR_CONST, d.r, d.g, d.b);
    m.specular = cf(RT_COLOR_CONST, s.r, s.g, s.b);
    m.power = power; m.reflectivity = refl; m.refraction_index = idx;
    return m;
}
void set_transform(rt_shape &s, const Matrix &t) { // Sphere::set_transform (sphere.rs:100-103)
    mat_out(t, s.transform);
    mat_out(t.inverse(), s.inv_transform);
}
rt_shape blank_shape(int kind, int material) {
    rt_shape s;
    std::memset(&s, 0, sizeof(s));
    s.kind = kind;
    s.material = material;
    mat_out(Matrix::identity(), s.transform);
    mat_out(Matrix::identity(), s.inv_transform);
    return s;
}
rt_shape plane_shape(Point3 o, Vector3 n, int material) { // Plane::new (plane.rs:22-42)
    rt_shape s = blank_shape(RT_SHAPE_PLANE, material);
    Vector3 u, v;
    plane_axes(n, u, v);
    float p[12] = {o.x, o.y, o.z, n.x, n.y, n.z, u.x, u.y, u.z, v.x, v.y, v.z};
    std::memcpy(s.p, p, sizeof(p));
    return s;
}

struct Job {
    const Scene *scene;
    const rt_camera *cam;
    int depth, spp, mode;
    uint64_t seed;
    const int32_t *rows;
    int n_rows;
    float *rgb_f32;
    uint8_t *rgb_u8;
};

Color render_pixel(const Job &j, int u, int v, Counters &cnt) {
    const Scene &scene = *j.scene;
    auto shade = [&](const Ray &ray) -> Color {
        if (j.mode == 0) return trace_ray(scene, ray, j.depth, cnt);
        RayTree tree; // generate_ray_forest + render_forest for one pixel
        tree.root = build_ray_tree(scene, ray, j.depth, tree, cnt);
        Vector3 dir;
        return render_ray_tree(scene, tree, tree.root, scene.ambient, dir);
    };
    if (j.spp <= 1) return shade(camera_ray(*j.cam, u, v, 0.f, 0.f, false));
    Color sum = BLACK;
    for (int s = 0; s < 4; ++s) {
        float ox, oy;
        jitter_offset(j.seed, *j.cam, u, v, s, ox, oy);
        sum = sum + shade(camera_ray(*j.cam, u, v, ox, oy, true));
    }
    return 0.25f * sum;
}

void worker(const Job &j, std::atomic<int> &next, Counters &total_out) {
    Counters cnt;
    const int w = j.cam->x_res;
    for (;;) {
        int k = next.fetch_add(1);
        if (k >= j.n_rows) break;
        int v = j.rows[k];
        for (int u = 0; u < w; ++u) {
            Color c = render_pixel(j, u, v, cnt);
            size_t o = ((size_t)k * (size_t)w + (size_t)u) * 3;
            if (j.rgb_f32) { j.rgb_f32[o] = c.r; j.rgb_f32[o + 1] = c.g; j.rgb_f32[o + 2] = c.b; }
            if (j.rgb_u8) {
                j.rgb_u8[o] = as_u8_channel(c.r);
                j.rgb_u8[o + 1] = as_u8_channel(c.g);
                j.rgb_u8[o + 2] = as_u8_channel(c.b);
            }
        }
    }
    total_out = cnt;
}

} // namespace

extern "C" {

// counters layout shared with ctypes: 7 x uint64
struct orc_counters {
    uint64_t rays, shadow_rays, nodes, sphere_tests, plane_tests, triangle_tests, cube_tests;
};

// ---- src/my_scene.rs:45-120 restated: 6 shapes, 6 materials, 3 lights ------------------
int orc_default_scene(rt_shape *shapes, rt_material *materials, rt_light *lights, float *ambient) {
    const Color RED(1.f, 0.f, 0.f), BLUE(0.f, 0.f, 1.f);
    const Color DIM_WHITE(0.1f, 0.1f, 0.1f), DIM_BLUE(0.f, 0.f, 0.1f);
    Color dim_white_fn = 0.1f * WHITE; // fn dim_white (my_scene.rs:11-13)

    materials[0] = phong_mat(DIM_WHITE, RED, WHITE, 60.f, 0.5f, 0.f);
    shapes[0] = blank_shape(RT_SHAPE_SPHERE, 0);
    set_transform(shapes[0], Matrix::translate(-1.0f, 0.f, 0.f) * Matrix::rotate_z(75.f) * Matrix::scale(1.0f, 0.25f, 1.0f));

    materials[1] = phong_mat(BLACK, BLUE, DIM_BLUE, 600.f, 0.4f, 0.f);
    shapes[1] = blank_shape(RT_SHAPE_SPHERE, 1); // "blue"
    set_transform(shapes[1], Matrix::translate(1.f, -1.f, 0.f));

    materials[2] = phong_mat(BLACK, WHITE, WHITE, 60.f, 0.7f, 1.333f);
    shapes[2] = blank_shape(RT_SHAPE_SPHERE, 2);
    set_transform(shapes[2], Matrix::translate(0.f, -0.5f, -3.f) * Matrix::scale(0.6f, 0.6f, 0.6f));

    for (int k = 3; k <= 4; ++k) {
        rt_material m;
        m.kind = RT_MATERIAL_TEXTURE_PHONG;
        m.ambient = cf(RT_COLOR_CONST, dim_white_fn.r, dim_white_fn.g, dim_white_fn.b);
        m.diffuse = cf(RT_COLOR_CHECKERBOARD, 1.f, 1.f, 1.f);
        m.specular = cf(RT_COLOR_CONST, dim_white_fn.r, dim_white_fn.g, dim_white_fn.b);
        m.power = 60.f; m.reflectivity = 0.f; m.refraction_index = 0.f;
        materials[k] = m;
    }
    shapes[3] = plane_shape(Point3(0.f, -2.f, 2.f), Vector3(0.f, 0.f, -1.f), 3);
    shapes[4] = plane_shape(Point3(0.f, -2.f, 0.f), Vector3(0.f, 1.f, 0.f), 4);

    materials[5] = phong_mat(BLACK, WHITE, WHITE, 60.f, 0.f, 1.333f);
    shapes[5] = blank_shape(RT_SHAPE_CUBE, 5);
    set_transform(shapes[5], Matrix::translate(-1.f, -1.0f, -4.f) * Matrix::rotate_x(-45.0f));

    const float lp[3][3] = {{4.f, 4.0f, 0.f}, {-1.f, 2.0f, -4.f}, {0.f, 8.0f, -4.f}};
    const float lc[3][3] = {{1.f, 0.f, 0.f}, {0.f, 1.f, 0.f}, {0.f, 0.f, 1.f}};
    for (int k = 0; k < 3; ++k) {
        lights[k].kind = RT_LIGHT_POINT;
        std::memcpy(lights[k].pos, lp[k], sizeof(lp[k]));
        std::memcpy(lights[k].color, lc[k], sizeof(lc[k]));
    }
    ambient[0] = ambient[1] = ambient[2] = 0.1f;
    return 0;
}

// render::render / render_tree::{generate_ray_forest + render_forest} over the listed rows.
// mode 0 = Basic, 1 = RayForest.  Output rows are packed in the order of rows[].
int orc_render(const rt_scene_desc *desc, const rt_camera *cam, int depth, int spp, uint64_t seed,
               const int32_t *rows, int n_rows, int mode, int threads, float *rgb_f32, uint8_t *rgb_u8,
               orc_counters *counters) {
    Scene scene;
    if (!scene_from_desc(*desc, scene)) return -1;
    if (spp != 1 && spp != 4) return -1;
    Job j{&scene, cam, depth, spp, mode, seed, rows, n_rows, rgb_f32, rgb_u8};
    if (threads < 1) threads = 1;
    std::atomic<int> next(0);
    std::vector<Counters> per(threads);
    if (threads == 1) {
        worker(j, next, per[0]);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back([&, t] { worker(j, next, per[t]); });
        for (auto &th : pool) th.join();
    }
    Counters total;
    for (auto &c : per) total.add(c);
    if (counters) {
        counters->rays = total.rays; counters->shadow_rays = total.shadow_rays; counters->nodes = total.nodes;
        counters->sphere_tests = total.sphere_tests; counters->plane_tests = total.plane_tests;
        counters->triangle_tests = total.triangle_tests; counters->cube_tests = total.cube_tests;
    }
    return 0;
}

// Scene::intersect for one ray: returns 1 on hit and fills rec[15] =
// {id, t, material, point(3), eye_dir(3), normal(3), entering, tex_u, tex_v} (ints as floats).
int orc_intersect(const rt_scene_desc *desc, const float *o, const float *d, float *rec) {
    Scene scene;
    if (!scene_from_desc(*desc, scene)) return -1;
    Counters cnt;
    Intersection i;
    Ray ray(Point3(o[0], o[1], o[2]), Vector3(d[0], d[1], d[2]));
    if (!scene.intersect(ray, i, cnt)) return 0;
    float r[14] = {(float)i.id, i.t, (float)i.material, i.point.x, i.point.y, i.point.z, i.eye_dir.x, i.eye_dir.y,
                   i.eye_dir.z, i.normal.x, i.normal.y, i.normal.z, i.entering ? 1.f : 0.f, i.tex_u};
    std::memcpy(rec, r, sizeof(r));
    rec[14] = i.tex_v;
    return 1;
}

// Batched nearest-hit for differential tests: hit[k] = shape id or -1, t[k].
int orc_intersect_batch(const rt_scene_desc *desc, const float *o, const float *d, int n, int32_t *hit, float *t) {
    Scene scene;
    if (!scene_from_desc(*desc, scene)) return -1;
    Counters cnt;
    for (int k = 0; k < n; ++k) {
        Intersection i;
        Ray ray(Point3(o[3 * k], o[3 * k + 1], o[3 * k + 2]), Vector3(d[3 * k], d[3 * k + 1], d[3 * k + 2]));
        if (scene.intersect(ray, i, cnt)) { hit[k] = i.id; t[k] = i.t; }
        else { hit[k] = -1; t[k] = 0.f; }
    }
    return 0;
}

// trace_ray for one ray (src/render.rs:40-103).
int orc_trace(const rt_scene_desc *desc, const float *o, const float *d, int depth, float *rgb) {
    Scene scene;
    if (!scene_from_desc(*desc, scene)) return -1;
    Counters cnt;
    Color c = trace_ray(scene, Ray(Point3(o[0], o[1], o[2]), Vector3(d[0], d[1], d[2])), depth, cnt);
    rgb[0] = c.r; rgb[1] = c.g; rgb[2] = c.b;
    return 0;
}

// ---- known-answer helpers (one per reference unit test family) -------------------------
void orc_mat_identity(float *out) { mat_out(Matrix::identity(), out); }
void orc_mat_scale(float x, float y, float z, float *out) { mat_out(Matrix::scale(x, y, z), out); }
void orc_mat_translate(float x, float y, float z, float *out) { mat_out(Matrix::translate(x, y, z), out); }
void orc_mat_rotate(int axis, float angle, float *out) {
    mat_out(axis == 0 ? Matrix::rotate_x(angle) : axis == 1 ? Matrix::rotate_y(angle) : Matrix::rotate_z(angle), out);
}
void orc_mat_mul(const float *a, const float *b, float *out) { mat_out(mat_in(a) * mat_in(b), out); }
void orc_mat_transpose(const float *a, float *out) { mat_out(mat_in(a).transpose(), out); }
int orc_mat_inverse(const float *a, float *out) {
    bool ok = true;
    Matrix inv = mat_in(a).inverse(&ok);
    mat_out(inv, out);
    return ok ? 0 : 1; // 1 where the reference panics "Singular Matrix"
}
void orc_mat_pt_mul(const float *a, const float *p, float *out) {
    Point3 r = mat_in(a).pt_mul(Point3(p[0], p[1], p[2]));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void orc_mat_vec3_mul(const float *a, const float *v, float *out) {
    Vector3 r = mat_in(a).vec3_mul(Vector3(v[0], v[1], v[2]));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void orc_vec3_reflect(const float *v, const float *about, float *out) {
    Vector3 r = Vector3(v[0], v[1], v[2]).reflect(Vector3(about[0], about[1], about[2]));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void orc_vec3_norm(const float *v, float *out) {
    Vector3 r = Vector3(v[0], v[1], v[2]).norm();
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void orc_vec3_cross(const float *a, const float *b, float *out) {
    Vector3 r = Vector3(a[0], a[1], a[2]).cross(Vector3(b[0], b[1], b[2]));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
// Triangle::new's normal and Plane::new's axes, as the reference constructors compute them
void orc_triangle_normal(const float *v0, const float *v1, const float *v2, float *n) {
    Triangle t = Triangle::make(Point3(v0[0], v0[1], v0[2]), Point3(v1[0], v1[1], v1[2]), Point3(v2[0], v2[1], v2[2]), 0);
    n[0] = t.normal.x; n[1] = t.normal.y; n[2] = t.normal.z;
}
void orc_plane_axes(const float *n, float *u, float *v) {
    Vector3 uu, vv;
    plane_axes(Vector3(n[0], n[1], n[2]), uu, vv);
    u[0] = uu.x; u[1] = uu.y; u[2] = uu.z;
    v[0] = vv.x; v[1] = vv.y; v[2] = vv.z;
}
// The 12 cube triangles: out[12][12] = v0, v1, v2, normal
void orc_cube_triangles(float *out) {
    std::vector<Triangle> t = cube_triangles(0);
    for (int k = 0; k < 12; ++k) {
        float r[12] = {t[k].v0.x, t[k].v0.y, t[k].v0.z, t[k].v1.x, t[k].v1.y, t[k].v1.z,
                       t[k].v2.x, t[k].v2.y, t[k].v2.z, t[k].normal.x, t[k].normal.y, t[k].normal.z};
        std::memcpy(out + 12 * k, r, sizeof(r));
    }
}
// Material::get_reflected_energy for a hand-built intersection (triangle.rs:193 "shading")
void orc_reflected_energy(const rt_material *m, const float *incoming, const float *light_dir, const float *eye_dir,
                          const float *normal, float tex_u, float tex_v, float *out) {
    Intersection i;
    i.id = 0; i.t = 0.f; i.material = 0; i.entering = true;
    i.eye_dir = Vector3(eye_dir[0], eye_dir[1], eye_dir[2]);
    i.normal = Vector3(normal[0], normal[1], normal[2]);
    i.tex_u = tex_u; i.tex_v = tex_v;
    Color c = get_reflected_energy(*m, Color(incoming[0], incoming[1], incoming[2]),
                                   Vector3(light_dir[0], light_dir[1], light_dir[2]), i);
    out[0] = c.r; out[1] = c.g; out[2] = c.b;
}
void orc_color_ops(const float *a, const float *b, float s, float *mul, float *smul, float *add) {
    Color ca(a[0], a[1], a[2]), cb(b[0], b[1], b[2]);
    Color m = ca * cb, sm = s * cb, ad = ca + cb;
    mul[0] = m.r; mul[1] = m.g; mul[2] = m.b;
    smul[0] = sm.r; smul[1] = sm.g; smul[2] = sm.b;
    add[0] = ad.r; add[1] = ad.g; add[2] = ad.b;
}
uint8_t orc_as_u8(float c) { return as_u8_channel(c); }
void orc_camera_ray(const rt_camera *cam, int u, int v, float *o, float *d) {
    Ray r = camera_ray(*cam, u, v, 0.f, 0.f, false);
    o[0] = r.origin.x; o[1] = r.origin.y; o[2] = r.origin.z;
    d[0] = r.direction.x; d[1] = r.direction.y; d[2] = r.direction.z;
}
float orc_fresnel_reflection(const float *dir, const float *n, float n1, float n2) {
    return fresnel_reflection(Vector3(dir[0], dir[1], dir[2]), Vector3(n[0], n[1], n[2]), n1, n2);
}

} // extern "C"
