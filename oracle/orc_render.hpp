// ORACLE -- TEST INFRASTRUCTURE ONLY (see orc_math.hpp).
// Restates src/render.rs (Basic renderer) and src/render_tree.rs (RayForest renderer).
#pragma once
#include <algorithm>
#include <set>
#include <vector>

#include "orc_scene.hpp"

namespace orc {

// ---- src/render.rs --------------------------------------------------------------------

// Camera::get_ray (src/render.rs:178-185).  (ox, oy) is the sub-pixel offset in pixel
// units; the reference has none (pixel corner) -> 0, 0.
inline Ray camera_ray(const rt_camera &cam, int u, int v, float ox, float oy, bool jitter) {
    float x_delta = (cam.x_max - cam.x_min) / (float)cam.x_res;
    float y_delta = (cam.y_max - cam.y_min) / (float)cam.y_res;
    float fu = (float)u, fv = (float)v;
    if (jitter) { fu = fu + ox; fv = fv + oy; }
    float x = cam.x_min + fu * x_delta;
    float y = cam.y_max - fv * y_delta;
    Point3 origin(cam.origin[0], cam.origin[1], cam.origin[2]);
    Point3 viewpoint(x, y, 0.f);
    return Ray(origin, (viewpoint - origin).norm());
}

// 4-spp extension (NO reference behaviour): SplitMix64 stream per (pixel, sample).
inline uint64_t splitmix64(uint64_t &state) {
    uint64_t z = (state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline float u01(uint64_t x) { return (float)(x >> 40) * (1.0f / 16777216.0f); }
inline void jitter_offset(uint64_t seed, const rt_camera &cam, int u, int v, int s, float &ox, float &oy) {
    uint64_t key = ((uint64_t)v * (uint64_t)cam.x_res + (uint64_t)u) * 4ull + (uint64_t)s;
    uint64_t st = seed ^ (key * 0xD1B54A32D192ED03ull);
    float xi1 = u01(splitmix64(st));
    float xi2 = u01(splitmix64(st));
    ox = ((float)(s & 1) + xi1) * 0.5f;
    oy = ((float)(s >> 1) + xi2) * 0.5f;
}

// src/render.rs:129-134   powi(5) = x * ((x*x) * (x*x))  (compiler-rt __powisf2 order)
inline float powi5(float x) {
    float x2 = x * x;
    float x4 = x2 * x2;
    return x * x4;
}
inline float fresnel_reflection(const Vector3 &light_dir, const Vector3 &normal, float n1, float n2) {
    float m_dot_r = light_dir.dot(normal);
    float r0 = ((n1 - n2) / (n1 + n2)) * ((n1 - n2) / (n1 + n2));
    return r0 + (1.f - r0) * powi5(1.f - m_dot_r);
}
// src/render.rs:138-140
inline float fresnel_refraction(const Vector3 &light_dir, const Vector3 &normal, float n1, float n2) {
    return 1.f - fresnel_reflection(light_dir, normal, n1, n2);
}

// src/render.rs:105-110
inline Ray reflect_ray(const Ray &ray, const Intersection &i) {
    Vector3 reflected_dir = -(ray.direction.reflect(i.normal).norm());
    Point3 p = i.point + 0.0002f * reflected_dir;
    return Ray(p, reflected_dir);
}

// src/render.rs:112-125
inline bool refract_ray(const Ray &ray, const Intersection &i, float n1, float n2, Ray &out) {
    float ratio = n1 / n2;
    float m_dot_r = -(ray.direction.dot(i.normal));
    float cos_theta_sqrd = 1.f - ratio * ratio * (1.f - m_dot_r * m_dot_r);
    if (cos_theta_sqrd > 0.f) {
        float cos_theta = std::sqrt(cos_theta_sqrd);
        Vector3 refract_dir = ray.direction * ratio + i.normal * (ratio * m_dot_r - cos_theta);
        Point3 p = i.point + 0.0002f * refract_dir;
        out = Ray(p, refract_dir);
        return true;
    }
    return false;
}

struct LightSample {
    Vector3 dir;
    Color energy;
};

// PointLight::get_energy (src/scene/mod.rs:189-206), AmbientLight::get_energy (:236-238)
inline LightSample light_energy(const Scene &scene, const Light &l, const Point3 &point, Counters &cnt) {
    LightSample out;
    if (l.kind == RT_LIGHT_AMBIENT) {
        out.dir = Vector3(0.f, 0.f, 0.f);
        out.energy = l.color;
        return out;
    }
    Vector3 dir_to_light = (l.pos - point).norm();
    Ray ray(point, dir_to_light);
    Intersection i;
    cnt.shadow_rays += 1;
    if (scene.intersect(ray, i, cnt)) {
        if ((i.point - point).len2() < (l.pos - point).len2()) out.energy = BLACK;
        else out.energy = l.color;
    } else {
        out.energy = l.color;
    }
    out.dir = dir_to_light;
    return out;
}

// src/render.rs:142-153
inline void get_light_energy(const Scene &scene, const Intersection &i, std::vector<LightSample> &out, Counters &cnt) {
    Point3 p = i.point + 0.0002f * i.normal;
    out.clear();
    for (const Light &l : scene.lights) out.push_back(light_energy(scene, l, p, cnt));
}

// src/render.rs:40-103
inline Color trace_ray(const Scene &scene, const Ray &ray, int depth, Counters &cnt) {
    if (depth == 0) return BLACK;
    Intersection i;
    if (!scene.intersect(ray, i, cnt)) return BLACK;
    cnt.nodes += 1;
    const rt_material &m = scene.materials[i.material];
    float n1, n2;
    if (i.entering) { n1 = 1.f; n2 = m.refraction_index; }
    else { n1 = m.refraction_index; n2 = 1.f; }

    Color ambient = eval_colorfn(m.ambient, i.tex_u, i.tex_v) * scene.ambient;

    std::vector<LightSample> ls;
    get_light_energy(scene, i, ls, cnt);
    Color lights = BLACK; // Sum for Color (src/scene/color.rs:97-107)
    for (const LightSample &s : ls) {
        float fresnel = fresnel_reflection(s.dir, i.normal, n1, n2);
        Color c = fresnel * get_reflected_energy(m, s.energy, s.dir, i);
        lights = lights + c;
    }

    Color reflected = BLACK;
    if (m.reflectivity > EPS) {
        Ray rr = reflect_ray(ray, i);
        Color energy = trace_ray(scene, rr, depth - 1, cnt);
        float fresnel = fresnel_reflection(rr.direction, i.normal, n1, n2);
        reflected = fresnel * get_reflected_energy(m, energy, rr.direction, i);
    }

    Color refracted = BLACK;
    if (m.refraction_index > EPS) {
        Ray tr;
        Color inner = BLACK;
        if (refract_ray(ray, i, n1, n2, tr)) {
            float fresnel = fresnel_refraction(tr.direction, i.normal.neg(), n1, n2);
            inner = fresnel * trace_ray(scene, tr, depth - 1, cnt);
        }
        refracted = eval_colorfn(m.diffuse, i.tex_u, i.tex_v) * inner;
    }

    return ((ambient + lights) + reflected) + refracted;
}

// ---- src/render_tree.rs ---------------------------------------------------------------

struct TreeNode { // RayTreeNode::Branch (:12-21); child index -1 = RayTreeNode::None
    Intersection i;
    std::vector<LightSample> lights;
    int reflected = -1, refracted = -1;
};

struct RayTree { // :23-28
    std::vector<TreeNode> nodes;
    int root = -1;
    std::set<int32_t> shapes;
    size_t size() const { return nodes.size(); } // every stored node is reachable from root
};

// build_ray_tree (:166-212)
inline int build_ray_tree(const Scene &scene, const Ray &ray, int depth, RayTree &tree, Counters &cnt) {
    if (depth == 0) return -1;
    Intersection i;
    if (!scene.intersect(ray, i, cnt)) return -1;
    cnt.nodes += 1;
    tree.shapes.insert(i.id);
    const rt_material &m = scene.materials[i.material];
    float n1, n2;
    if (i.entering) { n1 = 1.f; n2 = m.refraction_index; }
    else { n1 = m.refraction_index; n2 = 1.f; }

    int me = (int)tree.nodes.size();
    tree.nodes.emplace_back();
    tree.nodes[me].i = i;
    {
        std::vector<LightSample> ls;
        get_light_energy(scene, i, ls, cnt);
        tree.nodes[me].lights = ls;
    }
    int refl = -1;
    if (m.reflectivity > EPS) {
        Ray rr = reflect_ray(ray, i);
        refl = build_ray_tree(scene, rr, depth - 1, tree, cnt);
    }
    int refr = -1;
    if (m.refraction_index > EPS) {
        Ray tr;
        if (refract_ray(ray, i, n1, n2, tr)) refr = build_ray_tree(scene, tr, depth - 1, tree, cnt);
    }
    tree.nodes[me].reflected = refl;
    tree.nodes[me].refracted = refr;
    return me;
}

// render_ray_tree (:214-255): returns (colour, -eye_dir); None -> (BLACK, (0,0,0))
inline Color render_ray_tree(const Scene &scene, const RayTree &tree, int node, const Color &ambient_in, Vector3 &dir_out) {
    if (node < 0) {
        dir_out = Vector3(0.f, 0.f, 0.f);
        return BLACK;
    }
    const TreeNode &nd = tree.nodes[node];
    const Intersection &i = nd.i;
    const rt_material &m = scene.materials[i.material];
    float n1, n2;
    if (i.entering) { n1 = 1.f; n2 = m.refraction_index; }
    else { n1 = m.refraction_index; n2 = 1.f; }

    Color lights = BLACK;
    for (const LightSample &s : nd.lights) {
        float fresnel = fresnel_reflection(s.dir, i.normal, n1, n2);
        Color c = fresnel * get_reflected_energy(m, s.energy, s.dir, i);
        lights = lights + c;
    }
    Color reflected;
    {
        Vector3 dir;
        Color energy = render_ray_tree(scene, tree, nd.reflected, ambient_in, dir);
        float fresnel = fresnel_reflection(dir, i.normal, n1, n2);
        reflected = fresnel * get_reflected_energy(m, energy, i.eye_dir, i);
    }
    Color refracted;
    {
        Vector3 dir;
        Color energy = render_ray_tree(scene, tree, nd.refracted, ambient_in, dir);
        float fresnel = fresnel_refraction(dir, i.normal.neg(), n1, n2);
        refracted = fresnel * energy;
    }
    Color ambient = eval_colorfn(m.ambient, i.tex_u, i.tex_v) * ambient_in;
    dir_out = -i.eye_dir;
    return ((ambient + lights) + reflected) + refracted;
}

} // namespace orc
