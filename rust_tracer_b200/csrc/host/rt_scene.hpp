// rt_scene.hpp -- host-side mirror of the reference's `scene` module (src/scene/mod.rs):
// Scene, Renderable (Sphere, Plane, Triangle, Cube), Material (Phong, TexturePhong),
// LightSource (PointLight, AmbientLight).  Same names and constructor arguments as the
// reference so that src/my_scene.rs reads the same here (builtin_scenes.cpp).
//
// The one addition is `flatten`, the accessor the reference's trait objects lack
// (Renderable exposes no geometry, src/scene/mod.rs:147-165): it writes the shape into the
// flat rt_scene_desc the device library consumes.  There is NO host intersect(): tracing
// exists on the device only.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../../include/rt_b200.h"
#include "rt_math.hpp"

namespace rust_tracer {

using TextureCoords = std::pair<float, float>;

// ColorFun = fn((f32,f32)) -> Color (src/scene/material.rs:3).  Function pointers cannot
// cross the FFI, so a ColorFun is one of the device texture functions.
struct ColorFun {
    rt_colorfn fn{};
    static ColorFun constant(const Color &c) {
        ColorFun f;
        f.fn.kind = RT_COLOR_CONST;
        f.fn.rgb[0] = c.r; f.fn.rgb[1] = c.g; f.fn.rgb[2] = c.b;
        return f;
    }
    // src/my_scene.rs:26-43 (`base` is WHITE there)
    static ColorFun checkerboard(const Color &base = colors::WHITE) {
        ColorFun f = constant(base);
        f.fn.kind = RT_COLOR_CHECKERBOARD;
        return f;
    }
};

class Material { // src/scene/material.rs:9-22
  public:
    virtual ~Material() = default;
    virtual void set_diffuse(const Color &c) = 0;
    virtual float reflectivity() const = 0;
    virtual float refraction_index() const = 0;
    virtual std::string to_string() const = 0;
    virtual void flatten(rt_material &out) const = 0;
};

class Phong : public Material { // src/scene/material.rs:24-101
  public:
    Color specular, ambient, diffuse;
    float power, reflectivity_, refraction_index_;
    Phong(const Color &ambient_, const Color &diffuse_, const Color &specular_, float power_, float reflectivity,
          float refraction_index)
        : specular(specular_), ambient(ambient_), diffuse(diffuse_), power(power_), reflectivity_(reflectivity),
          refraction_index_(refraction_index) {}
    void set_diffuse(const Color &c) override { diffuse = c; }
    float reflectivity() const override { return reflectivity_; }
    float refraction_index() const override { return refraction_index_; }
    std::string to_string() const override { return "Phong"; }
    void flatten(rt_material &out) const override {
        out.kind = RT_MATERIAL_PHONG;
        out.ambient = ColorFun::constant(ambient).fn;
        out.diffuse = ColorFun::constant(diffuse).fn;
        out.specular = ColorFun::constant(specular).fn;
        out.power = power; out.reflectivity = reflectivity_; out.refraction_index = refraction_index_;
    }
};

class TexturePhong : public Material { // src/scene/material.rs:103-191
  public:
    ColorFun specular, ambient, diffuse;
    float power, reflectivity_, refraction_index_;
    TexturePhong(const ColorFun &ambient_, const ColorFun &diffuse_, const ColorFun &specular_, float power_,
                 float reflectivity, float refraction_index)
        : specular(specular_), ambient(ambient_), diffuse(diffuse_), power(power_), reflectivity_(reflectivity),
          refraction_index_(refraction_index) {}
    void set_diffuse(const Color &) override {} // material.rs:149: no-op
    float reflectivity() const override { return reflectivity_; }
    float refraction_index() const override { return refraction_index_; }
    std::string to_string() const override { return "Phong"; }
    void flatten(rt_material &out) const override {
        out.kind = RT_MATERIAL_TEXTURE_PHONG;
        out.ambient = ambient.fn; out.diffuse = diffuse.fn; out.specular = specular.fn;
        out.power = power; out.reflectivity = reflectivity_; out.refraction_index = refraction_index_;
    }
};

using MaterialRef = std::shared_ptr<Material>; // Rc<RefCell<dyn Material>>

class Renderable { // src/scene/mod.rs:147-165
  public:
    virtual ~Renderable() = default;
    int32_t id() const { return id_; }
    void set_id(int32_t id) { id_ = id; }
    virtual void set_transform(const Matrix &m) {
        transform = m;
        inv_transform = m.inverse();
    }
    virtual MaterialRef get_material() const = 0; // None for Cube in the reference (cube.rs:120-126)
    virtual MaterialRef flatten_material() const { return get_material(); }
    virtual std::string get_name() const = 0;
    virtual std::string to_string() const { return get_name(); }
    virtual size_t size() const { return 1; }
    virtual void flatten(rt_shape &out) const = 0;

    Matrix transform = Matrix::identity();
    Matrix inv_transform = Matrix::identity();

  protected:
    void flatten_common(rt_shape &out, int kind) const {
        out = rt_shape{};
        out.kind = kind;
        transform.store(out.transform);
        inv_transform.store(out.inv_transform);
    }
    int32_t id_ = 0;
};

class Sphere : public Renderable { // src/scene/sphere.rs
  public:
    explicit Sphere(MaterialRef material) : material_(std::move(material)), name_("Sphere") {}
    static Sphere new_with_name(const std::string &name, MaterialRef material) {
        Sphere s(std::move(material));
        s.name_ = name;
        return s;
    }
    MaterialRef get_material() const override { return material_; }
    std::string get_name() const override { return name_; }
    void flatten(rt_shape &out) const override { flatten_common(out, RT_SHAPE_SPHERE); }

  private:
    MaterialRef material_;
    std::string name_;
};

class Plane : public Renderable { // src/scene/plane.rs
  public:
    Plane(const Point3 &origin, const Vector3 &normal, MaterialRef material)
        : origin_(origin), normal_(normal), material_(std::move(material)) {
        // texture axes (plane.rs:23-30)
        Vector3 w = (normal.cross(Vector3(1.f, 0.f, 0.f)).len() <= F32_EPSILON) ? Vector3(0.f, 1.f, 0.f)
                                                                                  : Vector3(1.f, 0.f, 0.f);
        u_ = normal.cross(w).norm();
        v_ = normal.cross(u_).norm();
    }
    MaterialRef get_material() const override { return material_; }
    std::string get_name() const override { return "Plane"; }
    const Vector3 &u() const { return u_; }
    const Vector3 &v() const { return v_; }
    void flatten(rt_shape &out) const override {
        flatten_common(out, RT_SHAPE_PLANE);
        const float p[12] = {origin_.x, origin_.y, origin_.z, normal_.x, normal_.y, normal_.z,
                             u_.x, u_.y, u_.z, v_.x, v_.y, v_.z};
        for (int k = 0; k < 12; ++k) out.p[k] = p[k];
    }

  private:
    Point3 origin_;
    Vector3 normal_, u_, v_;
    MaterialRef material_;
};

class Triangle : public Renderable { // src/scene/triangle.rs
  public:
    Triangle(const Point3 &v0, const Point3 &v1, const Point3 &v2, MaterialRef material)
        : material_(std::move(material)) {
        verts_[0] = v0; verts_[1] = v1; verts_[2] = v2;
        normal_ = (v1 - v0).cross(v2 - v1).norm(); // triangle.rs:25-29
    }
    MaterialRef get_material() const override { return material_; }
    std::string get_name() const override { return "Triable"; } // sic, triangle.rs:113
    const Vector3 &normal() const { return normal_; }
    void flatten(rt_shape &out) const override {
        flatten_common(out, RT_SHAPE_TRIANGLE);
        for (int k = 0; k < 3; ++k) {
            out.p[3 * k] = verts_[k].x; out.p[3 * k + 1] = verts_[k].y; out.p[3 * k + 2] = verts_[k].z;
        }
        out.p[9] = normal_.x; out.p[10] = normal_.y; out.p[11] = normal_.z;
    }

  private:
    Point3 verts_[3];
    Vector3 normal_;
    MaterialRef material_;
};

class Cube : public Renderable { // src/scene/cube.rs: 12 fixed triangles sharing one material
  public:
    explicit Cube(MaterialRef material) : material_(std::move(material)) {}
    MaterialRef get_material() const override { return nullptr; } // cube.rs:120-126
    MaterialRef flatten_material() const override { return material_; }
    std::string get_name() const override { return "Cube"; }
    size_t size() const override { return 12; }
    void flatten(rt_shape &out) const override { flatten_common(out, RT_SHAPE_CUBE); }

  private:
    MaterialRef material_;
};

class LightSource { // src/scene/mod.rs:169-172
  public:
    virtual ~LightSource() = default;
    virtual std::string to_string() const = 0;
    virtual void flatten(rt_light &out) const = 0;
};

class PointLight : public LightSource { // src/scene/mod.rs:178-221
  public:
    PointLight(const Point3 &pos, const Color &color) : pos_(pos), color_(color) {}
    std::string to_string() const override { return "PointLight"; }
    void flatten(rt_light &out) const override {
        out.kind = RT_LIGHT_POINT;
        out.pos[0] = pos_.x; out.pos[1] = pos_.y; out.pos[2] = pos_.z;
        out.color[0] = color_.r; out.color[1] = color_.g; out.color[2] = color_.b;
    }

  private:
    Point3 pos_;
    Color color_;
};

class AmbientLight : public LightSource { // src/scene/mod.rs:223-246
  public:
    explicit AmbientLight(const Color &c) : color_(c) {}
    std::string to_string() const override { return "AmbientLight"; }
    void flatten(rt_light &out) const override {
        out.kind = RT_LIGHT_AMBIENT;
        out.pos[0] = out.pos[1] = out.pos[2] = 0.f;
        out.color[0] = color_.r; out.color[1] = color_.g; out.color[2] = color_.b;
    }

  private:
    Color color_;
};

// The flat description plus the storage behind its pointers.
struct FlatScene {
    std::vector<rt_shape> shapes;
    std::vector<rt_material> materials;
    std::vector<rt_light> lights;
    rt_scene_desc desc{};
    void bind() {
        desc.shapes = shapes.data(); desc.n_shapes = (int32_t)shapes.size();
        desc.materials = materials.data(); desc.n_materials = (int32_t)materials.size();
        desc.lights = lights.data(); desc.n_lights = (int32_t)lights.size();
    }
};

// A heap FlatScene whose first member is reachable from the rt_scene_desc pointer: what
// rt_builtin_scene / rt_synthetic_scene / rt_scene_desc_load hand out and rt_scene_desc_free takes.
struct OwnedDesc {
    rt_scene_desc desc; // must stay first
    FlatScene flat;
};

class Scene { // src/scene/mod.rs:22-85
  public:
    void add_shape(std::unique_ptr<Renderable> shape) { // id = index (mod.rs:40-44)
        shape->set_id((int32_t)shapes_.size());
        shapes_.push_back(std::move(shape));
    }
    template <typename T> void add_shape(T shape) { add_shape(std::unique_ptr<Renderable>(new T(std::move(shape)))); }
    void add_light(std::unique_ptr<LightSource> light) { lights_.push_back(std::move(light)); }
    template <typename T> void add_light(T light) { add_light(std::unique_ptr<LightSource>(new T(std::move(light)))); }
    void set_ambient(const Color &c) { ambient_ = c; }
    const Color &ambient() const { return ambient_; }
    const std::vector<std::unique_ptr<Renderable>> &shapes() const { return shapes_; }
    const std::vector<std::unique_ptr<LightSource>> &lights() const { return lights_; }
    Renderable *find_shape_mut(const std::string &name) {
        for (auto &s : shapes_)
            if (s->get_name() == name) return s.get();
        return nullptr;
    }
    const Renderable *find_shape(const std::string &name) const {
        for (auto &s : shapes_)
            if (s->get_name() == name) return s.get();
        return nullptr;
    }
    size_t size() const { // mod.rs:122-124
        size_t n = 0;
        for (auto &s : shapes_) n += s->size();
        return n;
    }
    // Materials are shared by identity (Rc): one table entry per distinct Material object,
    // in order of first use, so a later set_diffuse maps to one rt_material.
    void flatten(FlatScene &out) const {
        out = FlatScene();
        std::map<const Material *, int32_t> index;
        for (auto &s : shapes_) {
            rt_shape fs;
            s->flatten(fs);
            MaterialRef m = s->flatten_material();
            auto it = index.find(m.get());
            if (it == index.end()) {
                it = index.emplace(m.get(), (int32_t)out.materials.size()).first;
                rt_material fm{};
                m->flatten(fm);
                out.materials.push_back(fm);
            }
            fs.material = it->second;
            out.shapes.push_back(fs);
        }
        for (auto &l : lights_) {
            rt_light fl{};
            l->flatten(fl);
            out.lights.push_back(fl);
        }
        out.desc.ambient[0] = ambient_.r; out.desc.ambient[1] = ambient_.g; out.desc.ambient[2] = ambient_.b;
        out.bind();
    }

  private:
    Color ambient_ = colors::BLACK;
    std::vector<std::unique_ptr<LightSource>> lights_;
    std::vector<std::unique_ptr<Renderable>> shapes_;
};

} // namespace rust_tracer
