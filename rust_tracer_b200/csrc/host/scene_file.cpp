// scene_file.cpp -- the scene description file: rt_scene_desc_save / rt_scene_desc_load.
//
// The reference has no scene format: its one scene is Rust code (src/my_scene.rs:45-120) and
// Scene/Renderable expose no accessors (src/scene/mod.rs:147-165).  The file is therefore the
// flat rt_scene_desc of include/rt_b200.h written out as it is, so that the oracle, the GPU
// driver and a Rust host exchange scenes whose f32 values (host-computed inverse matrices
// included) are bit-identical on every side.
//
//   offset  size  field                       (little-endian; f32 = IEEE-754 bit patterns)
//        0     8  magic "RTSCENE\0"
//        8     4  u32 format version (1)
//       12     4  u32 RT_B200_ABI_VERSION of the writer (informative)
//       16    12  u32 sizeof(rt_shape), sizeof(rt_material), sizeof(rt_light)   (184, 64, 28)
//       28    12  i32 n_shapes, n_materials, n_lights
//       40    12  f32 ambient[3]                                  Scene::ambient, src/scene/mod.rs:50-56
//       52     4  u32 0
//       56     8  u64 FNV-1a 64 of the payload
//       64     .  payload: rt_shape[n_shapes], rt_material[n_materials], rt_light[n_lights]
//
// Pure host code: works without a GPU.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <new>
#include <vector>

#include "../../../include/rt_b200_host.h"
#include "rt_scene.hpp"

namespace rt {
int fail(int code, const char *fmt, ...); // api_scene.cu: sets rt_last_error()
}

namespace {

using rust_tracer::OwnedDesc;

constexpr char MAGIC[8] = {'R', 'T', 'S', 'C', 'E', 'N', 'E', '\0'};
constexpr uint32_t FORMAT_VERSION = 1;
constexpr size_t HEADER_BYTES = 64;

static_assert(sizeof(rt_shape) == 184 && sizeof(rt_material) == 64 && sizeof(rt_light) == 28,
              "record sizes are part of the file format");

struct Header {
    char magic[8];
    uint32_t version, abi;
    uint32_t shape_bytes, material_bytes, light_bytes;
    int32_t n_shapes, n_materials, n_lights;
    float ambient[3];
    uint32_t zero;
    uint64_t checksum;
};
static_assert(sizeof(Header) == HEADER_BYTES, "header layout");

uint64_t fnv1a(uint64_t h, const void *data, size_t n) {
    const unsigned char *p = static_cast<const unsigned char *>(data);
    for (size_t i = 0; i < n; ++i) h = (h ^ p[i]) * 0x100000001B3ull;
    return h;
}
constexpr uint64_t FNV_BASIS = 0xCBF29CE484222325ull;

bool little_endian() {
    const uint32_t one = 1;
    unsigned char b;
    std::memcpy(&b, &one, 1);
    return b == 1;
}

// what rt_scene_create rejects as malformed, checked here so that a bad file fails at load
int check_records(const rt_shape *shapes, int ns, const rt_material *mats, int nm, const rt_light *lights, int nl) {
    for (int i = 0; i < ns; ++i) {
        if (shapes[i].kind < RT_SHAPE_SPHERE || shapes[i].kind > RT_SHAPE_CUBE)
            return rt::fail(RT_ERR_INVALID, "shape %d: unknown kind %d", i, shapes[i].kind);
        if (shapes[i].material < 0 || shapes[i].material >= nm)
            return rt::fail(RT_ERR_INVALID, "shape %d: material %d out of range", i, shapes[i].material);
    }
    for (int i = 0; i < nm; ++i) {
        if (mats[i].kind < RT_MATERIAL_PHONG || mats[i].kind > RT_MATERIAL_TEXTURE_PHONG)
            return rt::fail(RT_ERR_INVALID, "material %d: unknown kind %d", i, mats[i].kind);
        const rt_colorfn *fns[3] = {&mats[i].ambient, &mats[i].diffuse, &mats[i].specular};
        for (const rt_colorfn *f : fns)
            if (f->kind < RT_COLOR_CONST || f->kind > RT_COLOR_CHECKERBOARD)
                return rt::fail(RT_ERR_UNSUPPORTED, "material %d: texture function %d is not in rt_colorfn_kind", i, f->kind);
    }
    for (int i = 0; i < nl; ++i)
        if (lights[i].kind < RT_LIGHT_POINT || lights[i].kind > RT_LIGHT_AMBIENT)
            return rt::fail(RT_ERR_INVALID, "light %d: unknown kind %d", i, lights[i].kind);
    return RT_OK;
}

struct File {
    std::FILE *f = nullptr;
    ~File() { if (f) std::fclose(f); }
};

} // namespace

extern "C" {

int rt_scene_desc_save(const rt_scene_desc *desc, const char *path) {
    if (!desc || !path) return rt::fail(RT_ERR_INVALID, "null argument");
    if (!little_endian()) return rt::fail(RT_ERR_UNSUPPORTED, "scene files are little-endian");
    if (desc->n_shapes < 0 || desc->n_materials < 0 || desc->n_lights < 0 || (desc->n_shapes && !desc->shapes) ||
        (desc->n_materials && !desc->materials) || (desc->n_lights && !desc->lights))
        return rt::fail(RT_ERR_INVALID, "malformed scene description");
    int rc = check_records(desc->shapes, desc->n_shapes, desc->materials, desc->n_materials, desc->lights, desc->n_lights);
    if (rc != RT_OK) return rc;

    Header h{};
    std::memcpy(h.magic, MAGIC, 8);
    h.version = FORMAT_VERSION;
    h.abi = RT_B200_ABI_VERSION;
    h.shape_bytes = sizeof(rt_shape); h.material_bytes = sizeof(rt_material); h.light_bytes = sizeof(rt_light);
    h.n_shapes = desc->n_shapes; h.n_materials = desc->n_materials; h.n_lights = desc->n_lights;
    std::memcpy(h.ambient, desc->ambient, sizeof(h.ambient));
    const size_t sb = sizeof(rt_shape) * (size_t)desc->n_shapes, mb = sizeof(rt_material) * (size_t)desc->n_materials,
                 lb = sizeof(rt_light) * (size_t)desc->n_lights;
    uint64_t sum = FNV_BASIS;
    if (sb) sum = fnv1a(sum, desc->shapes, sb);
    if (mb) sum = fnv1a(sum, desc->materials, mb);
    if (lb) sum = fnv1a(sum, desc->lights, lb);
    h.checksum = sum;

    File out;
    out.f = std::fopen(path, "wb");
    if (!out.f) return rt::fail(RT_ERR_INVALID, "cannot open %s for writing", path);
    bool ok = std::fwrite(&h, sizeof(h), 1, out.f) == 1;
    ok = ok && (!sb || std::fwrite(desc->shapes, sb, 1, out.f) == 1);
    ok = ok && (!mb || std::fwrite(desc->materials, mb, 1, out.f) == 1);
    ok = ok && (!lb || std::fwrite(desc->lights, lb, 1, out.f) == 1);
    ok = ok && std::fflush(out.f) == 0;
    if (!ok) return rt::fail(RT_ERR_INVALID, "short write to %s", path);
    return RT_OK;
}

int rt_scene_desc_load(const char *path, rt_scene_desc **out) {
    if (!path || !out) return rt::fail(RT_ERR_INVALID, "null argument");
    *out = nullptr;
    if (!little_endian()) return rt::fail(RT_ERR_UNSUPPORTED, "scene files are little-endian");
    File in;
    in.f = std::fopen(path, "rb");
    if (!in.f) return rt::fail(RT_ERR_INVALID, "cannot open %s", path);
    Header h{};
    if (std::fread(&h, sizeof(h), 1, in.f) != 1) return rt::fail(RT_ERR_INVALID, "%s: shorter than a scene header", path);
    if (std::memcmp(h.magic, MAGIC, 8) != 0) return rt::fail(RT_ERR_INVALID, "%s: not a scene file (bad magic)", path);
    if (h.version != FORMAT_VERSION) return rt::fail(RT_ERR_UNSUPPORTED, "%s: format version %u (this library reads %u)", path, h.version, FORMAT_VERSION);
    if (h.shape_bytes != sizeof(rt_shape) || h.material_bytes != sizeof(rt_material) || h.light_bytes != sizeof(rt_light))
        return rt::fail(RT_ERR_UNSUPPORTED, "%s: record sizes %u/%u/%u differ from this ABI's %zu/%zu/%zu", path, h.shape_bytes,
                        h.material_bytes, h.light_bytes, sizeof(rt_shape), sizeof(rt_material), sizeof(rt_light));
    if (h.n_shapes < 0 || h.n_materials < 0 || h.n_lights < 0) return rt::fail(RT_ERR_INVALID, "%s: negative record count", path);
    if (h.n_shapes >= (1 << 24)) return rt::fail(RT_ERR_UNSUPPORTED, "more than 2^24 shapes");

    // the counts must account for the file to the byte before anything is allocated from them
    const uint64_t want = HEADER_BYTES + (uint64_t)h.n_shapes * sizeof(rt_shape) + (uint64_t)h.n_materials * sizeof(rt_material) +
                          (uint64_t)h.n_lights * sizeof(rt_light);
    if (std::fseek(in.f, 0, SEEK_END) != 0) return rt::fail(RT_ERR_INVALID, "%s: not seekable", path);
    const long size = std::ftell(in.f);
    if (size < 0 || (uint64_t)size != want)
        return rt::fail(RT_ERR_INVALID, "%s: %ld bytes, header describes %llu", path, size, (unsigned long long)want);
    if (std::fseek(in.f, (long)HEADER_BYTES, SEEK_SET) != 0) return rt::fail(RT_ERR_INVALID, "%s: not seekable", path);

    OwnedDesc *o = new (std::nothrow) OwnedDesc();
    if (!o) return rt::fail(RT_ERR_NOMEM, "out of host memory");
    int rc = RT_OK;
    try {
        o->flat.shapes.resize((size_t)h.n_shapes);
        o->flat.materials.resize((size_t)h.n_materials);
        o->flat.lights.resize((size_t)h.n_lights);
    } catch (const std::bad_alloc &) {
        rc = rt::fail(RT_ERR_NOMEM, "out of host memory");
    }
    auto read_all = [&](void *dst, size_t bytes) { return !bytes || std::fread(dst, bytes, 1, in.f) == 1; };
    if (rc == RT_OK &&
        !(read_all(o->flat.shapes.data(), o->flat.shapes.size() * sizeof(rt_shape)) &&
          read_all(o->flat.materials.data(), o->flat.materials.size() * sizeof(rt_material)) &&
          read_all(o->flat.lights.data(), o->flat.lights.size() * sizeof(rt_light))))
        rc = rt::fail(RT_ERR_INVALID, "%s: short read", path);
    if (rc == RT_OK) {
        uint64_t sum = FNV_BASIS;
        sum = fnv1a(sum, o->flat.shapes.data(), o->flat.shapes.size() * sizeof(rt_shape));
        sum = fnv1a(sum, o->flat.materials.data(), o->flat.materials.size() * sizeof(rt_material));
        sum = fnv1a(sum, o->flat.lights.data(), o->flat.lights.size() * sizeof(rt_light));
        if (sum != h.checksum) rc = rt::fail(RT_ERR_INVALID, "%s: payload checksum mismatch (file corrupted)", path);
    }
    if (rc == RT_OK)
        rc = check_records(o->flat.shapes.data(), h.n_shapes, o->flat.materials.data(), h.n_materials, o->flat.lights.data(),
                           h.n_lights);
    if (rc != RT_OK) {
        delete o;
        return rc;
    }
    o->flat.bind();
    std::memcpy(o->flat.desc.ambient, h.ambient, sizeof(h.ambient));
    o->desc = o->flat.desc;
    *out = &o->desc;
    return RT_OK;
}

} // extern "C"
