/*
 * rt_b200_host.h -- host-side helpers exported by librt_b200.so next to the render ABI:
 * the built-in scenes as flat rt_scene_desc values, the scene description file and the
 * reference's matrix helpers.  Pure host code (no CUDA call), so these work on a machine
 * without a GPU.
 *
 *   "default"  the reference's one scene, src/my_scene.rs:45-120, value for value
 *   "s1k"      1,000 spheres + checkerboard floor, 4 point lights   (BASELINE config 3)
 *   "m10k"     5,000 spheres + 5,000 triangles + floor, 4 lights    (BASELINE configs 4, 5)
 * The synthetic scenes have no counterpart in the reference (its scene is compiled in);
 * their generator is specified in SURVEY.md 8(d) and DESIGN.md.
 */
#ifndef RT_B200_HOST_H
#define RT_B200_HOST_H

#include "rt_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* name: "default" | "s1k" | "m10k".  seed 0 = the scene's canonical seed (1000 / 10000). */
int rt_builtin_scene(const char *name, uint64_t seed, rt_scene_desc **out);
/* n_spheres spheres with radius in [r_lo, r_hi] + n_triangles triangles + the floor. */
int rt_synthetic_scene(int32_t n_spheres, int32_t n_triangles, uint64_t seed, float r_lo, float r_hi,
                       rt_scene_desc **out);
void rt_scene_desc_free(rt_scene_desc *desc);

/* Scene description file.  The reference has no scene format -- its one scene is
 * Rust code, src/my_scene.rs:45-120, built through Scene::add_shape / add_light / set_ambient
 * (src/scene/mod.rs:40-56) -- so this is what stands in for "edit my_scene.rs and recompile":
 * the flat rt_scene_desc written out as it is (64-byte little-endian header with record sizes,
 * counts, ambient and an FNV-1a checksum, then the rt_shape / rt_material / rt_light arrays;
 * layout in csrc/host/scene_file.cpp).  f32 values are stored as bit patterns, so the oracle,
 * the GPU library and any other host see identical matrices.
 * rt_scene_desc_load validates magic, version, record sizes, file length, checksum, kinds and
 * material indices (RT_ERR_INVALID / RT_ERR_UNSUPPORTED + rt_last_error()); the result is
 * released with rt_scene_desc_free.                                                        */
int rt_scene_desc_save(const rt_scene_desc *desc, const char *path);
int rt_scene_desc_load(const char *path, rt_scene_desc **out);

/* Matrix helpers of the host API (reference src/math/matrix.rs), for bindings that build
 * their own scenes: out = inverse(m) with the reference's Gauss-Jordan; returns
 * RT_ERR_INVALID where the reference panics "Singular Matrix".                            */
int rt_matrix_inverse(const float *m16, float *out16);
/* out = T(tx,ty,tz) * Rz(rz) * Ry(ry) * Rx(rx) * S(sx,sy,sz), angles in degrees. */
void rt_matrix_compose(float tx, float ty, float tz, float rx, float ry, float rz, float sx, float sy, float sz,
                       float *out16);
/* Plane::new's texture axes and Triangle::new's normal, as the reference computes them. */
void rt_plane_axes(const float *normal3, float *u3, float *v3);
void rt_triangle_normal(const float *v0, const float *v1, const float *v2, float *n3);

#ifdef __cplusplus
}
#endif
#endif
