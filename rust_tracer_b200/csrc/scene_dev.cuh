// scene_dev.cuh -- device-resident scene tables (structure of arrays, 16-byte records).
//
// Layout in HBM (all immutable during a render except `materials`, which
// rt_scene_update_materials may replace between calls):
//   bound[j]       float4 (cx, cy, cz, R^2)   conservative bounding sphere of the j-th
//                                             bounded shape (sphere / triangle / cube),
//                                             sorted by kind so candidate batches are
//                                             kind-coherent; the ONLY table the sweep's
//                                             inner loop reads (staged through smem)
//   bound_shape[j] int32                      shape id of bound j
//   shape_meta[s]  int4  (kind, material, plane index, 0)
//   geom[3*s..]    3 x float4                 sphere/cube/plane: rows 0..2 of inv_transform
//                                             triangle: v0.xyz v1.x | v1.yz v2.xy | v2.z n.xyz
//   plane_data[7*p..] 7 x float4              origin, normal, u axis, v axis, then rows 0..2 of the
//                                             forward transform (xyz used, w unused)
//   cube_tris[36]  float4                     the 12 unit-cube triangles, same packing as geom
//   materials[m]   DevMaterial (64 B)
//   lights[l]      DevLight (32 B)
#pragma once
#include "exact.cuh"

namespace rt {

enum : int { KIND_SPHERE = 0, KIND_PLANE = 1, KIND_TRIANGLE = 2, KIND_CUBE = 3 };

struct DevColorFn {
    int kind; // 0 const, 1 checkerboard
    float r, g, b;
};

struct DevMaterial { // 64 bytes
    DevColorFn ambient, diffuse, specular;
    float power, reflectivity, refraction_index;
    int kind;
};

struct DevLight { // 32 bytes
    float px, py, pz;
    int kind; // 0 point, 1 ambient
    float cr, cg, cb;
    int point_index; // index among the point lights (shadow slot), -1 for ambient lights
};

struct DevScene {
    int n_shapes;
    int n_bounded;
    int n_planes;
    int n_lights;
    int n_point_lights;
    int first_nan_shape; // first sphere/triangle/cube in id order, -1 if none (NaN rays)
    float amb_r, amb_g, amb_b;
    const float4 *bound;
    const int *bound_shape;
    const int4 *shape_meta; // (kind, material, plane index or -1, 0)
    const float4 *geom;
    const int *plane_shape;
    const float4 *plane_data;
    const float4 *cube_tris;
    const DevMaterial *materials;
    const DevLight *lights;
};

} // namespace rt
