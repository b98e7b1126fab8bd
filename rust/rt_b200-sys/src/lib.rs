//! Raw bindings to `librt_b200.so`: one item per declaration of `include/rt_b200.h` and
//! `include/rt_b200_host.h` (ABI version 3).  Source only: this repository has no Rust
//! toolchain; `tests/test_rust_binding_cpu.py` checks this file against the C headers.
//!
//! Which reference call each entry point replaces is cited in the headers
//! (e.g. `rt_render` <- `render::render`, src/render.rs:31-38).
#![allow(non_camel_case_types)]

use std::os::raw::{c_char, c_int, c_void};

pub const RT_B200_ABI_VERSION: c_int = 4;

// rt_status
pub const RT_OK: c_int = 0;
pub const RT_ERR_INVALID: c_int = -1;
pub const RT_ERR_UNSUPPORTED: c_int = -2;
pub const RT_ERR_CUDA: c_int = -3;
pub const RT_ERR_NOMEM: c_int = -4;
pub const RT_ERR_INTERNAL: c_int = -5;
pub const RT_ERR_TIMEOUT: c_int = -6;

// rt_shape_kind
pub const RT_SHAPE_SPHERE: i32 = 0;
pub const RT_SHAPE_PLANE: i32 = 1;
pub const RT_SHAPE_TRIANGLE: i32 = 2;
pub const RT_SHAPE_CUBE: i32 = 3;
// rt_colorfn_kind
pub const RT_COLOR_CONST: i32 = 0;
pub const RT_COLOR_CHECKERBOARD: i32 = 1;
// rt_material_kind
pub const RT_MATERIAL_PHONG: i32 = 0;
pub const RT_MATERIAL_TEXTURE_PHONG: i32 = 1;
// rt_light_kind
pub const RT_LIGHT_POINT: i32 = 0;
pub const RT_LIGHT_AMBIENT: i32 = 1;
// rt_render_flags
pub const RT_FLAG_NONE: i32 = 0;
pub const RT_FLAG_AUDIT_ALL_PAIRS: i32 = 1;
pub const RT_FLAG_FOREST_SHADING: i32 = 2;
pub const RT_FLAG_FRAME_ROWS: i32 = 4;

pub const RT_FRAME_HANDLE_BYTES: usize = 64;
pub const RT_FRAME_SIGNAL_WORDS: usize = 256;
pub const RT_FRAME_WORD_RELEASED: i32 = 255;

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rt_shape {
    pub kind: i32,
    pub material: i32,
    pub transform: [f32; 16],
    pub inv_transform: [f32; 16],
    pub p: [f32; 12],
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_colorfn {
    pub kind: i32,
    pub rgb: [f32; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_material {
    pub kind: i32,
    pub ambient: rt_colorfn,
    pub diffuse: rt_colorfn,
    pub specular: rt_colorfn,
    pub power: f32,
    pub reflectivity: f32,
    pub refraction_index: f32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_light {
    pub kind: i32,
    pub pos: [f32; 3],
    pub color: [f32; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rt_scene_desc {
    pub shapes: *const rt_shape,
    pub n_shapes: i32,
    pub materials: *const rt_material,
    pub n_materials: i32,
    pub lights: *const rt_light,
    pub n_lights: i32,
    pub ambient: [f32; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_camera {
    pub origin: [f32; 3],
    pub x_min: f32,
    pub x_max: f32,
    pub y_min: f32,
    pub y_max: f32,
    pub x_res: i32,
    pub y_res: i32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_render_params {
    pub depth: i32,
    pub spp: i32,
    pub seed: u64,
    pub rank: i32,
    pub world: i32,
    pub tile_rows: i32,
    pub flags: i32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_stats {
    pub pixels: u64,
    pub rays: u64,
    pub shadow_rays: u64,
    pub nodes: u64,
    pub sphere_tests: u64,
    pub plane_tests: u64,
    pub triangle_tests: u64,
    pub cube_tests: u64,
    pub exact_tests: u64,
    pub algorithmic_flops: f64,
    pub sweep_ms: f32,
    pub shade_ms: f32,
    pub total_ms: f32,
    pub kernel_launches: u32,
    pub filter_flops: f64,
    pub tested_pairs: u64,
    pub bound_evals_sphere: u64,
    pub bound_evals_triangle: u64,
    pub bound_evals_cube: u64,
    pub hier_evals: u64,
    pub exact_sphere_tests: u64,
    pub exact_triangle_tests: u64,
    pub exact_cube_tests: u64,
    pub exact_flops: f64,
    pub executed_flops: f64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_forest_stats {
    pub num_trees: u64,
    pub num_intersections: u64,
    pub smallest_tree: u64,
    pub largest_tree: u64,
    pub median: u64,
    pub p90: u64,
    pub p95: u64,
    pub p99: u64,
}

/// Opaque: device-resident scene tables.
#[repr(C)]
pub struct rt_scene {
    _private: [u8; 0],
}

/// Opaque: one scene resident on several devices, rendered by one call.
#[repr(C)]
pub struct rt_multi {
    _private: [u8; 0],
}

/// Opaque: device-resident ray forest.
#[repr(C)]
pub struct rt_forest {
    _private: [u8; 0],
}

extern "C" {
    // ---- library
    pub fn rt_abi_version() -> c_int;
    pub fn rt_last_error() -> *const c_char;
    pub fn rt_device_count() -> c_int;
    pub fn rt_camera_default(x_res: i32, y_res: i32, out: *mut rt_camera);
    pub fn rt_tile_rows_for_rank(height: i32, rank: i32, world: i32, tile_rows: i32, rows: *mut i32, cap: i32) -> i32;

    // ---- scene
    pub fn rt_scene_create(desc: *const rt_scene_desc, device: i32, out: *mut *mut rt_scene) -> c_int;
    pub fn rt_scene_update_materials(scene: *mut rt_scene, materials: *const rt_material, n: i32) -> c_int;
    pub fn rt_scene_set_stream(scene: *mut rt_scene, stream: *mut c_void) -> c_int;
    pub fn rt_scene_destroy(scene: *mut rt_scene);
    pub fn rt_trim(device: i32) -> c_int;

    // ---- Basic renderer
    pub fn rt_render(
        scene: *mut rt_scene,
        camera: *const rt_camera,
        params: *const rt_render_params,
        rgb_f32: *mut f32,
        rgb_u8: *mut u8,
        stats: *mut rt_stats,
    ) -> c_int;
    pub fn rt_render_device(
        scene: *mut rt_scene,
        camera: *const rt_camera,
        params: *const rt_render_params,
        d_rgb_f32: *mut c_void,
        d_rgb_u8: *mut c_void,
        stats: *mut rt_stats,
    ) -> c_int;

    // ---- one host call, several GPUs
    pub fn rt_multi_create(desc: *const rt_scene_desc, devices: *const i32, n_devices: i32, out: *mut *mut rt_multi) -> c_int;
    pub fn rt_multi_device_count(multi: *const rt_multi) -> i32;
    pub fn rt_multi_update_materials(multi: *mut rt_multi, materials: *const rt_material, n: i32) -> c_int;
    pub fn rt_multi_render(
        multi: *mut rt_multi,
        camera: *const rt_camera,
        params: *const rt_render_params,
        rgb_f32: *mut f32,
        rgb_u8: *mut u8,
        stats: *mut rt_stats,
    ) -> c_int;
    pub fn rt_multi_render_device(
        multi: *mut rt_multi,
        camera: *const rt_camera,
        params: *const rt_render_params,
        d_rgb_f32: *mut c_void,
        d_rgb_u8: *mut c_void,
        stats: *mut rt_stats,
    ) -> c_int;
    pub fn rt_multi_destroy(multi: *mut rt_multi);
    pub fn rt_render_multi(
        desc: *const rt_scene_desc,
        devices: *const i32,
        n_devices: i32,
        camera: *const rt_camera,
        params: *const rt_render_params,
        rgb_f32: *mut f32,
        rgb_u8: *mut u8,
        stats: *mut rt_stats,
    ) -> c_int;

    // ---- multi-GPU frame
    pub fn rt_frame_alloc(device: i32, bytes: usize, d_ptr: *mut *mut c_void) -> c_int;
    pub fn rt_frame_export(device: i32, d_ptr: *mut c_void, handle: *mut u8) -> c_int;
    pub fn rt_frame_import(device: i32, handle: *const u8, d_ptr: *mut *mut c_void) -> c_int;
    pub fn rt_frame_release(device: i32, d_ptr: *mut c_void, imported: i32) -> c_int;
    pub fn rt_frame_signal(device: i32, d_ptr: *mut c_void, bytes: usize, word: i32, value: u32, stream: *mut c_void) -> c_int;
    pub fn rt_frame_wait(
        device: i32,
        d_ptr: *mut c_void,
        bytes: usize,
        first_word: i32,
        n_words: i32,
        value: u32,
        timeout_ms: u32,
        stream: *mut c_void,
    ) -> c_int;

    // ---- RayForest renderer
    pub fn rt_forest_build(
        scene: *mut rt_scene,
        camera: *const rt_camera,
        params: *const rt_render_params,
        out: *mut *mut rt_forest,
        stats: *mut rt_stats,
    ) -> c_int;
    pub fn rt_forest_eval(
        forest: *mut rt_forest,
        ambient: *const f32,
        rgb_f32: *mut f32,
        rgb_u8: *mut u8,
        stats: *mut rt_stats,
    ) -> c_int;
    pub fn rt_forest_eval_filtered(
        forest: *mut rt_forest,
        ambient: *const f32,
        mutated_ids: *const i32,
        n_ids: i32,
        rgb_f32_inout: *mut f32,
        rgb_u8_inout: *mut u8,
        n_evaluated: *mut u64,
        stats: *mut rt_stats,
    ) -> c_int;
    pub fn rt_forest_get_stats(forest: *mut rt_forest, out: *mut rt_forest_stats) -> c_int;
    pub fn rt_forest_size(forest: *mut rt_forest, out: *mut u64) -> c_int;
    pub fn rt_forest_trees_with(forest: *mut rt_forest, shape_id: i32, out: *mut u64) -> c_int;
    /// Not in the reference: `rt_forest_trees_with` on static device scratch, no allocation
    /// (at most `RT_STATIC_TREES_MAX` = 4 Mi trees).
    pub fn rt_forest_trees_with_static(forest: *mut rt_forest, shape_id: i32, out: *mut u64) -> c_int;
    pub fn rt_forest_destroy(forest: *mut rt_forest);

    // ---- measurement
    pub fn rt_measure_fp32_peak(device: i32, tflops_measured: *mut f64, sm_count: *mut i32, sm_clock_khz: *mut i32) -> c_int;

    // ---- host helpers (rt_b200_host.h): no CUDA call, usable without a GPU
    pub fn rt_builtin_scene(name: *const c_char, seed: u64, out: *mut *mut rt_scene_desc) -> c_int;
    pub fn rt_synthetic_scene(n_spheres: i32, n_triangles: i32, seed: u64, r_lo: f32, r_hi: f32, out: *mut *mut rt_scene_desc) -> c_int;
    pub fn rt_scene_desc_free(desc: *mut rt_scene_desc);
    pub fn rt_scene_desc_save(desc: *const rt_scene_desc, path: *const c_char) -> c_int;
    pub fn rt_scene_desc_load(path: *const c_char, out: *mut *mut rt_scene_desc) -> c_int;
    pub fn rt_matrix_inverse(m16: *const f32, out16: *mut f32) -> c_int;
    pub fn rt_matrix_compose(tx: f32, ty: f32, tz: f32, rx: f32, ry: f32, rz: f32, sx: f32, sy: f32, sz: f32, out16: *mut f32);
    pub fn rt_plane_axes(normal3: *const f32, u3: *mut f32, v3: *mut f32);
    pub fn rt_triangle_normal(v0: *const f32, v1: *const f32, v2: *const f32, n3: *mut f32);
}
