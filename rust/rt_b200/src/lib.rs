//! Safe wrapper over `rt_b200-sys`.  Source only (no Rust toolchain in this repository).
//!
//! One type per handle of the C ABI, `Result` instead of status codes, and the two layout
//! conversions the reference needs: `FlatScene` (what `Scene::flatten` fills, see
//! `reference_patch/gpu.rs`) and the column-major `RenderBuffer` (src/render.rs:5-19).
//! Nothing here computes pixels; without a CUDA device every constructor returns
//! `Err(Error { code: RT_ERR_CUDA, .. })` -- there is no CPU fallback to fall back to.

use rt_b200_sys as sys;
use std::ffi::{CStr, CString};
use std::os::raw::c_int;
use std::path::Path;
use std::ptr;

pub use sys::{rt_camera, rt_colorfn, rt_forest_stats, rt_light, rt_material, rt_render_params, rt_shape, rt_stats};

/// A failed call: the `rt_status` and `rt_last_error()` of the calling thread.
#[derive(Debug, Clone)]
pub struct Error {
    pub code: c_int,
    pub message: String,
}

impl std::fmt::Display for Error {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result {
        write!(f, "librt_b200 error {}: {}", self.code, self.message)
    }
}
impl std::error::Error for Error {}

pub type Result<T> = std::result::Result<T, Error>;

fn check(rc: c_int) -> Result<()> {
    if rc >= 0 {
        return Ok(());
    }
    let message = unsafe { CStr::from_ptr(sys::rt_last_error()) }.to_string_lossy().into_owned();
    Err(Error { code: rc, message })
}

fn c_path(path: &Path) -> Result<CString> {
    CString::new(path.to_string_lossy().as_bytes())
        .map_err(|_| Error { code: sys::RT_ERR_INVALID, message: "path contains a NUL byte".into() })
}

/// The flat scene description and the storage behind its pointers (`rt_scene_desc`).
/// Shapes are in `Scene::add_shape` order: the index is the shape id and the tie-break order of
/// `Scene::intersect` (src/scene/mod.rs:40-44, 98-116).
#[derive(Clone, Default)]
pub struct FlatScene {
    pub shapes: Vec<rt_shape>,
    pub materials: Vec<rt_material>,
    pub lights: Vec<rt_light>,
    pub ambient: [f32; 3],
}

impl FlatScene {
    /// Borrowed view for the FFI; valid while `self` is neither moved nor mutated.
    pub fn as_desc(&self) -> sys::rt_scene_desc {
        sys::rt_scene_desc {
            shapes: self.shapes.as_ptr(),
            n_shapes: self.shapes.len() as i32,
            materials: self.materials.as_ptr(),
            n_materials: self.materials.len() as i32,
            lights: self.lights.as_ptr(),
            n_lights: self.lights.len() as i32,
            ambient: self.ambient,
        }
    }

    unsafe fn from_owned_desc(p: *mut sys::rt_scene_desc) -> FlatScene {
        // an empty array may come with a null pointer, which from_raw_parts does not accept
        unsafe fn copy<T: Copy>(ptr: *const T, n: i32) -> Vec<T> {
            if ptr.is_null() || n <= 0 { Vec::new() } else { std::slice::from_raw_parts(ptr, n as usize).to_vec() }
        }
        let d = &*p;
        let flat = FlatScene {
            shapes: copy(d.shapes, d.n_shapes),
            materials: copy(d.materials, d.n_materials),
            lights: copy(d.lights, d.n_lights),
            ambient: d.ambient,
        };
        sys::rt_scene_desc_free(p);
        flat
    }

    /// "default" = src/my_scene.rs value for value; "s1k", "m10k" = the synthetic benchmark scenes.
    pub fn builtin(name: &str, seed: u64) -> Result<FlatScene> {
        let name = CString::new(name).map_err(|_| Error { code: sys::RT_ERR_INVALID, message: "bad scene name".into() })?;
        let mut p = ptr::null_mut();
        check(unsafe { sys::rt_builtin_scene(name.as_ptr(), seed, &mut p) })?;
        Ok(unsafe { FlatScene::from_owned_desc(p) })
    }

    /// Scene description file (include/rt_b200_host.h).
    pub fn load(path: &Path) -> Result<FlatScene> {
        let path = c_path(path)?;
        let mut p = ptr::null_mut();
        check(unsafe { sys::rt_scene_desc_load(path.as_ptr(), &mut p) })?;
        Ok(unsafe { FlatScene::from_owned_desc(p) })
    }

    pub fn save(&self, path: &Path) -> Result<()> {
        let path = c_path(path)?;
        let desc = self.as_desc();
        check(unsafe { sys::rt_scene_desc_save(&desc, path.as_ptr()) })
    }
}

/// `Matrix::inverse` with the reference's Gauss-Jordan (src/math/matrix.rs:88-142); `Err` where
/// the reference panics "Singular Matrix".
pub fn matrix_inverse(m: &[f32; 16]) -> Result<[f32; 16]> {
    let mut out = [0f32; 16];
    check(unsafe { sys::rt_matrix_inverse(m.as_ptr(), out.as_mut_ptr()) })?;
    Ok(out)
}

/// `Camera::new(x_res, y_res)` (src/render.rs:166-176).
pub fn camera(x_res: usize, y_res: usize) -> rt_camera {
    let mut cam = rt_camera::default();
    unsafe { sys::rt_camera_default(x_res as i32, y_res as i32, &mut cam) };
    cam
}

fn params(depth: usize, flags: i32) -> rt_render_params {
    rt_render_params { depth: depth as i32, spp: 1, seed: 0, rank: 0, world: 1, tile_rows: 8, flags }
}

/// Row-major RGB f32 image, 3 floats per pixel: what `rt_render` / `rt_forest_eval` fill.
pub struct Image {
    pub width: usize,
    pub height: usize,
    pub rgb: Vec<f32>,
}

impl Image {
    pub fn new(width: usize, height: usize) -> Image {
        Image { width, height, rgb: vec![0.0; width * height * 3] }
    }
    /// Pixel (u, v) = column u, row v (the reference's `buffer.buf[u][v]`).
    pub fn pixel(&self, u: usize, v: usize) -> [f32; 3] {
        let i = (v * self.width + u) * 3;
        [self.rgb[i], self.rgb[i + 1], self.rgb[i + 2]]
    }
}

/// Device-resident scene tables (`rt_scene`).
pub struct Scene {
    handle: *mut sys::rt_scene,
    ambient: [f32; 3],
}

impl Scene {
    /// Replaces `Scene::new` + `create_scene` for the device (src/main.rs:33-34).
    pub fn new(flat: &FlatScene, device: i32) -> Result<Scene> {
        let desc = flat.as_desc();
        let mut handle = ptr::null_mut();
        check(unsafe { sys::rt_scene_create(&desc, device, &mut handle) })?;
        Ok(Scene { handle, ambient: flat.ambient })
    }

    /// `Material::set_diffuse` after upload (src/gui.rs:221-235): the whole material table again.
    pub fn update_materials(&mut self, materials: &[rt_material]) -> Result<()> {
        check(unsafe { sys::rt_scene_update_materials(self.handle, materials.as_ptr(), materials.len() as i32) })
    }

    /// `render::render(&camera, &scene, &mut buffer, depth)` (src/render.rs:31-38) with
    /// `Camera::new(width, height)`.
    pub fn render(&mut self, width: usize, height: usize, depth: usize) -> Result<(Image, rt_stats)> {
        self.render_with(&camera(width, height), depth)
    }

    /// The same with the caller's camera: every field of the reference's `Camera` is public
    /// (src/render.rs:155-163) and `rt_camera` carries them all.
    pub fn render_with(&mut self, cam: &rt_camera, depth: usize) -> Result<(Image, rt_stats)> {
        let p = params(depth, sys::RT_FLAG_NONE);
        let mut img = Image::new(cam.x_res as usize, cam.y_res as usize);
        let mut stats = rt_stats::default();
        check(unsafe { sys::rt_render(self.handle, cam, &p, img.rgb.as_mut_ptr(), ptr::null_mut(), &mut stats) })?;
        Ok((img, stats))
    }

    /// Same, quantised with `Color::as_u8` (src/scene/color.rs:43-46): what bmp.rs writes.
    pub fn render_u8(&mut self, width: usize, height: usize, depth: usize) -> Result<Vec<u8>> {
        let cam = camera(width, height);
        let p = params(depth, sys::RT_FLAG_NONE);
        let mut out = vec![0u8; width * height * 3];
        check(unsafe { sys::rt_render(self.handle, &cam, &p, ptr::null_mut(), out.as_mut_ptr(), ptr::null_mut()) })?;
        Ok(out)
    }

    /// `render_tree::generate_ray_forest` (src/render_tree.rs:147-164).
    pub fn generate_ray_forest(&mut self, width: usize, height: usize, depth: usize) -> Result<Forest> {
        let cam = camera(width, height);
        let p = params(depth, sys::RT_FLAG_NONE);
        let mut handle = ptr::null_mut();
        check(unsafe { sys::rt_forest_build(self.handle, &cam, &p, &mut handle, ptr::null_mut()) })?;
        Ok(Forest { handle, width, height, ambient: self.ambient })
    }
}

impl Drop for Scene {
    fn drop(&mut self) {
        unsafe { sys::rt_scene_destroy(self.handle) }
    }
}

/// One scene resident on several GPUs (`rt_multi`): the reference's single blocking
/// `render::render` call (src/render.rs:31-38) for a host that owns 1-8 devices.  The library
/// drives every device from its own worker thread; the frame is split into interleaved row tiles.
pub struct MultiScene {
    handle: *mut sys::rt_multi,
}

impl MultiScene {
    pub fn new(flat: &FlatScene, devices: &[i32]) -> Result<MultiScene> {
        let desc = flat.as_desc();
        let mut handle = ptr::null_mut();
        check(unsafe { sys::rt_multi_create(&desc, devices.as_ptr(), devices.len() as i32, &mut handle) })?;
        Ok(MultiScene { handle })
    }

    pub fn devices(&self) -> usize {
        unsafe { sys::rt_multi_device_count(self.handle) as usize }
    }

    pub fn update_materials(&mut self, materials: &[rt_material]) -> Result<()> {
        check(unsafe { sys::rt_multi_update_materials(self.handle, materials.as_ptr(), materials.len() as i32) })
    }

    /// `render::render(&camera, &scene, &mut buffer, depth)` on all devices of the scene.
    pub fn render_with(&mut self, cam: &rt_camera, depth: usize) -> Result<(Image, rt_stats)> {
        let p = params(depth, sys::RT_FLAG_NONE);
        let mut img = Image::new(cam.x_res as usize, cam.y_res as usize);
        let mut stats = rt_stats::default();
        check(unsafe { sys::rt_multi_render(self.handle, cam, &p, img.rgb.as_mut_ptr(), ptr::null_mut(), &mut stats) })?;
        Ok((img, stats))
    }
}

impl Drop for MultiScene {
    fn drop(&mut self) {
        unsafe { sys::rt_multi_destroy(self.handle) }
    }
}

/// One-shot `rt_render_multi`: scene description in, image out, on the listed devices.
pub fn render_multi(flat: &FlatScene, devices: &[i32], cam: &rt_camera, depth: usize) -> Result<(Image, rt_stats)> {
    let desc = flat.as_desc();
    let p = params(depth, sys::RT_FLAG_NONE);
    let mut img = Image::new(cam.x_res as usize, cam.y_res as usize);
    let mut stats = rt_stats::default();
    check(unsafe {
        sys::rt_render_multi(&desc, devices.as_ptr(), devices.len() as i32, cam, &p, img.rgb.as_mut_ptr(), ptr::null_mut(), &mut stats)
    })?;
    Ok((img, stats))
}

/// Number of CUDA devices the library sees (`rt_device_count`).
pub fn device_count() -> Result<usize> {
    let n = unsafe { sys::rt_device_count() };
    check(n.min(0))?;
    Ok(n as usize)
}

/// Device-resident ray forest (`rt_forest`).  May be dropped before or after the `Scene` it was built
/// from: the library keeps the scene's tables alive until its last forest is gone.
pub struct Forest {
    handle: *mut sys::rt_forest,
    pub width: usize,
    pub height: usize,
    ambient: [f32; 3],
}

impl Forest {
    /// `render_tree::render_forest(&forest, &mut buffer, &ambient)` (src/render_tree.rs:121-127).
    pub fn render(&mut self) -> Result<Image> {
        let mut img = Image::new(self.width, self.height);
        check(unsafe { sys::rt_forest_eval(self.handle, self.ambient.as_ptr(), img.rgb.as_mut_ptr(), ptr::null_mut(), ptr::null_mut()) })?;
        Ok(img)
    }

    /// `render_tree::render_forest_filter` (src/render_tree.rs:129-145): re-shades in place the
    /// pixels whose tree touches one of `mutated`; returns how many trees were evaluated.
    pub fn render_filter(&mut self, mutated: &[i32], img: &mut Image) -> Result<u64> {
        assert!(img.width == self.width && img.height == self.height);
        let mut n = 0u64;
        check(unsafe {
            sys::rt_forest_eval_filtered(self.handle, self.ambient.as_ptr(), mutated.as_ptr(), mutated.len() as i32,
                                         img.rgb.as_mut_ptr(), ptr::null_mut(), &mut n, ptr::null_mut())
        })?;
        Ok(n)
    }

    /// `RayForest::size` (src/render_tree.rs:62-64).
    pub fn size(&self) -> Result<u64> {
        let mut n = 0u64;
        check(unsafe { sys::rt_forest_size(self.handle, &mut n) })?;
        Ok(n)
    }

    /// `RayForest::trees_with` (src/render_tree.rs:68-70).
    pub fn trees_with(&self, shape_id: i32) -> Result<u64> {
        let mut n = 0u64;
        check(unsafe { sys::rt_forest_trees_with(self.handle, shape_id, &mut n) })?;
        Ok(n)
    }

    /// `trees_with` through the entry point that allocates nothing (static device scratch; forests of
    /// at most `RT_STATIC_TREES_MAX` = 4 Mi trees; calls on one GPU must not overlap).  Not in the reference.
    pub fn trees_with_static(&self, shape_id: i32) -> Result<u64> {
        let mut n = 0u64;
        check(unsafe { sys::rt_forest_trees_with_static(self.handle, shape_id, &mut n) })?;
        Ok(n)
    }

    /// `RayForest::stats` (src/render_tree.rs:73-105).
    pub fn stats(&self) -> Result<rt_forest_stats> {
        let mut st = rt_forest_stats::default();
        check(unsafe { sys::rt_forest_get_stats(self.handle, &mut st) })?;
        Ok(st)
    }
}

impl Drop for Forest {
    fn drop(&mut self) {
        unsafe { sys::rt_forest_destroy(self.handle) }
    }
}
