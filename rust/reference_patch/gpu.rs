//! src/gpu.rs -- the module a maintainer adds to rust_tracer to put the B200 library under its
//! render path.  SOURCE ONLY: never compiled in this repository (no Rust toolchain here).
//!
//! Cargo.toml:   rt_b200 = { path = "../rust_tracer_b200/rust/rt_b200" }
//! src/main.rs:  mod gpu;
//!
//! Call sites that change:
//!   src/main.rs:251   render::render(&camera, &scene, &mut buffer, depth)         -> gpu::render(..)
//!   src/main.rs:269   render_tree::generate_ray_forest(&camera, scene, w, h, d)   -> gpu::GpuForest::generate(..)
//!   src/main.rs:286   render_tree::render_forest(&forest, &mut buffer, &ambient)  -> forest.render(&mut buffer)
//!   src/main.rs:187, src/gui.rs:168   render_tree::render_forest_filter(..)       -> forest.render_filter(&ids, &mut buffer)
//!   src/gui.rs:221-235   material.set_diffuse(c)   -> also forest.scene_mut().update_materials(&scene.flatten().materials)
//!
//! The one thing the reference's types lack is a way to read geometry back out of a
//! `Box<dyn Renderable>` (src/scene/mod.rs:147-165 has no accessor), so each shape and material
//! gains one method, `flatten`, written below next to the fields it reads.

use crate::math::Matrix;
use crate::render::{Camera, RenderBuffer};
use crate::scene::{Color, ColorFun, Scene};
use rt_b200::{rt_colorfn, rt_light, rt_material, rt_shape, FlatScene};
use std::collections::HashSet;

// ---- what each type writes ---------------------------------------------------------------------

fn mat16(m: &Matrix) -> [f32; 16] {
    let mut out = [0f32; 16];
    for r in 0..4 {
        for c in 0..4 {
            out[r * 4 + c] = m.get(r, c); // row-major, column-vector convention (src/math/matrix.rs:5-9)
        }
    }
    out
}

fn shape(kind: i32, material: i32, transform: &Matrix, inv_transform: &Matrix, p: [f32; 12]) -> rt_shape {
    rt_shape { kind, material, transform: mat16(transform), inv_transform: mat16(inv_transform), p }
}

/// New trait method of `Renderable` (src/scene/mod.rs:147): append this shape (and its material)
/// to `out`.  Implementations, one per shape, each inside the shape's own module because the
/// fields are private:
///
/// ```ignore
/// // src/scene/sphere.rs      fields: transform, inv_transform, material (:11-17)
/// fn flatten(&self, out: &mut FlatScene) -> Result<(), String> {
///     let m = gpu::push_material(out, &*self.material.borrow())?;
///     out.shapes.push(gpu::shape(0, m, &self.transform, &self.inv_transform, [0.0; 12]));
///     Ok(())
/// }
/// // src/scene/plane.rs       fields: origin, normal, u, v, transform, inv_transform (:8-19)
/// fn flatten(&self, out: &mut FlatScene) -> Result<(), String> {
///     let m = gpu::push_material(out, &*self.material.borrow())?;
///     let p = [self.origin.x(), self.origin.y(), self.origin.z(), self.normal.x(), self.normal.y(), self.normal.z(),
///              self.u.x(), self.u.y(), self.u.z(), self.v.x(), self.v.y(), self.v.z()];
///     out.shapes.push(gpu::shape(1, m, &self.transform, &self.inv_transform, p));
///     Ok(())
/// }
/// // src/scene/triangle.rs    fields: verts[0..3], normal (:7-14); transform is ignored by intersect (:51)
/// fn flatten(&self, out: &mut FlatScene) -> Result<(), String> {
///     let m = gpu::push_material(out, &*self.material.borrow())?;
///     let (a, b, c, n) = (&self.verts[0], &self.verts[1], &self.verts[2], &self.normal);
///     let p = [a.x(), a.y(), a.z(), b.x(), b.y(), b.z(), c.x(), c.y(), c.z(), n.x(), n.y(), n.z()];
///     out.shapes.push(gpu::shape(2, m, &self.transform, &self.inv_transform, p));
///     Ok(())
/// }
/// // src/scene/cube.rs        fields: transform, inv_transform; the 12 unit triangles are constants (:21-77)
/// fn flatten(&self, out: &mut FlatScene) -> Result<(), String> {
///     let m = gpu::push_material(out, &*self.triangles.shapes()[0].get_material().unwrap())?;
///     out.shapes.push(gpu::shape(3, m, &self.transform, &self.inv_transform, [0.0; 12]));
///     Ok(())
/// }
/// ```
pub fn _renderable_flatten_doc() {}

/// `ColorFun = fn((f32, f32)) -> Color` (src/scene/material.rs:3) cannot cross the FFI: the device
/// has a closed set of texture functions.  `checkerboard` (src/my_scene.rs:26-43) is recognised by
/// address; anything else must be constant (the reference's only other one is `dim_white`,
/// src/my_scene.rs:11-13) or flattening fails -- the C side would answer RT_ERR_UNSUPPORTED.
pub fn colorfn(f: ColorFun, checkerboard: ColorFun) -> Result<rt_colorfn, String> {
    if f as usize == checkerboard as usize {
        return Ok(rt_colorfn { kind: 1, rgb: [1.0, 1.0, 1.0] }); // WHITE or 0.5 WHITE by cell parity
    }
    let c = f((0.0, 0.0));
    for tx in &[(0.5f32, 0.5f32), (-1.5, 2.5), (7.25, -3.75), (-0.1, -0.1)] {
        let d = f(*tx);
        if d.r != c.r || d.g != c.g || d.b != c.b {
            return Err("texture function is neither constant nor my_scene::checkerboard".into());
        }
    }
    Ok(constant(&c))
}

pub fn constant(c: &Color) -> rt_colorfn {
    rt_colorfn { kind: 0, rgb: [c.r, c.g, c.b] }
}

/// New trait method of `Material` (src/scene/material.rs:9):
///
/// ```ignore
/// // Phong (:24-101)
/// fn flatten(&self) -> Result<rt_material, String> {
///     Ok(rt_material { kind: 0, ambient: gpu::constant(&self.ambient), diffuse: gpu::constant(&self.diffuse),
///                      specular: gpu::constant(&self.specular), power: self.power,
///                      reflectivity: self.reflectivity, refraction_index: self.refraction_index })
/// }
/// // TexturePhong (:103-191)
/// fn flatten(&self) -> Result<rt_material, String> {
///     let cb = crate::my_scene::checkerboard as ColorFun;
///     Ok(rt_material { kind: 1, ambient: gpu::colorfn(self.ambient, cb)?, diffuse: gpu::colorfn(self.diffuse, cb)?,
///                      specular: gpu::colorfn(self.specular, cb)?, power: self.power,
///                      reflectivity: self.reflectivity, refraction_index: self.refraction_index })
/// }
/// ```
pub fn push_material(out: &mut FlatScene, m: &dyn crate::scene::Material) -> Result<i32, String> {
    out.materials.push(m.flatten()?);
    Ok(out.materials.len() as i32 - 1)
}

/// `impl Scene { pub fn flatten(&self) -> FlatScene }` (src/scene/mod.rs): shapes in add order
/// (= ids, mod.rs:40-44), lights in add order, ambient.  Each light gains
/// `fn flatten(&self) -> rt_light`: PointLight -> kind 0 + pos + color (mod.rs:178-187),
/// AmbientLight -> kind 1 + color (mod.rs:223-233).
pub fn flatten(scene: &Scene) -> FlatScene {
    let mut out = FlatScene::default();
    for s in scene.shapes() {
        s.flatten(&mut out).expect("scene uses something the device library has no counterpart for");
    }
    for l in scene.lights() {
        let l: rt_light = l.flatten();
        out.lights.push(l);
    }
    let a = scene.ambient();
    out.ambient = [a.r, a.g, a.b];
    out
}

// ---- drop-in bodies ----------------------------------------------------------------------------

fn store(img: &rt_b200::Image, buffer: &mut RenderBuffer) {
    // the library's images are row-major; RenderBuffer is column-major buf[u][v] (src/render.rs:5-19)
    for v in 0..buffer.h {
        for u in 0..buffer.w {
            let p = img.pixel(u, v);
            buffer.buf[u][v] = Color::new(p[0], p[1], p[2]);
        }
    }
}

/// The reference's `Camera` (src/render.rs:155-163, every field public) as the ABI's `rt_camera`.
pub fn flatten_camera(camera: &Camera) -> rt_b200::rt_camera {
    rt_b200::rt_camera {
        origin: [camera.origin.x(), camera.origin.y(), camera.origin.z()],
        x_min: camera.x_min,
        x_max: camera.x_max,
        y_min: camera.y_min,
        y_max: camera.y_max,
        x_res: camera.x_res as i32,
        y_res: camera.y_res as i32,
    }
}

/// Replaces the body of `render::render(camera, scene, buffer, depth)` (src/render.rs:31-38) for
/// `--method basic`, same arguments.  Renders on every GPU of the box (`--gpus N` of the patched
/// CLI narrows that; one device = the single-GPU path of the library).
/// Panics on failure like the reference does on its own errors (src/cli.rs:123).
pub fn render(camera: &Camera, scene: &Scene, buffer: &mut RenderBuffer, depth: usize) {
    let n = rt_b200::device_count().unwrap_or_else(|e| panic!("{}", e));
    let devices: Vec<i32> = (0..n as i32).collect();
    render_on(&devices, camera, scene, buffer, depth)
}

/// `render` on the listed CUDA devices: one blocking call, the library drives them all.
pub fn render_on(devices: &[i32], camera: &Camera, scene: &Scene, buffer: &mut RenderBuffer, depth: usize) {
    let flat = flatten(scene);
    let cam = flatten_camera(camera);
    let (img, _stats) = rt_b200::render_multi(&flat, devices, &cam, depth).unwrap_or_else(|e| panic!("{}", e));
    store(&img, buffer);
}

/// Stands where `RayForest` stood (src/render_tree.rs:53-119): the trees stay on the device.
pub struct GpuForest {
    forest: Option<rt_b200::Forest>, // dropped before the scene it was built from
    scene: rt_b200::Scene,
}

impl GpuForest {
    /// `render_tree::generate_ray_forest` (src/render_tree.rs:147-164).
    pub fn generate(scene: &Scene, width: usize, height: usize, depth: usize) -> GpuForest {
        let flat = flatten(scene);
        let mut gpu = rt_b200::Scene::new(&flat, 0).unwrap_or_else(|e| panic!("{}", e));
        let forest = gpu.generate_ray_forest(width, height, depth).unwrap_or_else(|e| panic!("{}", e));
        GpuForest { forest: Some(forest), scene: gpu }
    }

    pub fn scene_mut(&mut self) -> &mut rt_b200::Scene {
        &mut self.scene
    }

    /// `render_tree::render_forest` (src/render_tree.rs:121-127).
    pub fn render(&mut self, buffer: &mut RenderBuffer) {
        let img = self.forest.as_mut().unwrap().render().unwrap_or_else(|e| panic!("{}", e));
        store(&img, buffer);
    }

    /// `render_tree::render_forest_filter` (src/render_tree.rs:129-145): only the trees that touch a
    /// mutated shape are shaded again; the other pixels of `buffer` keep their colour.
    pub fn render_filter(&mut self, mutated: &HashSet<i32>, buffer: &mut RenderBuffer) -> u64 {
        let ids: Vec<i32> = mutated.iter().cloned().collect();
        let mut img = rt_b200::Image::new(buffer.w, buffer.h);
        for v in 0..buffer.h {
            for u in 0..buffer.w {
                let c = &buffer.buf[u][v];
                let i = (v * buffer.w + u) * 3;
                img.rgb[i] = c.r;
                img.rgb[i + 1] = c.g;
                img.rgb[i + 2] = c.b;
            }
        }
        let n = self.forest.as_mut().unwrap().render_filter(&ids, &mut img).unwrap_or_else(|e| panic!("{}", e));
        store(&img, buffer);
        n
    }

    pub fn size(&self) -> usize {
        self.forest.as_ref().unwrap().size().unwrap_or_else(|e| panic!("{}", e)) as usize
    }

    pub fn trees_with(&self, shape_id: i32) -> usize {
        self.forest.as_ref().unwrap().trees_with(shape_id).unwrap_or_else(|e| panic!("{}", e)) as usize
    }
}

impl Drop for GpuForest {
    fn drop(&mut self) {
        self.forest.take(); // the forest first, then `scene`
    }
}
