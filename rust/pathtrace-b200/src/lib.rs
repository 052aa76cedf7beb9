//! Safe wrapper over `pathtrace-b200-sys`.  NOT compiled on the build box of this repository (rust/README.md).
//!
//! The reference's `World` keeps building the scene (`World::new`, src/world.rs:65-241) and owning the two output
//! buffers (`data`, `luminance_data`); this crate replaces only the body of its render thread (src/main.rs:42-67):
//!
//! ```ignore
//! let shapes: Vec<ShapeDesc> = world.objects.iter().map(|o| o.shape.describe()).collect();      // the additive trait method
//! let mats:   Vec<MaterialDesc> = world.objects.iter().map(|o| o.material.describe()).collect(); // (reference-describe.patch)
//! let ctx = Context::new(0)?;                       // no sm_100 device => Err: there is no CPU fallback
//! let scene = ctx.scene(&shapes, &mats)?;           // upload + LBVH build on the device
//! let frame = ctx.render(&scene, &camera_desc(&world.camera), &RenderParams { spp: SAMPLE_NUM, ..Default::default() })?;
//! ```
use pathtrace_b200_sys as sys;
use std::ffi::CStr;
use std::ptr;

pub type ShapeDesc = sys::pt_shape;
pub type MaterialDesc = sys::pt_material;
pub type CameraDesc = sys::pt_camera;
pub type Stats = sys::pt_stats;

/// `SphereShape::describe` (shape.rs:38-43)
pub fn sphere(center: [f64; 3], radius: f64) -> ShapeDesc {
    let mut v = [0f32; 9];
    for k in 0..3 {
        v[k] = center[k] as f32;
    }
    v[3] = radius as f32;
    ShapeDesc { kind: sys::PT_SHAPE_SPHERE, v }
}
/// `TriangleShape::describe` (shape.rs:148-152)
pub fn triangle(v0: [f64; 3], v1: [f64; 3], v2: [f64; 3]) -> ShapeDesc {
    let mut v = [0f32; 9];
    for k in 0..3 {
        v[k] = v0[k] as f32;
        v[3 + k] = v1[k] as f32;
        v[6 + k] = v2[k] as f32;
    }
    ShapeDesc { kind: sys::PT_SHAPE_TRIANGLE, v }
}
fn mat(kind: u32, vals: &[f64]) -> MaterialDesc {
    let mut p = [0f32; 8];
    for (d, s) in p.iter_mut().zip(vals) {
        *d = *s as f32;
    }
    MaterialDesc { kind, p }
}
/// `LambertianCosineWeighted::describe` (material.rs:67-69)
pub fn lambertian(albedo: [f64; 3]) -> MaterialDesc { mat(sys::PT_MAT_LAMBERTIAN, &albedo) }
/// `Emissive::describe` (material.rs:126-129)
pub fn emissive(emission: [f64; 3]) -> MaterialDesc { mat(sys::PT_MAT_EMISSIVE, &emission) }
/// `Mirror::describe` (mirror.rs:5-14)
pub fn mirror(roughness: f64, color: [f64; 3], metallic: f64, ior: f64) -> MaterialDesc {
    mat(sys::PT_MAT_MIRROR, &[roughness, color[0], color[1], color[2], metallic, ior])
}
/// `OrenNayar::describe` (material.rs:166-174; a and b as `OrenNayar::new` computes them, :182-196)
pub fn oren_nayar(albedo: [f64; 3], roughness: f64, a: f64, b: f64) -> MaterialDesc {
    mat(sys::PT_MAT_OREN_NAYAR, &[albedo[0], albedo[1], albedo[2], roughness, a, b])
}

#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum Strategy { Mis = 0, BrdfOnly = 1 }

/// The reference's compile-time constants as run-time fields (world.rs:18, rendering.rs:6-10, Cargo.toml:6-10).
#[derive(Clone, Copy, Debug)]
pub struct RenderParams {
    pub spp: u32,
    pub sample_offset: u32,
    pub min_depth: u32,
    pub max_depth: u32,
    pub strategy: Strategy,
    pub seed: u64,
}
impl Default for RenderParams {
    fn default() -> Self {
        RenderParams { spp: 3000, sample_offset: 0, min_depth: 4, max_depth: 50,
                       strategy: if cfg!(feature = "brdf_only") { Strategy::BrdfOnly } else { Strategy::Mis }, seed: 0x5EED }
    }
}
impl RenderParams {
    fn raw(&self) -> sys::pt_render_params {
        sys::pt_render_params { spp: self.spp, sample_offset: self.sample_offset, min_depth: self.min_depth, max_depth: self.max_depth,
                                n_light_samples: 1, strategy: self.strategy as u32, traversal: sys::PT_TRAVERSAL_AUTO, flags: 0,
                                seed: self.seed, t_min: 0.001, path_pool: 0 }
    }
}

/// What `World.luminance_data` (linear mean radiance, W*H*3) and `World.data` (RGBA8, W*H*4) receive.
pub struct Frame { pub radiance: Vec<f32>, pub rgba8: Vec<u8>, pub stats: Stats }

fn message(p: *const std::os::raw::c_char) -> String {
    if p.is_null() { String::new() } else { unsafe { CStr::from_ptr(p) }.to_string_lossy().into_owned() }
}

pub struct Context { raw: *mut sys::pt_context }
pub struct Scene<'c> { raw: *mut sys::pt_scene, _ctx: &'c Context }

impl Context {
    pub fn new(device: i32) -> Result<Self, String> {
        let mut raw = ptr::null_mut();
        let rc = unsafe { sys::pt_context_create(device, &mut raw) };
        if rc != sys::PT_OK { return Err(format!("pt_context_create ({rc}): {}", message(unsafe { sys::pt_last_error(ptr::null()) }))); }
        Ok(Context { raw })
    }
    fn check(&self, rc: i32, what: &str) -> Result<(), String> {
        if rc == sys::PT_OK { Ok(()) } else { Err(format!("{what} ({rc}): {}", message(unsafe { sys::pt_last_error(self.raw) }))) }
    }
    pub fn scene(&self, shapes: &[ShapeDesc], materials: &[MaterialDesc]) -> Result<Scene<'_>, String> {
        if shapes.len() != materials.len() { return Err("shapes and materials differ in length".into()); }
        let mut raw = ptr::null_mut();
        self.check(unsafe { sys::pt_scene_create(self.raw, shapes.as_ptr(), materials.as_ptr(), shapes.len() as u32, &mut raw) }, "pt_scene_create")?;
        Ok(Scene { raw, _ctx: self })
    }
    /// Replaces src/main.rs:43-60 + src/world.rs:293-333 for the whole frame.
    pub fn render(&self, scene: &Scene<'_>, cam: &CameraDesc, p: &RenderParams) -> Result<Frame, String> {
        let n = cam.width as usize * cam.height as usize;
        let mut f = Frame { radiance: vec![0f32; 3 * n], rgba8: vec![0u8; 4 * n], stats: Stats::default() };
        let raw = p.raw();
        self.check(unsafe { sys::pt_render(self.raw, scene.raw, cam, &raw, f.radiance.as_mut_ptr(), f.rgba8.as_mut_ptr(), &mut f.stats) }, "pt_render")?;
        Ok(f)
    }
}
impl Drop for Context { fn drop(&mut self) { unsafe { sys::pt_context_destroy(self.raw) } } }
impl Scene<'_> { pub fn num_lights(&self) -> u32 { unsafe { sys::pt_scene_num_lights(self.raw) } } }
impl Drop for Scene<'_> { fn drop(&mut self) { unsafe { sys::pt_scene_destroy(self.raw) } } }

/// All GPUs of one box from one process: spp split + one ncclReduce inside the library.
pub struct Group { raw: *mut sys::pt_group }
pub struct GroupScene<'g> { raw: *mut sys::pt_group_scene, _g: &'g Group }
impl Group {
    pub fn new(devices: &[i32]) -> Result<Self, String> {
        let mut raw = ptr::null_mut();
        let rc = unsafe { sys::pt_group_create(devices.as_ptr(), devices.len() as i32, &mut raw) };
        if rc != sys::PT_OK { return Err(format!("pt_group_create ({rc}): {}", message(unsafe { sys::pt_group_last_error(ptr::null()) }))); }
        Ok(Group { raw })
    }
    fn check(&self, rc: i32, what: &str) -> Result<(), String> {
        if rc == sys::PT_OK { Ok(()) } else { Err(format!("{what} ({rc}): {}", message(unsafe { sys::pt_group_last_error(self.raw) }))) }
    }
    pub fn scene(&self, shapes: &[ShapeDesc], materials: &[MaterialDesc]) -> Result<GroupScene<'_>, String> {
        let mut raw = ptr::null_mut();
        self.check(unsafe { sys::pt_group_scene_create(self.raw, shapes.as_ptr(), materials.as_ptr(), shapes.len() as u32, &mut raw) }, "pt_group_scene_create")?;
        Ok(GroupScene { raw, _g: self })
    }
    pub fn render(&self, scene: &GroupScene<'_>, cam: &CameraDesc, p: &RenderParams) -> Result<Frame, String> {
        let n = cam.width as usize * cam.height as usize;
        let mut f = Frame { radiance: vec![0f32; 3 * n], rgba8: vec![0u8; 4 * n], stats: Stats::default() };
        let raw = p.raw();
        self.check(unsafe { sys::pt_group_render(self.raw, scene.raw, cam, &raw, f.radiance.as_mut_ptr(), f.rgba8.as_mut_ptr(), &mut f.stats) }, "pt_group_render")?;
        Ok(f)
    }
}
impl Drop for Group { fn drop(&mut self) { unsafe { sys::pt_group_destroy(self.raw) } } }
impl Drop for GroupScene<'_> { fn drop(&mut self) { unsafe { sys::pt_group_scene_destroy(self.raw) } } }
