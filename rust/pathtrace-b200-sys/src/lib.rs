//! Raw bindings to `libpathtrace_b200.so` — one declaration per export of `include/pathtrace_b200.h`, `#[repr(C)]` twins
//! of its structs.  NOT compiled on the build box of this repository (no Rust toolchain there); the layouts and the symbol
//! list are checked against the header by `tests/test_rust_layout.py`.
//!
//! What each call replaces in roxas1533/pathtrace is cited in the header: `pt_render` = the render thread's
//! `pixels_coords.par_iter().for_each(render_pixel)` (`src/main.rs:43-60` + `src/world.rs:293-333`).
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_void};

// enum values of the header
pub const PT_OK: i32 = 0;
pub const PT_ERR_INVALID: i32 = 1;
pub const PT_ERR_NO_DEVICE: i32 = 2;
pub const PT_ERR_CUDA: i32 = 3;
pub const PT_ERR_OOM: i32 = 4;
pub const PT_SHAPE_SPHERE: u32 = 0;
pub const PT_SHAPE_TRIANGLE: u32 = 1;
pub const PT_MAT_LAMBERTIAN: u32 = 0;
pub const PT_MAT_EMISSIVE: u32 = 1;
pub const PT_MAT_MIRROR: u32 = 2;
pub const PT_MAT_OREN_NAYAR: u32 = 3;
pub const PT_STRATEGY_MIS: u32 = 0;
pub const PT_STRATEGY_BRDF_ONLY: u32 = 1;
pub const PT_TRAVERSAL_AUTO: u32 = 0;
pub const PT_TRAVERSAL_BVH: u32 = 1;
pub const PT_TRAVERSAL_BRUTE: u32 = 2;
pub const PT_TRAVERSAL_BRUTE_TABLE: u32 = 3;
pub const PT_TRAVERSAL_BRUTE_PARAM_CHECKED: u32 = 4;
pub const PT_TRAVERSAL_BRUTE_EXACT: u32 = 5;
pub const PT_TRAVERSAL_BRUTE_FAST: u32 = 6;
pub const PT_FLAG_PROFILE: u32 = 1;
pub const PT_FLAG_HBM_QUEUES: u32 = 2;
pub const PT_FLAG_COUNTERS: u32 = 4;
pub const PT_FLAG_PHASE_CLOCKS: u32 = 8;
pub const PT_NOALLOC_MAX_PIXELS: u32 = 65536;

/// `SphereShape{center, radius}` (shape.rs:38-43): kind 0, v = [cx, cy, cz, r, ..]; `TriangleShape{v0, v1, v2}`
/// (shape.rs:148-152): kind 1, v = [v0.xyz, v1.xyz, v2.xyz].
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct pt_shape {
    pub kind: u32,
    pub v: [f32; 9],
}

/// `LambertianCosineWeighted` kind 0 p = [albedo.rgb]; `Emissive` kind 1 p = [emission.rgb]; `Mirror` kind 2
/// p = [roughness, color.rgb, metallic, ior]; `OrenNayar` kind 3 p = [albedo.rgb, roughness, a, b].
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct pt_material {
    pub kind: u32,
    pub p: [f32; 8],
}

/// The four cached vectors of `Camera` (camera.rs:27-39) + frame size.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct pt_camera {
    pub origin: [f32; 3],
    pub lower_left_corner: [f32; 3],
    pub horizontal: [f32; 3],
    pub vertical: [f32; 3],
    pub width: u32,
    pub height: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct pt_render_params {
    pub spp: u32,
    pub sample_offset: u32,
    pub min_depth: u32,
    pub max_depth: u32,
    pub n_light_samples: u32,
    pub strategy: u32,
    pub traversal: u32,
    pub flags: u32,
    pub seed: u64,
    pub t_min: f32,
    pub path_pool: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct pt_stats {
    pub samples: u64,
    pub rays_primary: u64,
    pub rays_shadow: u64,
    pub rays_scattered: u64,
    pub path_vertices: u64,
    pub wavefront_iters: u64,
    pub kernel_launches: u64,
    pub nodes_visited: u64,
    pub prim_tests: u64,
    pub ms_total: f32,
    pub ms_upload: f32,
    pub ms_render: f32,
    pub ms_download: f32,
    pub ms_kernel: [f32; 5],
    pub n_kernel: [u32; 5],
    pub path_pool: u32,
    pub phase_share: [f32; 3],
    pub scan_mode: u32,
}

#[repr(C)]
pub struct pt_context {
    _private: [u8; 0],
}
#[repr(C)]
pub struct pt_scene {
    _private: [u8; 0],
}
#[repr(C)]
pub struct pt_group {
    _private: [u8; 0],
}
#[repr(C)]
pub struct pt_group_scene {
    _private: [u8; 0],
}

extern "C" {
    pub fn pt_version() -> *const c_char;
    pub fn pt_context_create(device_id: i32, out: *mut *mut pt_context) -> i32;
    pub fn pt_context_destroy(ctx: *mut pt_context);
    pub fn pt_last_error(ctx_or_null: *const pt_context) -> *const c_char;
    pub fn pt_context_stream(ctx: *mut pt_context) -> *mut c_void;
    pub fn pt_context_synchronize(ctx: *mut pt_context) -> i32;

    pub fn pt_scene_create(ctx: *mut pt_context, shapes: *const pt_shape, materials: *const pt_material, n_objects: u32,
                           out: *mut *mut pt_scene) -> i32;
    pub fn pt_scene_destroy(scene: *mut pt_scene);
    pub fn pt_scene_num_objects(scene: *const pt_scene) -> u32;
    pub fn pt_scene_num_lights(scene: *const pt_scene) -> u32;
    pub fn pt_scene_build_ms(scene: *const pt_scene) -> f32;
    pub fn pt_scene_build_launches(scene: *const pt_scene) -> u32;

    pub fn pt_render(ctx: *mut pt_context, scene: *const pt_scene, cam: *const pt_camera, p: *const pt_render_params,
                     radiance_rgb: *mut f32, rgba8: *mut u8, stats_or_null: *mut pt_stats) -> i32;
    pub fn pt_render_moments(ctx: *mut pt_context, scene: *const pt_scene, cam: *const pt_camera, p: *const pt_render_params,
                             batches: u32, radiance_rgb: *mut f32, moment2_rgb: *mut f32, stats_or_null: *mut pt_stats) -> i32;
    pub fn pt_render_accumulate_device(ctx: *mut pt_context, scene: *const pt_scene, cam: *const pt_camera,
                                       p: *const pt_render_params, d_accum_f4: *mut c_void, stats_or_null: *mut pt_stats) -> i32;
    pub fn pt_resolve_device(ctx: *mut pt_context, d_accum_f4: *const c_void, width: u32, height: u32, total_spp: u32,
                             d_radiance_rgb: *mut c_void, d_rgba8: *mut c_void) -> i32;

    /// No counterpart in the reference: `pt_render` into the caller's device buffers with the library's static scratch array
    /// (no cudaMalloc on the path); `width * height <= PT_NOALLOC_MAX_PIXELS`, `flags == 0`.
    pub fn pt_render_device_noalloc(ctx: *mut pt_context, scene: *const pt_scene, cam: *const pt_camera, p: *const pt_render_params,
                                    d_radiance_rgb: *mut c_void, d_rgba8: *mut c_void) -> i32;

    pub fn pt_trace_primary(ctx: *mut pt_context, scene: *const pt_scene, cam: *const pt_camera, p: *const pt_render_params,
                            jitter: i32, ox: f32, oy: f32, sample: u32, obj_id: *mut i32, t: *mut f32, dir_or_null: *mut f32) -> i32;
    pub fn pt_primary_cull_rects(shapes: *const pt_shape, n_objects: u32, cam: *const pt_camera, rects: *mut i32, entry_of_object: *mut i32,
                                 n_entries: *mut u32) -> i32;
    pub fn pt_bvh_download(ctx: *mut pt_context, scene: *const pt_scene, codes_sorted: *mut u32, prim_sorted: *mut u32,
                           nodes: *mut u32, bounds: *mut f32) -> i32;

    pub fn pt_group_create(device_ids: *const i32, n: i32, out: *mut *mut pt_group) -> i32;
    pub fn pt_group_destroy(g: *mut pt_group);
    pub fn pt_group_size(g: *const pt_group) -> i32;
    pub fn pt_group_context(g: *mut pt_group, i: i32) -> *mut pt_context;
    pub fn pt_group_last_error(g_or_null: *const pt_group) -> *const c_char;
    pub fn pt_group_scene_create(g: *mut pt_group, shapes: *const pt_shape, materials: *const pt_material, n_objects: u32,
                                 out: *mut *mut pt_group_scene) -> i32;
    pub fn pt_group_scene_destroy(scene: *mut pt_group_scene);
    pub fn pt_group_render(g: *mut pt_group, scene: *const pt_group_scene, cam: *const pt_camera, p: *const pt_render_params,
                           radiance_rgb: *mut f32, rgba8: *mut u8, stats_or_null: *mut pt_stats) -> i32;

    pub fn pt_export_luminance(path: *const c_char, radiance_rgb: *const f32, width: u32, height: u32) -> i32;
    pub fn pt_write_ppm(path: *const c_char, rgba8: *const u8, width: u32, height: u32) -> i32;
    pub fn pt_measure_fp32_peak(ctx: *mut pt_context, tflops_out: *mut f32, ms_out: *mut f32) -> i32;
}
