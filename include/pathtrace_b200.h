/* pathtrace_b200.h — C ABI of the B200-native render path (libpathtrace_b200.so).
 *
 * Drop-in boundary for the per-pixel Monte Carlo sample loop of roxas1533/pathtrace.  The reference has
 * no FFI seam (src/main.rs:1 forbids unsafe); the seam is introduced where its render thread runs
 *     pixels_coords.par_iter().for_each(|&(x, y)| { ... world.render_pixel(x, y, &mut rng) ... })
 * (src/main.rs:43-60, src/world.rs:293-333): ONE pt_render call replaces that whole loop.
 * Plain pointers and sizes only; no C++ / torch types.  All functions return 0 on success, a
 * pt_status code otherwise; pt_last_error() gives the message.  There is NO CPU fallback: without a
 * usable sm_100 device pt_context_create fails.
 *
 * Ownership: handles are owned by the library and freed by *_destroy.  Arrays passed in are copied
 * before return; output buffers are caller-allocated; no caller pointer is retained.
 * Threading: a context is driven by one host thread at a time; calls block until results are in the
 * caller's buffers (except the *_device entry points, which are stream-ordered on the context stream).
 */
#ifndef PATHTRACE_B200_H
#define PATHTRACE_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    PT_OK = 0,
    PT_ERR_INVALID = 1,    /* bad argument */
    PT_ERR_NO_DEVICE = 2,  /* no CUDA device / not sm_100 */
    PT_ERR_CUDA = 3,       /* CUDA runtime error (message in pt_last_error) */
    PT_ERR_OOM = 4
} pt_status;

/* src/objects/shape.rs:38-43 (SphereShape{center,radius}) and :148-152 (TriangleShape{v0,v1,v2}).
 * kind 0: v = {cx,cy,cz, radius, -,-,-,-,-};  kind 1: v = {v0.xyz, v1.xyz, v2.xyz}. */
enum { PT_SHAPE_SPHERE = 0, PT_SHAPE_TRIANGLE = 1 };
typedef struct { uint32_t kind; float v[9]; } pt_shape;

/* src/objects/material.rs:67-69 LambertianCosineWeighted{albedo}        p = {albedo.rgb}
 * src/objects/material.rs:126-129 Emissive{emission}                    p = {emission.rgb}
 * src/objects/mirror.rs:5-14 Mirror{roughness,color,metallic,ior}       p = {roughness, color.rgb, metallic, ior}
 * src/objects/material.rs:166-174 OrenNayar{albedo,roughness,a,b}       p = {albedo.rgb, roughness, a, b} */
enum { PT_MAT_LAMBERTIAN = 0, PT_MAT_EMISSIVE = 1, PT_MAT_MIRROR = 2, PT_MAT_OREN_NAYAR = 3 };
typedef struct { uint32_t kind; float p[8]; } pt_material;

/* src/camera.rs:27-39: the four cached vectors + frame size (what get_ray_with_offset reads). */
typedef struct {
    float origin[3], lower_left_corner[3], horizontal[3], vertical[3];
    uint32_t width, height;
} pt_camera;

/* Compile-time constants / cargo features of the reference turned into run-time fields:
 * SAMPLE_NUM world.rs:18; MIN_DEPTH/MAX_DEPTH rendering.rs:6-7; features mis|brdf_only Cargo.toml:6-10. */
enum { PT_STRATEGY_MIS = 0, PT_STRATEGY_BRDF_ONLY = 1 };
enum { PT_TRAVERSAL_AUTO = 0, PT_TRAVERSAL_BVH = 1, PT_TRAVERSAL_BRUTE = 2 };
typedef struct {
    uint32_t spp;            /* samples per pixel rendered by THIS call */
    uint32_t sample_offset;  /* global index of the first sample (spp split across GPUs / progressive chunks) */
    uint32_t min_depth;      /* 4  */
    uint32_t max_depth;      /* 50 (soft: extra 0.5^(depth-4) roulette factor beyond it, rendering.rs:93-95) */
    uint32_t strategy;       /* PT_STRATEGY_* */
    uint32_t traversal;      /* PT_TRAVERSAL_* */
    uint64_t seed;           /* Philox key */
    float t_min;             /* 0.001 (rendering.rs:41,:64,:105) */
    uint32_t path_pool;      /* wavefront pool size in paths; 0 = default */
    uint32_t flags;          /* PT_FLAG_* */
} pt_render_params;
enum {
    PT_FLAG_PROFILE = 1,    /* time every kernel launch with CUDA events into pt_stats.ms_kernel (slower) */
    PT_FLAG_HBM_QUEUES = 2  /* force the HBM-queue wavefront (one launch per stage per bounce) even when the scene is small
                               enough for the shared-memory wavefront (one persistent launch per frame) */
};
enum { PT_K_GENERATE = 0, PT_K_TRACE_CLOSEST = 1, PT_K_SHADE = 2, PT_K_TRACE_SHADOW = 3, PT_K_SMEM_WAVEFRONT = 4, PT_K_COUNT = 5 };

typedef struct {
    uint64_t samples;           /* camera samples started */
    uint64_t rays_primary, rays_shadow, rays_scattered;   /* unique rays traced */
    uint64_t path_vertices;     /* shade invocations on non-emissive hits */
    uint64_t wavefront_iters;   /* wavefront iterations */
    uint64_t kernel_launches;   /* kernels launched by this call */
    float ms_total;             /* device time of the call, CUDA events */
    float ms_upload, ms_render, ms_download;
    float ms_kernel[5];         /* PT_FLAG_PROFILE only: summed CUDA-event time per kernel class (PT_K_*) */
    uint32_t n_kernel[5];       /* launches per kernel class */
    uint32_t path_pool;         /* pool size actually used */
    float phase_share[3];       /* PT_FLAG_PROFILE, shared-memory wavefront: share of warp cycles in closest-hit items, shading (+ shadow ray) items, waiting */
} pt_stats;

typedef struct pt_context pt_context;
typedef struct pt_scene pt_scene;

/* Binds one CUDA device (one process per GPU; multi-GPU = one context per rank + a reduce of the sum buffer). */
int32_t pt_context_create(int32_t device_id, pt_context** out);
void pt_context_destroy(pt_context* ctx);
const char* pt_last_error(const pt_context* ctx_or_null);

/* Replaces World::new's object list (world.rs:80-225): uploads n objects, finds lights (emission length > 0,
 * world.rs:214-225) and builds the LBVH on the device.  Object id == index in the arrays. */
int32_t pt_scene_create(pt_context* ctx, const pt_shape* shapes, const pt_material* materials, uint32_t n_objects, pt_scene** out);
void pt_scene_destroy(pt_scene* scene);

/* Replaces main.rs:43-60 + world.rs:293-333 for the whole frame.  Outputs (either may be null), row-major, y down:
 *   radiance_rgb: W*H*3 floats, linear mean radiance (== World.luminance_data, world.rs:318-319)
 *   rgba8:        W*H*4 bytes, sqrt gamma, clamp, *255 truncated, alpha 255 (== World.data, world.rs:321-332) */
int32_t pt_render(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p,
                  float* radiance_rgb, uint8_t* rgba8, pt_stats* stats_or_null);

/* Device-resident variant for the multi-GPU spp split: ADDS this call's samples into d_accum (device pointer,
 * W*H float4 = {sum r, sum g, sum b, reserved}); caller zeroes it, reduces it across ranks and resolves it. */
int32_t pt_render_accumulate_device(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p,
                                    void* d_accum_f4, pt_stats* stats_or_null);
/* mean = sum / total_spp; writes d_radiance_rgb (W*H*3 floats, may be null) and d_rgba8 (W*H*4 bytes, may be null). */
int32_t pt_resolve_device(pt_context* ctx, const void* d_accum_f4, uint32_t width, uint32_t height, uint32_t total_spp,
                          void* d_radiance_rgb, void* d_rgba8);
/* The CUDA stream (cudaStream_t) the *_device entry points are ordered on. */
void* pt_context_stream(pt_context* ctx);
int32_t pt_context_synchronize(pt_context* ctx);

/* Parity hooks.
 * pt_trace_primary: closest hit of one primary ray per pixel.  jitter==0: fixed offsets (ox, oy) (camera.rs:139-147);
 * jitter!=0: the offsets of sample `sample` from the render RNG.  obj_id: W*H int32 (-1 = miss); t: W*H floats (+inf = miss);
 * dir_or_null: W*H*3 floats, the normalised ray directions. */
int32_t pt_trace_primary(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p, int32_t jitter,
                         float ox, float oy, uint32_t sample, int32_t* obj_id, float* t, float* dir_or_null);
/* LBVH download: codes_sorted[n], prim_sorted[n], nodes[max(n-1,1)*16 words], bounds[6] = scene min.xyz, max.xyz. */
int32_t pt_bvh_download(pt_context* ctx, const pt_scene* scene, uint32_t* codes_sorted, uint32_t* prim_sorted, uint32_t* nodes,
                        float* bounds);
uint32_t pt_scene_num_objects(const pt_scene* scene);
uint32_t pt_scene_num_lights(const pt_scene* scene);

/* world.rs:344-369 — writes luminance.csv byte-compatibly ("x,y,r,g,b,luminance", {:.6}) from a radiance buffer. */
int32_t pt_export_luminance(const char* path, const float* radiance_rgb, uint32_t width, uint32_t height);
/* Headless stand-in for World::draw (world.rs:335-341): binary PPM (P6) of the RGBA8 buffer. */
int32_t pt_write_ppm(const char* path, const uint8_t* rgba8, uint32_t width, uint32_t height);

/* FP32 FMA microbenchmark on this device (roofline denominator measured in the same run): dense independent
 * FFMA chains on every SM; returns TFLOP/s counting an FMA as 2 flops. */
int32_t pt_measure_fp32_peak(pt_context* ctx, float* tflops_out, float* ms_out);

const char* pt_version(void);

#ifdef __cplusplus
}
#endif
#endif
