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
 * SAMPLE_NUM world.rs:18; MIN_DEPTH/MAX_DEPTH/NUM_LIGHT_SAMPLES rendering.rs:6-10; features mis|brdf_only Cargo.toml:6-10. */
enum { PT_STRATEGY_MIS = 0, PT_STRATEGY_BRDF_ONLY = 1 };
/* How hit_scene (world.rs:270-290) is evaluated.  PRIMARY rays get the same closest hit from every value except BRUTE_FAST (id
 * and t bit-equal to the FP32 restatement of shape.rs, tie rule included).
 *   AUTO        <= 16 objects: BRUTE; more: BVH
 *   BVH         GPU-built LBVH (exact primitive tests for every ray)
 *   BRUTE       linear scan in object order like the reference.  The library picks the fastest form the scene allows: triangle
 *               pairs from the constant bank, unchecked reciprocal inside the FP32 coordinate domain, and — for shadow and
 *               scattered rays, which the parity contract holds to 1e-5 on t rather than to the bit — a precomputed-transform
 *               test per triangle or parallelogram (pt_stats.scan_mode 4)
 *   BRUTE_EXACT as BRUTE with the bit-exact Moller-Trumbore test for every ray (scan_mode 3)
 *   BRUTE_PARAM_CHECKED  as BRUTE_EXACT with every range check (scan_mode 2)
 *   BRUTE_TABLE the scan table from shared memory with every range check (scan_mode 1)
 *   BRUTE_FAST  parity hook: the secondary-ray test applied to every ray, so that pt_trace_primary can put it in front of the
 *               oracle (ids equal except on shared edges, t within 1e-5) */
enum { PT_TRAVERSAL_AUTO = 0, PT_TRAVERSAL_BVH = 1, PT_TRAVERSAL_BRUTE = 2, PT_TRAVERSAL_BRUTE_TABLE = 3, PT_TRAVERSAL_BRUTE_PARAM_CHECKED = 4,
       PT_TRAVERSAL_BRUTE_EXACT = 5, PT_TRAVERSAL_BRUTE_FAST = 6, PT_TRAVERSAL_COUNT = 7 };
typedef struct {
    uint32_t spp;              /* samples per pixel rendered by THIS call */
    uint32_t sample_offset;    /* global index of the first sample (spp split across GPUs / progressive chunks); offset + spp <= 2^24 */
    uint32_t min_depth;        /* 4; <= 254 (the path state keeps the depth in 8 bits; paths are cut at depth 254) */
    uint32_t max_depth;        /* 50 (soft: extra 0.5^(depth-4) roulette factor beyond it, rendering.rs:93-95) */
    uint32_t n_light_samples;  /* NUM_LIGHT_SAMPLES, rendering.rs:10: the reference's constant 1; 0 means 1; anything else is PT_ERR_INVALID */
    uint32_t strategy;         /* PT_STRATEGY_* */
    uint32_t traversal;        /* PT_TRAVERSAL_* */
    uint32_t flags;            /* PT_FLAG_* */
    uint64_t seed;             /* Philox key */
    float t_min;               /* 0.001 (rendering.rs:41,:64,:105); finite, >= 0 */
    uint32_t path_pool;        /* wavefront pool size in paths; 0 = default */
} pt_render_params;
enum {
    PT_FLAG_PROFILE = 1,     /* time every kernel launch with CUDA events into pt_stats.ms_kernel (the kernels themselves are unchanged) */
    PT_FLAG_HBM_QUEUES = 2,  /* force the HBM-queue wavefront (one launch per stage per bounce) even when the scene is small
                                enough for the shared-memory wavefront (one persistent launch per frame) */
    PT_FLAG_PHASE_CLOCKS = 8,/* shared-memory wavefront: run the kernel variant that reads the cycle counter around its phases and fill
                                pt_stats.phase_share (a few percent slower; implies the events of PT_FLAG_PROFILE) */
    PT_FLAG_COUNTERS = 4     /* count LBVH node visits and primitive tests into the stats of the HBM-queue wavefront (slower) */
};
enum { PT_K_GENERATE = 0, PT_K_TRACE_CLOSEST = 1, PT_K_SHADE = 2, PT_K_TRACE_SHADOW = 3, PT_K_SMEM_WAVEFRONT = 4, PT_K_COUNT = 5 };

typedef struct {
    uint64_t samples;           /* camera samples started */
    uint64_t rays_primary, rays_shadow, rays_scattered;   /* unique rays traced */
    uint64_t path_vertices;     /* shade invocations on non-emissive hits */
    uint64_t wavefront_iters;   /* wavefront iterations */
    uint64_t kernel_launches;   /* kernels launched by this call */
    uint64_t nodes_visited;     /* PT_FLAG_COUNTERS: LBVH nodes fetched (64 B each), closest + shadow walks */
    uint64_t prim_tests;        /* PT_FLAG_COUNTERS: primitive records fetched and tested (48 B each) */
    float ms_total;             /* device time of the call, CUDA events */
    float ms_upload, ms_render, ms_download;
    float ms_kernel[5];         /* PT_FLAG_PROFILE only: summed CUDA-event time per kernel class (PT_K_*) */
    uint32_t n_kernel[5];       /* launches per kernel class */
    uint32_t path_pool;         /* pool size actually used */
    float phase_share[3];       /* PT_FLAG_PHASE_CLOCKS, shared-memory wavefront: share of warp cycles in closest-hit items, shading (+ shadow ray) items, waiting */
    uint32_t scan_mode;         /* shared-memory wavefront: 0 LBVH, 1 scan table, 2 constant-bank scan (checked), 3 constant-bank scan (in-domain),
                                   4 = 3 for primary rays + precomputed-transform test for shadow and scattered rays; 0xffffffff otherwise */
} pt_stats;

typedef struct pt_context pt_context;
typedef struct pt_scene pt_scene;

/* Binds one CUDA device.  Multi-GPU: either one context per process and rank with the caller reducing the sum buffers
 * (pt_render_accumulate_device / pt_resolve_device; what bench.py does under torchrun), or pt_group_* below. */
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

/* pt_render plus the second moment the per-pixel z-test needs (sits beside the sum of world.rs:296-313).  The spp
 * samples are rendered as `batches` consecutive batches of spp / batches samples (batches must divide spp); batch b's per-pixel
 * radiance SUM S_b goes into the frame sum and S_b^2 into the second moment:
 *   radiance_rgb = (sum_b S_b) / spp              moment2_rgb = sum_b S_b^2          (both W*H*3 floats)
 * batches == spp gives the literal per-sample sum of squares; with B batches the variance of the mean is estimated as
 * (moment2 / B - (spp * radiance / B)^2) / (B - 1) / (spp / B)^2 with B - 1 degrees of freedom.  Same samples as one
 * pt_render call of spp samples (the RNG is keyed by the global sample index). */
int32_t pt_render_moments(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p, uint32_t batches,
                          float* radiance_rgb, float* moment2_rgb, pt_stats* stats_or_null);

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

/* pt_render for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  The same samples and the same
 * kernels as pt_render, but the scratch memory (the float4 sum buffer of the frame and the kernel's work counter / per-CTA
 * statistics) is a static device array inside the library instead of context-owned cudaMalloc'd buffers, and the results are
 * written to the CALLER's device buffers d_radiance_rgb (W*H*3 floats) and d_rgba8 (W*H*4 bytes); either may be null.  The
 * call makes no cudaMalloc / cudaFree / event creation; it enqueues on pt_context_stream and returns without waiting
 * (pt_context_synchronize, or order later work on that stream).
 * Limits (PT_ERR_INVALID with a message otherwise): width * height <= PT_NOALLOC_MAX_PIXELS; p->flags == 0; a scene the
 * shared-memory wavefront serves (up to 1024 objects staged in shared memory; larger scenes need HBM path queues); at most
 * 1024 CTAs (only an explicit p->path_pool can ask for more).
 * The array (PT_NOALLOC_MAX_PIXELS * 16 B + 48 KiB, about 1.05 MiB) is held by every GPU that loads the library until the
 * process exits; all calls on a device share it, it is cleared at the start of each call, and two calls that overlap on the
 * GPU (two contexts of one device, or one context driven from two threads) race on it. */
#define PT_NOALLOC_MAX_PIXELS 65536u
int32_t pt_render_device_noalloc(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p,
                                 void* d_radiance_rgb, void* d_rgba8);

/* Parity hooks.
 * pt_trace_primary: closest hit of one primary ray per pixel.  jitter==0: fixed offsets (ox, oy) (camera.rs:139-147);
 * jitter!=0: the offsets of sample `sample` from the render RNG.  obj_id: W*H int32 (-1 = miss); t: W*H floats (+inf = miss);
 * dir_or_null: W*H*3 floats, the normalised ray directions. */
int32_t pt_trace_primary(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p, int32_t jitter,
                         float ox, float oy, uint32_t sample, int32_t* obj_id, float* t, float* dir_or_null);
/* LBVH download: codes_sorted[n], prim_sorted[n], nodes[max(n-1,1)*16 words], bounds[6] = scene min.xyz, max.xyz. */
int32_t pt_bvh_download(pt_context* ctx, const pt_scene* scene, uint32_t* codes_sorted, uint32_t* prim_sorted, uint32_t* nodes,
                        float* bounds);
/* Host-only hook (needs no device): the pixel rectangles with which the camera rays' constant-bank scan is culled
 * (csrc/pt_wave.cuh: make_cull_param, closest_hit_param_culled) for a scene of at most 8 triangle pairs.  Entries are in scan
 * order — the pairs of consecutive triangles first, then the unpaired primitives.  rects: 4 int32 per entry = x0, y0, x1, y1
 * (inclusive, y counted from the BOTTOM row like camera.rs:139-147 after the flip of world.rs:299; x0 > x1 = empty; the full
 * screen when the entry cannot be bounded); entry_of_object: n_objects int32.  A camera ray of pixel (x, y) with offsets in
 * [0, 1] can only hit objects whose entry's rectangle contains (x, y) — the property tests/test_host_and_abi.py checks against
 * the oracle.  Returns PT_ERR_INVALID when the scene has no such scan (more than 8 pairs or more than 1024 objects). */
int32_t pt_primary_cull_rects(const pt_shape* shapes, uint32_t n_objects, const pt_camera* cam, int32_t* rects, int32_t* entry_of_object,
                              uint32_t* n_entries);
uint32_t pt_scene_num_objects(const pt_scene* scene);
uint32_t pt_scene_num_lights(const pt_scene* scene);
/* Device time (CUDA events) of the LBVH build inside pt_scene_create — primitive boxes, Morton codes, radix sort, hierarchy,
 * refit; allocation and the host-to-device copy of the scene are outside — and the kernels it launched. */
float pt_scene_build_ms(const pt_scene* scene);
uint32_t pt_scene_build_launches(const pt_scene* scene);

/* Multi-GPU through the C ABI: one host process drives n devices.  The path shards by samples per pixel:
 * device g renders the whole frame for the global sample indices [off_g, off_g + spp_g) (spp_g = spp / n, the first spp % n
 * devices one more), scene and LBVH are replicated, and ONE ncclReduce(sum, float32, root = device_ids[0]) of the W*H float4
 * sum buffer combines them; the root resolves (world.rs:315-332).  NCCL is loaded at run time (libnccl.so.2); a group of one
 * device needs no NCCL.  pt_group_render blocks until the root's results are in the caller's buffers; stats (may be null) are
 * the sums over devices, ms_total the wall time of the slowest one. */
typedef struct pt_group pt_group;
typedef struct pt_group_scene pt_group_scene;
int32_t pt_group_create(const int32_t* device_ids, int32_t n, pt_group** out);
void pt_group_destroy(pt_group* g);
int32_t pt_group_size(const pt_group* g);
pt_context* pt_group_context(pt_group* g, int32_t i);
const char* pt_group_last_error(const pt_group* g_or_null);
int32_t pt_group_scene_create(pt_group* g, const pt_shape* shapes, const pt_material* materials, uint32_t n_objects, pt_group_scene** out);
void pt_group_scene_destroy(pt_group_scene* scene);
int32_t pt_group_render(pt_group* g, const pt_group_scene* scene, const pt_camera* cam, const pt_render_params* p,
                        float* radiance_rgb, uint8_t* rgba8, pt_stats* stats_or_null);

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
