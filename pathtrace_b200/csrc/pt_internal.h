// pt_internal.h — host/device structs shared by the .cu files of libpathtrace_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/pathtrace_b200.h"

constexpr float kFp32DomainMax = 4294967296.0f;   // 2^32, see SceneDev::fp32_domain
constexpr int kBruteParamPairs = 8;   // pair entries of the scan table that can travel as a kernel parameter (pt_wave.cuh: BruteParam)
constexpr int kFastParamPairs = 4;    // pairs of planar entries (parallelograms / triangles) of the secondary-ray scan (pt_wave.cuh: FastParam)
constexpr int kFastParamSpheres = 4;  // spheres that travel with them

// Scene resident in HBM (replicated per GPU).  All arrays are float4-packed for 128-bit loads.
struct SceneDev {
    uint32_t n_objects = 0, n_lights = 0, n_nodes = 0;
    float4* prims = nullptr;      // 4 x float4 per object (64 B): see pt_math.cuh
    float4* mats = nullptr;       // 2 x float4 per object (32 B): see pt_shade.cuh
    uint32_t* light_ids = nullptr;
    float4* nodes = nullptr;      // 4 x float4 per LBVH node (64 B): [Llo,left][Lhi,right][Rlo,0][Rhi,0]
    float4* brute = nullptr;      // brute-force scan table, 5 x float4 per entry (see k_brute_table); scenes <= 1024 objects only
    uint32_t n_brute = 0, n_pairs = 0;   // entries; the first n_pairs are triangle pairs
    // Every shape coordinate (vertices, centres, radii) is finite with magnitude <= kFp32DomainMax: no intermediate of the
    // triangle test can overflow, and every determinant is either below the 1e-8 rejection threshold or a normal number far
    // from 2^126 — the precondition of the unchecked pair test (pt_math.cuh: prcp_rn<false>, hit_triangle2<false, .>).
    bool fp32_domain = false;
    std::vector<pt_shape> shapes_host;   // the caller's shapes for scenes with a scan table (<= 1024 objects): the primary-ray cull projects them per camera
    std::vector<float4> brute_host;      // host copy of the pair entries that fit a kernel parameter (pt_wave.cuh: BruteParam)
    // Secondary-ray scan (pt_wave.cuh: fast_pair): every triangle of the scene as a planar entry with its world -> (u, v, w)
    // affine map; two consecutive triangles that form a parallelogram share ONE entry.  8 float4 per pair of entries; empty
    // when the scene's triangles do not fit kFastParamPairs pairs (the exact scan then serves every ray).
    std::vector<float4> fast_host;
    uint32_t n_fast_pairs = 0;
    float4 fast_spheres[kFastParamSpheres] = {};   // (center, radius) and object ids of the scene's spheres when they all fit
    int fast_sphere_ids[kFastParamSpheres] = {};
    uint32_t n_fast_spheres = 0;
    uint32_t* codes_sorted = nullptr;
    uint32_t* prim_sorted = nullptr;
    float bounds[6] = {0, 0, 0, 0, 0, 0};
};

struct pt_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    int sm_count = 0;
    std::string err;
    // wavefront pool (grown on demand, reused across calls)
    uint32_t pool_cap = 0;
    void* pool_mem = nullptr;
    size_t pool_bytes = 0;
    // frame buffers owned by the context for pt_render (grown on demand)
    float4* d_accum = nullptr; size_t accum_px = 0;
    float* d_radiance = nullptr; uint8_t* d_rgba = nullptr; size_t out_px = 0;
    uint64_t* d_counters = nullptr;   // device counters block
    uint64_t* h_counters = nullptr;   // pinned mirror
    std::vector<cudaEvent_t> prof_events;   // PT_FLAG_PROFILE: start/stop pairs per launch
    // HBM-queue wavefront: the shadow rays of bounce i are traced on a second stream while bounce i + 1 finds its closest hits
    cudaStream_t stream2 = nullptr;
    cudaEvent_t ev_shaded = nullptr, ev_shadowed = nullptr;
};

struct pt_scene {
    pt_context* ctx = nullptr;
    SceneDev dev;
    float build_ms = 0.f;          // scene packing + LBVH build, CUDA events
    uint32_t build_launches = 0;
};

// cudaMalloc'd temporaries that are released on every return path
struct DevTemps {
    std::vector<void*> ptrs;
    ~DevTemps() { for (void* p : ptrs) if (p) cudaFree(p); }
    template <class T> cudaError_t alloc(T** out, size_t bytes) {
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, bytes ? bytes : 1);
        if (e == cudaSuccess) ptrs.push_back(p);
        *out = static_cast<T*>(p);
        return e;
    }
};

// capi.cu: the context-owned frame buffers (d_accum, d_radiance, d_rgba) hold at least npix pixels afterwards
int ensure_frame(pt_context* ctx, size_t npix);
// lbvh.cu
int lbvh_build_device(pt_context* ctx, const pt_shape* d_shapes, uint32_t n, SceneDev& out, uint64_t* launches);
// wavefront.cu
struct RenderLaunch {
    const SceneDev* scene;
    pt_camera cam;
    pt_render_params prm;
    float4* d_accum;
};
int wavefront_render(pt_context* ctx, const RenderLaunch& rl, pt_stats* stats);
// smem_wavefront.cu
bool smem_wavefront_applicable(const pt_context* ctx, const SceneDev& scene, bool brute);
int smem_wavefront_render(pt_context* ctx, const RenderLaunch& rl, bool brute, pt_stats* stats);
// the shared-memory wavefront with its scratch in a file-scope device array (pt_render_device_noalloc)
int smem_wavefront_render_noalloc(pt_context* ctx, const SceneDev& scene, const pt_camera& cam, const pt_render_params& prm,
                                  float* d_radiance, uint8_t* d_rgba);
int wavefront_trace_primary(pt_context* ctx, const SceneDev& scene, const pt_camera& cam, const pt_render_params& prm, int jitter,
                            float ox, float oy, uint32_t sample, int32_t* d_obj, float* d_t, float* d_dir, uint64_t* launches);
// host only: the cull rectangles of make_cull_param in scan order (pt_primary_cull_rects); false when the scene has no constant-bank scan
bool primary_cull_rects_host(const pt_shape* shapes, uint32_t n, const pt_camera& cam, int32_t* rects, int32_t* entry_of_object, uint32_t* n_entries);
int resolve_launch(pt_context* ctx, const float4* d_accum, uint32_t w, uint32_t h, uint32_t spp, float* d_rad, uint8_t* d_rgba);
// total += batch; moment2 += batch^2 (per pixel and channel); batch is zeroed for the next one
int fold_moments_launch(pt_context* ctx, float4* d_batch, float4* d_total, float4* d_moment2, uint32_t npix);
int moments_out_launch(pt_context* ctx, const float4* d_moment2, uint32_t npix, float* d_out_rgb);
int fp32_peak_launch(pt_context* ctx, float* tflops, float* ms);
// capi.cu (host): the planar-entry table of the secondary-ray scan
void build_fast_table(const pt_shape* shapes, uint32_t n, SceneDev& out);
int scene_pack_launch(pt_context* ctx, const pt_shape* d_shapes, const pt_material* d_mats, uint32_t n, SceneDev& out);

#define PT_CUDA_CHECK(ctx, expr)                                                                          \
    do {                                                                                                  \
        cudaError_t _e = (expr);                                                                          \
        if (_e != cudaSuccess) {                                                                          \
            (ctx)->err = std::string(#expr) + ": " + cudaGetErrorString(_e);                              \
            return _e == cudaErrorMemoryAllocation ? PT_ERR_OOM : PT_ERR_CUDA;                            \
        }                                                                                                 \
    } while (0)
