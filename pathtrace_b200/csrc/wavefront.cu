// wavefront.cu — the Monte Carlo sample loop of roxas1533/pathtrace as sm_100a wavefront kernels.
//
// Replaces src/main.rs:43-60 (pixel loop) + src/world.rs:293-333 (spp loop) + src/rendering.rs:34-142
// (MisStrategy::ray_color) / :213-265 (BrdfOnlyStrategy) with the iterative form of SURVEY Appendix A.6:
//
//   K1 k_generate        camera rays for a pool of P path slots            camera.rs:139-147, world.rs:297-299
//   K3 k_trace_closest   closest hit (LBVH or brute force, smem-staged)    world.rs:270-290, shape.rs:53-89/:161-198
//   K4/K6/K7 k_shade     emission / look-ahead MIS resolve, NEE setup, BSDF sample, Russian roulette,
//                        shadow-queue compaction, per-pixel accumulate, in-place path regeneration
//   K5 k_trace_shadow    any-hit over the compacted shadow queue           rendering.rs:62-65
//   K8 k_resolve         mean, sqrt gamma, clamp, RGBA8                    world.rs:315-332
//
// Path state lives in HBM as SoA float4 / uint4 arrays indexed by slot; a slot whose path ends is
// re-seeded in place with the next (pixel, sample) so the pool stays full until the work runs out.
#include <cfloat>
#include <cstdio>
#include <vector>

#include "pt_internal.h"
#include "pt_shade.cuh"

using namespace ptd;

namespace {

constexpr int kBlock = 256;
#ifndef PT_SHADE_BLOCK
#define PT_SHADE_BLOCK 256
#endif
#ifndef PT_SHADE_SLOTS
#define PT_SHADE_SLOTS 1024
#endif
#ifndef PT_SHADE_MIN_BLOCKS
#define PT_SHADE_MIN_BLOCKS 4
#endif
constexpr int kShadeBlock = PT_SHADE_BLOCK, kShadeMinBlocks = PT_SHADE_MIN_BLOCKS;
constexpr int kShadeSlots = PT_SHADE_SLOTS;   // path slots owned by one shade CTA (a multiple of kShadeBlock and of kBlock)
constexpr int kItersPerBatch = 16;
constexpr int kStackDepth = 64;
constexpr uint32_t kMaxSmemPrims = 1024;   // scene staged in shared memory up to this many objects

// device counter block (uint64 words)
// device counter block (uint64 words); every counter sits in its own 256-byte line so their L2 atomics do not serialise
enum { C_STRIDE = 32, C_WORK_NEXT = 0, C_SHADOW_BASE = C_STRIDE, C_ALIVE_BASE = C_SHADOW_BASE + kItersPerBatch * C_STRIDE,
       C_TOTAL = C_ALIVE_BASE + kItersPerBatch * C_STRIDE };

struct Pool {
    uint32_t P;
    float4* ray_o;    // origin.xyz, eta_ratio carried by the ray (camera.rs:6)
    float4* ray_d;    // direction.xyz, bsdf pdf of the scatter that produced the ray (MIS at look-ahead)
    uint2* hit;       // t bits, object id (0xffffffff = miss)
    float4* beta;     // path throughput rgb
    uint2* meta;      // pixel | alive<<31, sample | depth<<24
    float4* sh_o;     // shadow queue: origin.xyz, t_max
    float4* sh_d;     // direction.xyz, slot bits
    float4* sh_c;     // contribution if unoccluded
    uint4* block_stats;   // per-CTA: primary, shadow, scattered, vertices (no atomics)
    uint32_t* seg_count;  // per shade-CTA: shadow rays emitted this iteration (its private segment of the shadow queue)
    unsigned long long* block_next;   // per shade-CTA: next work item of its private, contiguous work range
};

struct Frame {
    CameraDev cam;
    uint32_t npix;
    uint64_t total_work;     // npix * spp of this call
    uint64_t chunk;          // work items owned by each shade-CTA (contiguous range)
    uint32_t sample_offset, min_depth, max_depth, strategy;
    uint32_t k0, k1;         // Philox key
    float t_min;
    float4* accum;           // per pixel: sum r, g, b, (reserved)
};

struct SceneView {
    const float4* prims;
    const float4* mats;
    const uint32_t* light_ids;
    const float4* nodes;
    uint32_t n_objects, n_lights, n_nodes;
};

PT_DEV void make_primary(const Frame& f, uint64_t w, float4& ro, float4& rd, uint2& meta) {
    uint32_t sample = f.sample_offset + (uint32_t)(w / f.npix);
    uint32_t pixel = (uint32_t)(w % f.npix);
    uint32_t x = pixel % f.cam.width, y = pixel / f.cam.width;
    uint4 r = philox4x32_10(pixel, sample, 0xFFFFFFFFu, 0u, f.k0, f.k1);
    float3 d = camera_dir(f.cam, x, f.cam.height - 1 - y, u01(r.x), u01(r.y));   // world.rs:299
    ro = make_float4(f.cam.origin.x, f.cam.origin.y, f.cam.origin.z, 1.0f);      // eta_ratio = 1, camera.rs:14
    rd = make_float4(d.x, d.y, d.z, 0.0f);
    meta = make_uint2(pixel | 0x80000000u, sample & 0x00ffffffu);
}

// K1 — initial fill of the pool.  Work is partitioned statically: shade-CTA b owns the contiguous work range
// [b*chunk, (b+1)*chunk) and hands it out to its own kShadeBlock slots (no global work counter, no global atomics).
__global__ void __launch_bounds__(kBlock) k_generate(Pool pool, const __grid_constant__ Frame f) {
    const uint32_t seg = blockIdx.x / (kShadeSlots / kBlock), l = (blockIdx.x % (kShadeSlots / kBlock)) * kBlock + threadIdx.x;
    const uint32_t i = seg * kShadeSlots + l;
    uint64_t lo = (uint64_t)seg * f.chunk, hi = lo + f.chunk < f.total_work ? lo + f.chunk : f.total_work;
    uint64_t w = lo + l;
    if (l == 0) {
        pool.block_next[seg] = lo + kShadeSlots;
        uint64_t avail = hi > lo ? hi - lo : 0;
        pool.block_stats[seg] = make_uint4((uint32_t)(avail < kShadeSlots ? avail : kShadeSlots), 0u, 0u, 0u);
    }
    if (i >= pool.P) return;
    if (w < hi) {
        float4 ro, rd; uint2 m;
        make_primary(f, w, ro, rd, m);
        pool.ray_o[i] = ro; pool.ray_d[i] = rd; pool.meta[i] = m;
        pool.beta[i] = make_float4(1.f, 1.f, 1.f, 0.f);
    } else {
        pool.meta[i] = make_uint2(0u, 0u);
    }
}

// ------------------------------------------------------------------------------------ traversal
// Scene staged in shared memory: [nodes: 4 float4 each][prims: 3 float4 each (q0,q1,q2)]
template <bool SMEM>
struct Geo {
    const float4* nodes;
    const float4* prims;   // stride 3 (smem) or 4 (global)
    PT_DEV void node(int i, float4& a, float4& b, float4& c, float4& d) const {
        if (SMEM) { a = nodes[4 * i]; b = nodes[4 * i + 1]; c = nodes[4 * i + 2]; d = nodes[4 * i + 3]; }
        else { a = __ldg(nodes + 4 * i); b = __ldg(nodes + 4 * i + 1); c = __ldg(nodes + 4 * i + 2); d = __ldg(nodes + 4 * i + 3); }
    }
    PT_DEV void prim(int i, float4& q0, float4& q1, float4& q2) const {
        if (SMEM) { q0 = prims[3 * i]; q1 = prims[3 * i + 1]; q2 = prims[3 * i + 2]; }
        else { q0 = __ldg(prims + 4 * (size_t)i); q1 = __ldg(prims + 4 * (size_t)i + 1); q2 = __ldg(prims + 4 * (size_t)i + 2); }
    }
};
template <bool SMEM>
PT_DEV Geo<SMEM> stage_scene(const SceneView& s, float4* smem, bool want_nodes) {
    Geo<SMEM> g;
    if (SMEM) {
        uint32_t nn = want_nodes ? s.n_nodes * 4 : 0;
        for (uint32_t i = threadIdx.x; i < nn; i += blockDim.x) smem[i] = __ldg(s.nodes + i);
        for (uint32_t i = threadIdx.x; i < s.n_objects * 3; i += blockDim.x) smem[nn + i] = __ldg(s.prims + 4 * (i / 3) + (i % 3));
        __syncthreads();
        g.nodes = smem; g.prims = smem + nn;
    } else {
        g.nodes = s.nodes; g.prims = s.prims;
    }
    return g;
}

// Slab test against [lo,hi]; conservative: NaN slabs (0 * inf) are ignored, far side widened by 2 ulp.
PT_DEV bool slab(float3 lo, float3 hi, float3 ro, float3 inv, float t_min, float t_max, float& tnear) {
    float tx0 = (lo.x - ro.x) * inv.x, tx1 = (hi.x - ro.x) * inv.x;
    float ty0 = (lo.y - ro.y) * inv.y, ty1 = (hi.y - ro.y) * inv.y;
    float tz0 = (lo.z - ro.z) * inv.z, tz1 = (hi.z - ro.z) * inv.z;
    float tn = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), fmaxf(fminf(tz0, tz1), t_min));
    float tf = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), fminf(fmaxf(tz0, tz1), t_max));
    tnear = tn;
    return tn <= tf * 1.0000004f;
}

// Closest hit.  Winner = smallest t in [t_min, inf); equal t -> higher object index (world.rs:281-287 scans
// in index order with an inclusive t_max, so a later object replaces an earlier one on a tie).
template <bool SMEM, bool BRUTE>
PT_DEV void closest_hit(const Geo<SMEM>& g, uint32_t n_objects, float3 ro, float3 rd, float t_min, float& best_t, int& best_id) {
    best_t = FLT_MAX; best_id = -1;
    float tmax = __int_as_float(0x7f800000);
    auto test = [&](int id) {
        float4 q0, q1, q2;
        g.prim(id, q0, q1, q2);
        float t = hit_prim(q0, q1, q2, ro, rd, t_min, tmax);
        if (t >= 0.0f && (t < tmax || id > best_id)) { tmax = t; best_id = id; }
    };
    if (BRUTE) {
        for (int id = 0; id < (int)n_objects; ++id) test(id);
    } else {
        float3 inv = v3(1.0f / rd.x, 1.0f / rd.y, 1.0f / rd.z);
        int stack[kStackDepth];
        int sp = 0, node = 0;
        while (true) {
            float4 a, b, c, d;
            g.node(node, a, b, c, d);
            int cl = __float_as_int(a.w), cr = __float_as_int(b.w);
            float tl, tr;
            bool hl = slab(v3(a), v3(b), ro, inv, t_min, tmax, tl);
            bool hr = slab(v3(c), v3(d), ro, inv, t_min, tmax, tr);
            if (hl && cl < 0) { test(~cl); hl = false; }
            if (hr && cr < 0) { test(~cr); hr = false; }
            if (hl && hr) {
                if (tr < tl) { int t = cl; cl = cr; cr = t; }
                stack[sp++] = cr; node = cl;
            } else if (hl) node = cl;
            else if (hr) node = cr;
            else { if (sp == 0) break; node = stack[--sp]; }
        }
    }
    if (best_id >= 0) best_t = tmax;
}
// Any hit in [t_min, t_max] (rendering.rs:62-65)
template <bool SMEM, bool BRUTE>
PT_DEV bool any_hit(const Geo<SMEM>& g, uint32_t n_objects, float3 ro, float3 rd, float t_min, float t_max) {
    if (BRUTE) {
        bool occluded = false;
        for (int id = 0; id < (int)n_objects; ++id) {
            float4 q0, q1, q2;
            g.prim(id, q0, q1, q2);
            occluded |= hit_prim(q0, q1, q2, ro, rd, t_min, t_max) >= 0.0f;
        }
        return occluded;
    }
    float3 inv = v3(1.0f / rd.x, 1.0f / rd.y, 1.0f / rd.z);
    int stack[kStackDepth];
    int sp = 0, node = 0;
    while (true) {
        float4 a, b, c, d;
        g.node(node, a, b, c, d);
        int cl = __float_as_int(a.w), cr = __float_as_int(b.w);
        float tl, tr;
        bool hl = slab(v3(a), v3(b), ro, inv, t_min, t_max, tl);
        bool hr = slab(v3(c), v3(d), ro, inv, t_min, t_max, tr);
        if (hl && cl < 0) {
            float4 q0, q1, q2; g.prim(~cl, q0, q1, q2);
            if (hit_prim(q0, q1, q2, ro, rd, t_min, t_max) >= 0.0f) return true;
            hl = false;
        }
        if (hr && cr < 0) {
            float4 q0, q1, q2; g.prim(~cr, q0, q1, q2);
            if (hit_prim(q0, q1, q2, ro, rd, t_min, t_max) >= 0.0f) return true;
            hr = false;
        }
        if (hl && hr) { stack[sp++] = cr; node = cl; }
        else if (hl) node = cl;
        else if (hr) node = cr;
        else { if (sp == 0) return false; node = stack[--sp]; }
    }
}

// K3 — closest hit for every live slot
template <bool SMEM, bool BRUTE>
__global__ void __launch_bounds__(kBlock) k_trace_closest(Pool pool, SceneView s, float t_min) {
    extern __shared__ float4 smem[];
    Geo<SMEM> g = stage_scene<SMEM>(s, smem, !BRUTE);
    uint32_t i = blockIdx.x * kBlock + threadIdx.x;
    if (i >= pool.P) return;
    if ((pool.meta[i].x >> 31) == 0u) return;
    float4 ro = pool.ray_o[i], rd = pool.ray_d[i];
    float t; int id;
    closest_hit<SMEM, BRUTE>(g, s.n_objects, v3(ro), v3(rd), t_min, t, id);
    pool.hit[i] = make_uint2(__float_as_uint(t), (uint32_t)id);
}

// K5 — any-hit over the shadow queue.  The queue is compacted per shade-CTA (ballot/popc + shared-memory atomics):
// segment b holds seg_count[b] dense entries, so every warp here is full except the last one of a segment.
template <bool SMEM, bool BRUTE>
__global__ void __launch_bounds__(kBlock) k_trace_shadow(Pool pool, SceneView s, float t_min, float4* __restrict__ accum) {
    extern __shared__ float4 smem[];
    const uint32_t seg = blockIdx.x / (kShadeSlots / kBlock), l0 = (blockIdx.x % (kShadeSlots / kBlock)) * kBlock;
    const uint32_t n = pool.seg_count[seg];
    if (l0 >= n) return;
    Geo<SMEM> g = stage_scene<SMEM>(s, smem, !BRUTE);
    if (l0 + threadIdx.x >= n) return;
    uint32_t i = seg * kShadeSlots + l0 + threadIdx.x;
    float4 o = pool.sh_o[i], d = pool.sh_d[i];
    if (!any_hit<SMEM, BRUTE>(g, s.n_objects, v3(o), v3(d), t_min, o.w)) {
        float4 c = pool.sh_c[i];                       // K8: color_temp += ..., world.rs:296-313
        atomicAdd(accum + __float_as_uint(d.w), make_float4(c.x, c.y, c.z, 0.f));
    }
}

// ------------------------------------------------------------------------------------ shade
// CTA-wide stream compaction of a 0/1 flag: warp ballot + popc, one shared counter per warp, ONE global
// atomic per CTA.  Returns base (value of the global counter before this CTA's add), this thread's rank
// among the CTA's flagged threads, and the CTA total.
struct Compact { unsigned long long base; uint32_t rank, total; };
PT_DEV Compact block_compact(bool flag, uint64_t* counter, uint32_t* s_warp, unsigned long long* s_base, uint32_t* s_total) {
    uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t ballot = __ballot_sync(0xffffffffu, flag);
    if (lane == 0) s_warp[warp] = __popc(ballot);
    __syncthreads();
    if (warp == 0) {
        uint32_t nw = blockDim.x >> 5;
        uint32_t c = lane < nw ? s_warp[lane] : 0, x = c;
        for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane < nw) s_warp[lane] = x - c;
        if (lane == 31) {
            *s_total = x;
            *s_base = x ? atomicAdd((unsigned long long*)counter, (unsigned long long)x) : 0ull;
        }
    }
    __syncthreads();
    Compact r;
    r.base = *s_base; r.total = *s_total;
    r.rank = s_warp[warp] + __popc(ballot & ((1u << lane) - 1));
    __syncthreads();
    return r;
}

// rendering.rs:91-98
PT_DEV float rr_prob(const Frame& f, float3 tn, uint32_t depth) {
    if (depth < f.min_depth) return 1.0f;
    float l = fminf(luminance(tn), 1.0f);
    if (depth >= f.max_depth) l *= exp2f(-(float)(depth - f.min_depth));
    return l;
}

// meta packing: x = pixel | alive << 31, y = sample | depth << 24
PT_DEV uint2 pack_meta(uint32_t pixel, uint32_t sample, uint32_t depth, bool alive) {
    return make_uint2(pixel | (alive ? 0x80000000u : 0u), (sample & 0x00ffffffu) | (depth << 24));
}
#define META_ALIVE(m) (((m).x >> 31) != 0u)
#define META_PIXEL(m) ((m).x & 0x7fffffffu)
#define META_SAMPLE(m) ((m).y & 0x00ffffffu)
#define META_DEPTH(m) ((m).y >> 24)

enum { CLS_MIRROR = 0, CLS_DIFFUSE = 1, CLS_EMISSIVE = 2, CLS_MISS = 3, CLS_DEAD = 4, CLS_COUNT = 5 };

// K4 + K6 + K7 (+ K8 accumulate, + K1 regeneration).
// Work per path varies by an order of magnitude (miss ~ nothing, diffuse vertex ~ 500 instructions, GGX vertex ~ 1400), so
// a CTA owns kShadeSlots (> its thread count) path slots and schedules them dynamically:
// Phase 0: coalesced load of (meta, hit) of all its slots into shared memory; every slot is classified
//          (mirror / diffuse / emissive / miss / dead) and the CTA counting-sorts the slot list by class, heaviest
//          class first, with warp ballots + popc.
// Phase A: warps pull TILES of 32 consecutive entries of the sorted list from a shared-memory counter until none is left
//          (longest-processing-time-first list scheduling): a tile is (almost always) one class, so warps run coherent
//          code, and no warp idles behind a GGX warp.  Per tile: emission / look-ahead MIS resolve, NEE setup, BSDF sample,
//          Russian roulette; shadow rays are compacted (ballot/popc + one shared atomic per warp) into the CTA's private
//          segment of the shadow queue; finished paths are accumulated into the frame and their slots re-seeded from the
//          CTA's private contiguous work range.  No global atomics except the per-pixel accumulate.
__global__ void __launch_bounds__(kShadeBlock, kShadeMinBlocks) k_shade(Pool pool, SceneView s, const __grid_constant__ Frame f,
                                                                         uint64_t* __restrict__ counters, int iter_in_batch) {
    constexpr int NW = kShadeBlock / 32;
    constexpr int SPT = kShadeSlots / kShadeBlock;      // slots classified per thread
    constexpr int NTILES = kShadeSlots / 32;
    __shared__ uint2 s_hit[kShadeSlots], s_meta[kShadeSlots];
    __shared__ uint16_t s_cls_cnt[CLS_COUNT][SPT * NW];
    __shared__ uint16_t s_order[kShadeSlots];
    __shared__ uint32_t s_stats[5];            // primary, shadow(unused), scattered, vertices, alive
    __shared__ unsigned long long s_next;      // this CTA's next work item
    __shared__ uint32_t s_shadow;              // entries in this CTA's shadow-queue segment
    __shared__ uint32_t s_tile;                // next tile of the sorted list
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t base = blockIdx.x * kShadeSlots;
    const bool mis = f.strategy == PT_STRATEGY_MIS;
    const uint64_t work_lo = (uint64_t)blockIdx.x * f.chunk, work_hi = work_lo + f.chunk < f.total_work ? work_lo + f.chunk : f.total_work;

    // ---- phase 0
    {
        uint32_t cls[SPT], rank[SPT];
#pragma unroll
        for (int k = 0; k < SPT; ++k) {
            uint32_t l = k * kShadeBlock + tid, i = base + l;
            uint32_t c = CLS_DEAD;
            uint2 m = make_uint2(0u, 0u), h = make_uint2(0u, 0xffffffffu);
            if (i < pool.P) {
                m = pool.meta[i];
                h = pool.hit[i];
                if (META_ALIVE(m)) {
                    if (h.y == 0xffffffffu) c = CLS_MISS;
                    else {
                        uint32_t kind = __float_as_uint(__ldg(s.mats + 2 * (size_t)h.y).w);
                        c = kind == PT_MATK_MIRROR ? CLS_MIRROR : kind == PT_MATK_EMISSIVE ? CLS_EMISSIVE : CLS_DIFFUSE;
                    }
                }
            }
            s_meta[l] = m; s_hit[l] = h;
            cls[k] = c; rank[k] = 0;
#pragma unroll
            for (uint32_t cc = 0; cc < CLS_COUNT; ++cc) {
                uint32_t b = __ballot_sync(0xffffffffu, c == cc);
                if (lane == 0) s_cls_cnt[cc][k * NW + warp] = (uint16_t)__popc(b);
                if (c == cc) rank[k] = __popc(b & ((1u << lane) - 1));
            }
        }
        if (tid < 5) s_stats[tid] = 0;
        if (tid == 0) { s_next = pool.block_next[blockIdx.x]; s_shadow = 0; s_tile = 0; }
        __syncthreads();
        if (tid < 32) {   // exclusive prefix over the CLS_COUNT x (SPT*NW) counts in (class, group) order, one warp
            constexpr uint32_t N = CLS_COUNT * SPT * NW;
            uint16_t* flat = &s_cls_cnt[0][0];
            uint32_t run = 0;
            for (uint32_t k0 = 0; k0 < N; k0 += 32) {
                uint32_t k = k0 + tid;
                uint32_t v = k < N ? flat[k] : 0, x = v;
                for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
                if (k < N) flat[k] = (uint16_t)(run + x - v);
                run += __shfl_sync(0xffffffffu, x, 31);
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < SPT; ++k) s_order[s_cls_cnt[cls[k]][k * NW + warp] + rank[k]] = (uint16_t)(k * kShadeBlock + tid);
        __syncthreads();
    }

    // ---- phase A: dynamic tiles
    uint32_t c_primary = 0, c_scatter = 0, c_vertex = 0, c_alive = 0;
    while (true) {
        uint32_t tile = 0;
        if (lane == 0) tile = atomicAdd(&s_tile, 1u);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= NTILES) break;
        const uint32_t j = s_order[tile * 32 + lane];
        const uint32_t i = base + j;
        const uint2 meta = s_meta[j];
        bool alive = META_ALIVE(meta);
        if (!__any_sync(0xffffffffu, alive)) break;          // sorted: dead slots are last
        bool finished = false, emit_shadow = false;
        bool add_L = false;
        float3 Ladd = v3(0.f, 0.f, 0.f);
        float3 sh_o = v3(0.f, 0.f, 0.f), sh_d = v3(0.f, 0.f, 0.f), sh_c = v3(0.f, 0.f, 0.f);
        float sh_tmax = 0.f;
        if (alive) {
            const uint2 h = s_hit[j];
            const uint32_t depth = META_DEPTH(meta);
            if (h.y == 0xffffffffu) {
                finished = true;                                  // background is black, rendering.rs:140-141
            } else {
                const float4 ro4 = pool.ray_o[i], rd4 = pool.ray_d[i];
                const float3 ro = v3(ro4), rd = v3(rd4);
                const float3 beta = v3(pool.beta[i]);
                const int obj = (int)h.y;
                const float t = __uint_as_float(h.x);
                float4 q0 = __ldg(s.prims + 4 * (size_t)obj), q1 = __ldg(s.prims + 4 * (size_t)obj + 1);
                float4 q2 = __ldg(s.prims + 4 * (size_t)obj + 2), q3 = __ldg(s.prims + 4 * (size_t)obj + 3);
                Mat m = load_mat(s.mats, obj);
                float3 p = ro + rd * t;
                const bool is_tri = __float_as_uint(q0.w) == PT_KIND_TRIANGLE;
                float3 n_out = is_tri ? v3(q3) : (p - v3(q0)) / q1.x;     // shape.rs:195 / :86
                const bool front = dot(rd, n_out) < 0.0f;                 // base.rs:20
                float3 n = front ? n_out : -n_out;
                if (m.kind == PT_MATK_EMISSIVE && dot(m.color, m.color) > 0.0f) {
                    // rendering.rs:42-49 for a primary ray; :106-121 (look-ahead MIS) for a scattered ray
                    float3 L;
                    if (depth == 0) L = m.color;
                    else if (!mis) L = beta * m.color;                    // rendering.rs:226-228
                    else {
                        float ps;                                         // shape pdf from the previous vertex, NOT / n_lights (Q2)
                        if (is_tri) { float3 ld; float dd; ps = triangle_pdf_omega(v3(q1), v3(q2), v3(q3), ro, p, ld, dd); }
                        else ps = sphere_pdf_omega(v3(q0), q1.x, ro);
                        float pb = rd4.w;
                        L = beta * (m.color * (pb / (pb + ps)));
                    }
                    finished = true;
                    add_L = true; Ladd = L;
                } else {
                    c_vertex++;
                    const uint32_t pixel = META_PIXEL(meta), sample = META_SAMPLE(meta);
                    uint4 b0 = philox4x32_10(pixel, sample, depth, 0u, f.k0, f.k1);
                    uint4 b1 = philox4x32_10(pixel, sample, depth, 1u, f.k0, f.k1);
                    // ---- NEE (rendering.rs:53-81, world.rs:251-267); visibility resolved by K5
                    bool want_shadow = false;
                    if (mis && s.n_lights > 0) {
                        uint32_t li = __ldg(s.light_ids + urange(b0.x, s.n_lights));
                        float4 l0 = __ldg(s.prims + 4 * (size_t)li), l1 = __ldg(s.prims + 4 * (size_t)li + 1);
                        float4 l2 = __ldg(s.prims + 4 * (size_t)li + 2), l3 = __ldg(s.prims + 4 * (size_t)li + 3);
                        float3 le = v3(__ldg(s.mats + 2 * (size_t)li));
                        float3 ldir; float dist, pdf_shape;
                        sample_light_shape(l0, l1, l2, l3, p, u01(b0.y), u01(b0.z), ldir, dist, pdf_shape);
                        float pl = pdf_shape / (float)s.n_lights;         // world.rs:260
                        float cos_theta = fabsf(dot(n, ldir));            // rendering.rs:68
                        float3 fe; float pb;
                        bsdf_eval(m, rd, ro4.w, ldir, n, fe, pb);         // inherited eta_ratio (Q3)
                        float w_nee = pl / (pl + pb);
                        float3 c = beta * fe * le * (w_nee * cos_theta / pl);
                        if (finite3(c) && (c.x != 0.f || c.y != 0.f || c.z != 0.f)) {
                            want_shadow = true;
                            sh_o = p; sh_d = ldir; sh_c = c; sh_tmax = dist - 0.001f;   // rendering.rs:64
                        }
                    }
                    // ---- BSDF sample + Russian roulette (rendering.rs:83-102)
                    float eta_new = front ? 1.0f / mat_eta(m) : mat_eta(m);   // rendering.rs:20-25
                    float3 wo, fs; float pdf, cos_s;
                    bool ok = bsdf_sample(m, rd, eta_new, n, u01(b0.w), u01(b1.x), u01(b1.y), wo, fs, pdf, cos_s);
                    float3 tn = beta * fs * (cos_s / pdf);
                    ok = ok && finite3(tn);
                    float q = ok ? rr_prob(f, tn, depth) : 0.0f;
                    if (!ok || u01(b1.z) > q) {
                        finished = true;                                  // D of this vertex is dropped too (Q1)
                    } else {
                        emit_shadow = want_shadow;
                        float3 nb = tn / q;
                        pool.beta[i] = make_float4(nb.x, nb.y, nb.z, 0.f);
                        pool.ray_o[i] = make_float4(p.x, p.y, p.z, eta_new);      // rendering.rs:86-87
                        pool.ray_d[i] = make_float4(wo.x, wo.y, wo.z, pdf);
                        pool.meta[i] = pack_meta(pixel, sample, depth + 1, true);
                        c_scatter++;
                    }
                }
            }
        }
        // ---- shadow queue compaction into this CTA's segment: warp ballot/popc + one shared-memory atomic per warp
        {
            uint32_t bal = __ballot_sync(0xffffffffu, emit_shadow);
            uint32_t sbase = 0;
            if (lane == 0 && bal) sbase = atomicAdd(&s_shadow, (uint32_t)__popc(bal));
            sbase = __shfl_sync(0xffffffffu, sbase, 0);
            if (emit_shadow) {
                uint32_t sidx = base + sbase + __popc(bal & ((1u << lane) - 1));
                pool.sh_o[sidx] = make_float4(sh_o.x, sh_o.y, sh_o.z, sh_tmax);
                pool.sh_d[sidx] = make_float4(sh_d.x, sh_d.y, sh_d.z, __uint_as_float(META_PIXEL(meta)));
                pool.sh_c[sidx] = make_float4(sh_c.x, sh_c.y, sh_c.z, 0.f);
            }
        }
        // ---- K8 accumulate (world.rs:296-313) + regeneration from this CTA's private work range
        {
            uint32_t bal = __ballot_sync(0xffffffffu, finished);
            unsigned long long wbase = 0;
            if (lane == 0 && bal) wbase = atomicAdd(&s_next, (unsigned long long)__popc(bal));
            wbase = __shfl_sync(0xffffffffu, wbase, 0);
            if (finished) {
                if (add_L) atomicAdd(f.accum + META_PIXEL(meta), make_float4(Ladd.x, Ladd.y, Ladd.z, 0.f));
                uint64_t w = wbase + __popc(bal & ((1u << lane) - 1));
                if (w < work_hi) {
                    float4 ro, rd; uint2 m;
                    make_primary(f, w, ro, rd, m);
                    pool.ray_o[i] = ro; pool.ray_d[i] = rd; pool.meta[i] = m;
                    pool.beta[i] = make_float4(1.f, 1.f, 1.f, 0.f);
                    c_primary++;
                } else {
                    pool.meta[i] = make_uint2(0u, 0u);
                    alive = false;
                }
            }
        }
        c_alive += alive ? 1u : 0u;
    }
    // ---- bookkeeping: warp reductions -> shared atomics -> one global update per CTA
    {
        uint32_t a = __reduce_add_sync(0xffffffffu, c_alive), v = __reduce_add_sync(0xffffffffu, c_vertex);
        uint32_t c = __reduce_add_sync(0xffffffffu, c_scatter), pr = __reduce_add_sync(0xffffffffu, c_primary);
        if (lane == 0) { atomicAdd(&s_stats[0], pr); atomicAdd(&s_stats[2], c); atomicAdd(&s_stats[3], v); atomicAdd(&s_stats[4], a); }
    }
    __syncthreads();
    if (tid == 0) {
        if (s_stats[4]) atomicAdd((unsigned long long*)(counters + C_ALIVE_BASE + iter_in_batch * C_STRIDE), (unsigned long long)s_stats[4]);
        pool.block_next[blockIdx.x] = s_next;
        pool.seg_count[blockIdx.x] = s_shadow;
        uint4 st = pool.block_stats[blockIdx.x];
        st.x += s_stats[0]; st.y += s_shadow; st.z += s_stats[2]; st.w += s_stats[3];
        pool.block_stats[blockIdx.x] = st;
    }
}

// K8 — world.rs:315-332: mean, store linear radiance, sqrt gamma, clamp [0,1], *255 truncated, alpha 255
__global__ void k_resolve(const float4* __restrict__ accum, uint32_t npix, float inv_spp, float* __restrict__ radiance,
                          uchar4* __restrict__ rgba) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    float4 a = accum[i];
    float r = a.x * inv_spp, g = a.y * inv_spp, b = a.z * inv_spp;
    if (radiance) { radiance[3 * (size_t)i] = r; radiance[3 * (size_t)i + 1] = g; radiance[3 * (size_t)i + 2] = b; }
    if (rgba) {
        auto to8 = [](float v) -> unsigned char {
            float s = sqrtf(v);
            if (!(s == s)) return 0;   // Rust: NaN as u8 == 0
            s = fminf(fmaxf(s, 0.0f), 1.0f);
            return (unsigned char)(s * 255.0f);
        };
        rgba[i] = make_uchar4(to8(r), to8(g), to8(b), 255);
    }
}

// Parity hook: one primary ray per pixel, closest hit only (K1 + K3 without the pool)
template <bool SMEM, bool BRUTE>
__global__ void __launch_bounds__(kBlock) k_trace_primary(SceneView s, const __grid_constant__ Frame f, int jitter, float ox, float oy,
                                                          uint32_t sample, int32_t* __restrict__ obj, float* __restrict__ tt,
                                                          float* __restrict__ dirs) {
    extern __shared__ float4 smem[];
    Geo<SMEM> g = stage_scene<SMEM>(s, smem, !BRUTE);
    uint32_t i = blockIdx.x * kBlock + threadIdx.x;
    if (i >= f.npix) return;
    uint32_t x = i % f.cam.width, y = i / f.cam.width;
    if (jitter) {
        uint4 r = philox4x32_10(i, sample, 0xFFFFFFFFu, 0u, f.k0, f.k1);
        ox = u01(r.x); oy = u01(r.y);
    }
    float3 d = camera_dir(f.cam, x, f.cam.height - 1 - y, ox, oy);
    float t; int id;
    closest_hit<SMEM, BRUTE>(g, s.n_objects, f.cam.origin, d, f.t_min, t, id);
    obj[i] = id;
    tt[i] = id >= 0 ? t : __int_as_float(0x7f800000);
    if (dirs) { dirs[3 * (size_t)i] = d.x; dirs[3 * (size_t)i + 1] = d.y; dirs[3 * (size_t)i + 2] = d.z; }
}

SceneView make_view(const SceneDev& d) {
    SceneView v;
    v.prims = d.prims; v.mats = d.mats; v.light_ids = d.light_ids; v.nodes = d.nodes;
    v.n_objects = d.n_objects; v.n_lights = d.n_lights; v.n_nodes = d.n_nodes;
    return v;
}
Frame make_frame(const pt_camera& c, const pt_render_params& p, float4* accum) {
    Frame f;
    f.cam.origin = make_float3(c.origin[0], c.origin[1], c.origin[2]);
    f.cam.llc = make_float3(c.lower_left_corner[0], c.lower_left_corner[1], c.lower_left_corner[2]);
    f.cam.horizontal = make_float3(c.horizontal[0], c.horizontal[1], c.horizontal[2]);
    f.cam.vertical = make_float3(c.vertical[0], c.vertical[1], c.vertical[2]);
    f.cam.width = c.width; f.cam.height = c.height;
    f.npix = c.width * c.height;
    f.total_work = (uint64_t)f.npix * p.spp;
    f.sample_offset = p.sample_offset; f.min_depth = p.min_depth; f.max_depth = p.max_depth; f.strategy = p.strategy;
    f.k0 = (uint32_t)p.seed; f.k1 = (uint32_t)(p.seed >> 32);
    f.t_min = p.t_min;
    f.accum = accum;
    return f;
}
struct Mode { bool smem, brute; size_t smem_bytes; };
Mode pick_mode(const SceneDev& s, uint32_t traversal) {
    Mode m;
    m.brute = traversal == PT_TRAVERSAL_BRUTE || (traversal == PT_TRAVERSAL_AUTO && s.n_objects <= 16);
    m.smem = s.n_objects <= kMaxSmemPrims;
    m.smem_bytes = m.smem ? sizeof(float4) * ((m.brute ? 0 : (size_t)s.n_nodes * 4) + (size_t)s.n_objects * 3) : 0;
    return m;
}
template <class K> int set_smem(pt_context* ctx, K kernel, size_t bytes) {
    if (bytes > 48 * 1024) PT_CUDA_CHECK(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return PT_OK;
}

}  // namespace

#define PT_DISPATCH(mode, CALL)                                       \
    do {                                                              \
        if (mode.smem && mode.brute) { CALL(true, true); }            \
        else if (mode.smem) { CALL(true, false); }                    \
        else if (mode.brute) { CALL(false, true); }                   \
        else { CALL(false, false); }                                  \
    } while (0)

// FP32 roofline denominator: 8 independent FFMA chains per thread, 4096 FMAs each.
__global__ void __launch_bounds__(256) k_fp32_peak(float* out, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 1
    for (int i = 0; i < 512; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if (s == 12345.678f) out[0] = s;
}
int fp32_peak_launch(pt_context* ctx, float* tflops, float* ms_out) {
    float* d = nullptr;
    PT_CUDA_CHECK(ctx, cudaMalloc(&d, 4));
    int grid = ctx->sm_count * 8 * 16;
    for (int rep = 0; rep < 2; ++rep) {
        if (rep == 1) PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
        k_fp32_peak<<<grid, 256, 0, ctx->stream>>>(d, 0.999f, 0.001f);
    }
    PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
    PT_CUDA_CHECK(ctx, cudaEventSynchronize(ctx->ev[1]));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
    cudaFree(d);
    double flops = (double)grid * 256 * 512 * 8 * 8 * 2;
    if (tflops) *tflops = (float)(flops / (ms * 1e-3) / 1e12);
    if (ms_out) *ms_out = ms;
    return PT_OK;
}

int resolve_launch(pt_context* ctx, const float4* d_accum, uint32_t w, uint32_t h, uint32_t spp, float* d_rad, uint8_t* d_rgba) {
    uint32_t npix = w * h;
    k_resolve<<<(npix + 255) / 256, 256, 0, ctx->stream>>>(d_accum, npix, 1.0f / (float)spp, d_rad, reinterpret_cast<uchar4*>(d_rgba));
    PT_CUDA_CHECK(ctx, cudaGetLastError());
    return PT_OK;
}

int wavefront_trace_primary(pt_context* ctx, const SceneDev& scene, const pt_camera& cam, const pt_render_params& prm, int jitter,
                            float ox, float oy, uint32_t sample, int32_t* d_obj, float* d_t, float* d_dir, uint64_t* launches) {
    SceneView sv = make_view(scene);
    Frame f = make_frame(cam, prm, nullptr);
    Mode mode = pick_mode(scene, prm.traversal);
    uint32_t grid = (f.npix + kBlock - 1) / kBlock;
#define CALL(S, B)                                                                                                            \
    {                                                                                                                         \
        int rc = set_smem(ctx, k_trace_primary<S, B>, mode.smem_bytes);                                                       \
        if (rc) return rc;                                                                                                    \
        k_trace_primary<S, B><<<grid, kBlock, mode.smem_bytes, ctx->stream>>>(sv, f, jitter, ox, oy, sample, d_obj, d_t, d_dir); \
    }
    PT_DISPATCH(mode, CALL);
#undef CALL
    PT_CUDA_CHECK(ctx, cudaGetLastError());
    if (launches) *launches += 1;
    return PT_OK;
}

static int ensure_pool(pt_context* ctx, uint32_t P, Pool& pool) {
    uint32_t nblocks = (P + kBlock - 1) / kBlock;
    size_t per_slot = 16 * 6 + 8 + 8;   // 6 float4 arrays + hit (uint2) + meta (uint2)
    size_t bytes = per_slot * (size_t)P + 128 * (size_t)nblocks + 256 * 16;
    if (bytes > ctx->pool_bytes) {
        if (ctx->pool_mem) cudaFree(ctx->pool_mem);
        ctx->pool_mem = nullptr; ctx->pool_bytes = 0;
        PT_CUDA_CHECK(ctx, cudaMalloc(&ctx->pool_mem, bytes));
        ctx->pool_bytes = bytes;
    }
    char* p = static_cast<char*>(ctx->pool_mem);
    auto take = [&](size_t n) { char* r = p; p += (n + 255) & ~size_t(255); return r; };
    pool.P = P;
    pool.ray_o = (float4*)take(16 * (size_t)P); pool.ray_d = (float4*)take(16 * (size_t)P);
    pool.beta = (float4*)take(16 * (size_t)P);
    pool.sh_o = (float4*)take(16 * (size_t)P); pool.sh_d = (float4*)take(16 * (size_t)P); pool.sh_c = (float4*)take(16 * (size_t)P);
    pool.meta = (uint2*)take(8 * (size_t)P);
    pool.hit = (uint2*)take(8 * (size_t)P);
    pool.block_stats = (uint4*)take(16 * (size_t)nblocks * 4);
    pool.seg_count = (uint32_t*)take(4 * (size_t)nblocks * 4);
    pool.block_next = (unsigned long long*)take(8 * (size_t)nblocks * 4);
    return PT_OK;
}

int wavefront_render(pt_context* ctx, const RenderLaunch& rl, pt_stats* stats) {
    cudaStream_t st = ctx->stream;
    const SceneDev& scene = *rl.scene;
    SceneView sv = make_view(scene);
    Frame f = make_frame(rl.cam, rl.prm, rl.d_accum);
    Mode mode = pick_mode(scene, rl.prm.traversal);
    uint64_t launches = 0, iters = 0;
    if (f.total_work == 0) { if (stats) { stats->samples = 0; } return PT_OK; }

    // default pool: 4 full waves of shade CTAs (SMs x resident CTAs x slots per CTA), i.e. 2.4 M paths on a 148-SM B200
    uint32_t P = rl.prm.path_pool ? rl.prm.path_pool : (uint32_t)ctx->sm_count * kShadeMinBlocks * kShadeSlots * 4u;
    if ((uint64_t)P > f.total_work) P = (uint32_t)f.total_work;
    P = (P + kShadeSlots - 1) / kShadeSlots * kShadeSlots;
    Pool pool;
    int rc = ensure_pool(ctx, P, pool);
    if (rc) return rc;
    uint32_t grid = P / kBlock;
    if (!ctx->d_counters) {
        PT_CUDA_CHECK(ctx, cudaMalloc(&ctx->d_counters, 8 * C_TOTAL));
        PT_CUDA_CHECK(ctx, cudaMallocHost(&ctx->h_counters, 8 * C_TOTAL));
    }
#define SETS(S, B) { rc = set_smem(ctx, k_trace_closest<S, B>, mode.smem_bytes); if (rc) return rc; rc = set_smem(ctx, k_trace_shadow<S, B>, mode.smem_bytes); if (rc) return rc; }
    PT_DISPATCH(mode, SETS);
#undef SETS
    const uint32_t sgrid = P / kShadeSlots;
    f.chunk = (f.total_work + sgrid - 1) / sgrid;
    const bool profile = (rl.prm.flags & PT_FLAG_PROFILE) != 0;
    float ms_k[PT_K_COUNT] = {0, 0, 0, 0};
    uint32_t n_k[PT_K_COUNT] = {0, 0, 0, 0};
    std::vector<int> ev_class;
    size_t ev_used = 0;
    auto prof_begin = [&](int cls) {
        n_k[cls]++;
        if (!profile) return;
        if (ctx->prof_events.size() < ev_used + 2) { cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b); ctx->prof_events.push_back(a); ctx->prof_events.push_back(b); }
        cudaEventRecord(ctx->prof_events[ev_used], st);
        ev_class.push_back(cls);
    };
    auto prof_end = [&]() { if (profile) { cudaEventRecord(ctx->prof_events[ev_used + 1], st); ev_used += 2; } };
    auto prof_collect = [&]() {   // after a stream synchronize
        for (size_t k = 0; k < ev_class.size(); ++k) { float ms = 0; cudaEventElapsedTime(&ms, ctx->prof_events[2 * k], ctx->prof_events[2 * k + 1]); ms_k[ev_class[k]] += ms; }
        ev_class.clear(); ev_used = 0;
    };
    prof_begin(PT_K_GENERATE);
    k_generate<<<grid, kBlock, 0, st>>>(pool, f); launches++;
    prof_end();
    const bool mis = rl.prm.strategy == PT_STRATEGY_MIS && scene.n_lights > 0;
    bool done = false;
    while (!done) {
        PT_CUDA_CHECK(ctx, cudaMemsetAsync(ctx->d_counters + C_SHADOW_BASE, 0, 8 * (C_TOTAL - C_SHADOW_BASE), st));
        for (int it = 0; it < kItersPerBatch; ++it) {
#define TRACE(S, B) k_trace_closest<S, B><<<grid, kBlock, mode.smem_bytes, st>>>(pool, sv, f.t_min)
            prof_begin(PT_K_TRACE_CLOSEST);
            PT_DISPATCH(mode, TRACE);
            prof_end();
#undef TRACE
            prof_begin(PT_K_SHADE);
            k_shade<<<sgrid, kShadeBlock, 0, st>>>(pool, sv, f, ctx->d_counters, it);
            prof_end();
            launches += 2;
            if (mis) {
                prof_begin(PT_K_TRACE_SHADOW);
#define SHADOW(S, B) k_trace_shadow<S, B><<<grid, kBlock, mode.smem_bytes, st>>>(pool, sv, f.t_min, f.accum)
                PT_DISPATCH(mode, SHADOW);
#undef SHADOW
                prof_end();
                launches++;
            }
        }
        iters += kItersPerBatch;
        PT_CUDA_CHECK(ctx, cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 8 * C_TOTAL, cudaMemcpyDeviceToHost, st));
        PT_CUDA_CHECK(ctx, cudaStreamSynchronize(st));
        prof_collect();
        done = ctx->h_counters[C_ALIVE_BASE + (kItersPerBatch - 1) * C_STRIDE] == 0;
    }
    PT_CUDA_CHECK(ctx, cudaGetLastError());
    if (stats) {
        std::vector<uint4> bs(sgrid);
        PT_CUDA_CHECK(ctx, cudaMemcpy(bs.data(), pool.block_stats, 16 * (size_t)sgrid, cudaMemcpyDeviceToHost));
        uint64_t a = 0, b = 0, c = 0, d = 0;
        for (auto& v : bs) { a += v.x; b += v.y; c += v.z; d += v.w; }
        stats->samples = f.total_work;
        stats->rays_primary = a; stats->rays_shadow = b; stats->rays_scattered = c; stats->path_vertices = d;
        stats->wavefront_iters = iters; stats->kernel_launches = launches; stats->path_pool = P;
        for (int k = 0; k < PT_K_COUNT; ++k) { stats->ms_kernel[k] = ms_k[k]; stats->n_kernel[k] = n_k[k]; }
    }
    return PT_OK;
}
