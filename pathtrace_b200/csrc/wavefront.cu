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

#include "pt_wave.cuh"

using namespace ptw;

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
// queue slots per grab of a stream-traversal warp (divides kShadeSlots).  Measured (C3 shared-memory walk / C4 L2 walk, Msamples/s):
// static ranges 1004 / 398; chunks of 256: 1153 / 357, 128: 1254 / 426, 64: 1278 / 458, 32: 1223 / 462
constexpr uint32_t kStreamChunkSmem = 64, kStreamChunkL2 = 32;
static_assert(PT_SHADE_SLOTS % 64 == 0, "a chunk must lie inside one shadow-queue segment");

// device counter block (uint64 words)
// device counter block (uint64 words); every counter sits in its own 256-byte line so their L2 atomics do not serialise
enum { C_STRIDE = 32, C_WORK_NEXT = 0, C_SHADOW_BASE = C_STRIDE, C_ALIVE_BASE = C_SHADOW_BASE + kItersPerBatch * C_STRIDE,
       C_WALK = C_ALIVE_BASE + kItersPerBatch * C_STRIDE,   // PT_FLAG_COUNTERS: [0] LBVH node fetches, [1] primitive tests
       C_TOTAL = C_WALK + C_STRIDE };

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

// K3 / K5 for scenes that walk the LBVH (pt_wave.cuh: bvh_stream).  A resident grid of warps; each takes chunks of
// kStreamChunk* slots of the queue from a global counter and streams them through its lanes.
// Occupancy of the walk through L1/L2 (C4, bound by memory latency).  With a static range per warp every extra warp helped
// (8 CTAs = 32 registers per thread with ~100 bytes of spills: 290 -> 314 Msamples/s); with the queue handed out in chunks
// the spill traffic in the dependent chain costs more than the warps hide: 8 CTAs 465, 7: 468, 6 (40 registers): 477,
// 5 (48 registers, no spills): 512.  Walking the tree in shared memory (C3) is compute-bound and keeps its registers.
#ifndef PT_STREAM_MINB
#define PT_STREAM_MINB 5   // resident CTAs per SM asked of the L1/L2 walk
#endif
// The walk with the scene in shared memory runs larger CTAs than the rest: more warps share each staged copy of the scene
// (C3: 79 KB, two copies fit), so the warp count is set by the register file (32 warps at 64 registers) and not by the copies.
#ifndef PT_STREAM_BLOCK_SMEM
#define PT_STREAM_BLOCK_SMEM 512    // C3 at 16 spp: 256 threads (2 CTAs per SM) 1395, 512 (2 CTAs) 1443, 768 (1 CTA) 1371, 1024 (1 CTA) 1433 Msamples/s
#endif
constexpr int kStreamBlockSmem = PT_STREAM_BLOCK_SMEM;
template <bool SMEM, bool COUNT>
__global__ void __launch_bounds__(SMEM ? kStreamBlockSmem : kBlock, SMEM ? 1 : PT_STREAM_MINB) k_trace_closest_stream(Pool pool, SceneView s, float t_min, uint32_t chunk, unsigned long long* __restrict__ next_chunk,
                                                                                            unsigned long long* __restrict__ walk_counters) {
    extern __shared__ float4 smem[];
    Geo<SMEM> g = stage_scene<SMEM>(s, smem, true);
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    // The queue is handed out in chunks from one global counter: a warp whose rays end early takes the next chunk, so all
    // resident warps stay busy until the queue is empty (a static range per warp left 30 % of the warp slots idle on C4).
    uint32_t cursor = 0, end = 0;
    auto grab = [&]() {
        unsigned long long c = 0;
        if (lane == 0) c = atomicAdd(next_chunk, (unsigned long long)chunk);
        c = __shfl_sync(0xffffffffu, c, 0);
        cursor = c < pool.P ? (uint32_t)c : pool.P;
        end = cursor + chunk < pool.P ? cursor + chunk : pool.P;
    };
    auto next = [&](bool want, uint32_t idle, float3& o, float3& d, float& tm, uint32_t& slot, bool& exhausted) {
        if (cursor >= end) grab();                                   // warp-uniform
        bool got = false;
        const uint32_t i = cursor + __popc(idle & lt);
        if (want && i < end && (pool.meta[i].x >> 31) != 0u) {
            o = v3(pool.ray_o[i]); d = v3(pool.ray_d[i]); tm = __int_as_float(0x7f800000); slot = i; got = true;
        }
        cursor += __popc(idle);
        exhausted = cursor >= end && end >= pool.P;
        return got;
    };
    auto finish = [&](uint32_t slot, float t, int id) { pool.hit[slot] = make_uint2(__float_as_uint(t), (uint32_t)id); };
    bvh_stream<SMEM, false, COUNT>(g, t_min, next, finish, walk_counters);
}
template <bool SMEM, bool COUNT>
__global__ void __launch_bounds__(SMEM ? kStreamBlockSmem : kBlock, SMEM ? 1 : PT_STREAM_MINB) k_trace_shadow_stream(Pool pool, SceneView s, float t_min, float4* __restrict__ accum, uint32_t chunk,
                                                                                           unsigned long long* __restrict__ next_chunk, unsigned long long* __restrict__ walk_counters) {
    extern __shared__ float4 smem[];
    Geo<SMEM> g = stage_scene<SMEM>(s, smem, true);
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    uint32_t cursor = 0, end = 0;                                    // current chunk (a divisor of kShadeSlots: inside one segment)
    auto grab = [&]() {
        unsigned long long c = 0;
        if (lane == 0) c = atomicAdd(next_chunk, (unsigned long long)chunk);
        c = __shfl_sync(0xffffffffu, c, 0);
        cursor = c < pool.P ? (uint32_t)c : pool.P;
        end = cursor + chunk < pool.P ? cursor + chunk : pool.P;
    };
    // the shadow queue is dense per shade-CTA segment: entries [seg * kShadeSlots, seg * kShadeSlots + seg_count[seg])
    auto next = [&](bool want, uint32_t idle, float3& o, float3& d, float& tm, uint32_t& slot, bool& exhausted) {
        uint32_t take = 0;
        while (true) {                                               // warp-uniform: skip chunks in the empty tail of a segment
            if (cursor >= end) { if (end >= pool.P && end != 0u) break; grab(); if (cursor >= end) break; }
            const uint32_t seg = cursor / kShadeSlots, off = cursor % kShadeSlots, cnt = pool.seg_count[seg];
            if (off >= cnt) { cursor = end; continue; }
            take = min(min((uint32_t)__popc(idle), cnt - off), end - cursor);
            break;
        }
        bool got = false;
        const uint32_t r = __popc(idle & lt);
        if (want && r < take) {
            const uint32_t i = cursor + r;
            const float4 so = pool.sh_o[i], sd = pool.sh_d[i];
            o = v3(so); d = v3(sd); tm = so.w; slot = i; got = true;
        }
        cursor += take;
        exhausted = cursor >= end && end >= pool.P;
        return got;
    };
    auto finish = [&](uint32_t slot, float, int id) {
        if (id < 0) {                                                // unoccluded: K8, color_temp += ..., world.rs:296-313
            const float4 c = pool.sh_c[slot];
            atomicAdd(accum + __float_as_uint(pool.sh_d[slot].w), make_float4(c.x, c.y, c.z, 0.f));
        }
    };
    bvh_stream<SMEM, true, COUNT>(g, t_min, next, finish, walk_counters);
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
    const SceneGlobal sg{s.prims, s.mats, s.light_ids, s.n_lights};
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
        PT_CHECK(j < (uint32_t)kShadeSlots);
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
            if (h.y == 0xffffffffu) {
                finished = true;                                  // background is black, rendering.rs:140-141
            } else {
                const float4 ro4 = pool.ray_o[i], rd4 = pool.ray_d[i];
                const float3 beta = v3(pool.beta[i]);
                VertexOut vo = shade_vertex(sg, f, mis, ro4, rd4, beta, h, meta);
                finished = vo.finished; add_L = vo.add_L; Ladd = vo.Ladd;
                c_vertex += vo.vertex ? 1u : 0u;
                emit_shadow = vo.emit_shadow;                         // a finished path may still owe its last vertex's shadow ray
                sh_o = v3(vo.new_ro); sh_d = vo.sh_d; sh_c = vo.sh_c; sh_tmax = vo.sh_tmax;
                if (!vo.finished) {
                    pool.beta[i] = make_float4(vo.new_beta.x, vo.new_beta.y, vo.new_beta.z, 0.f);
                    pool.ray_o[i] = vo.new_ro;
                    pool.ray_d[i] = vo.new_rd;
                    pool.meta[i] = vo.new_meta;
                    c_scatter++;
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
                PT_CHECK(sidx < pool.P && sbase + __popc(bal) <= (uint32_t)kShadeSlots);
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

// K8 — world.rs:315-332: mean, store linear radiance, sqrt gamma, clamp [0,1], *255 truncated, alpha 255.
// The mean is the correctly rounded FP32 quotient; gamma runs in f64 on that value exactly as world.rs:321-332 does on its f64
// mean, so the bytes equal the reference's tone map of the same radiance (tests/test_gpu_outputs.py) whatever -prec-sqrt says.
__global__ void k_resolve(const float4* __restrict__ accum, uint32_t npix, float spp, float* __restrict__ radiance,
                          uchar4* __restrict__ rgba) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    float4 a = accum[i];
    float r = __fdiv_rn(a.x, spp), g = __fdiv_rn(a.y, spp), b = __fdiv_rn(a.z, spp);
    if (radiance) { radiance[3 * (size_t)i] = r; radiance[3 * (size_t)i + 1] = g; radiance[3 * (size_t)i + 2] = b; }
    if (rgba) {
        auto to8 = [](float v) -> unsigned char {
            double s = __dsqrt_rn((double)v);
            if (!(s == s)) return 0;   // Rust: NaN as u8 == 0
            s = s < 0.0 ? 0.0 : (s > 1.0 ? 1.0 : s);
            return (unsigned char)(s * 255.0);
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
        uint4 r = rng_block(i, sample, 0xFFFFFFFFu, 0u, f.k0, f.k1);
        ox = u01(r.x); oy = u01(r.y);
    }
    float3 d = camera_dir(f.cam, x, f.cam.height - 1 - y, ox, oy);
    float t; int id;
    closest_hit<SMEM, BRUTE>(g, s.n_objects, f.cam.origin, d, f.t_min, t, id);
    obj[i] = id;
    tt[i] = id >= 0 ? t : __int_as_float(0x7f800000);
    if (dirs) { dirs[3 * (size_t)i] = d.x; dirs[3 * (size_t)i + 1] = d.y; dirs[3 * (size_t)i + 2] = d.z; }
}

// The same hook through the constant-bank scan of the shared-memory wavefront (closest_hit_param, MODE 2 / 3 there).
template <bool CHECKED>
__global__ void __launch_bounds__(kBlock) k_trace_primary_param(SceneView s, const __grid_constant__ Frame f, int jitter, float ox, float oy, uint32_t sample,
                                                                int32_t* __restrict__ obj, float* __restrict__ tt, float* __restrict__ dirs,
                                                                const __grid_constant__ BruteParam bp, const __grid_constant__ CullParam cull) {
    extern __shared__ float4 smem[];
    Geo<true> g = stage_scene<true>(s, smem, false);
    uint32_t i = blockIdx.x * kBlock + threadIdx.x;
    if (i >= f.npix) return;
    uint32_t x = i % f.cam.width, y = i / f.cam.width;
    if (jitter) {
        uint4 r = rng_block(i, sample, 0xFFFFFFFFu, 0u, f.k0, f.k1);
        ox = u01(r.x); oy = u01(r.y);
    }
    float3 d = camera_dir(f.cam, x, f.cam.height - 1 - y, ox, oy);
    float t; int id;
    closest_hit_param_culled<CHECKED>(bp, cull, g, x, f.cam.height - 1 - y, f.cam.origin, d, f.t_min, t, id);   // what the wavefront's camera rays run
    obj[i] = id;
    tt[i] = id >= 0 ? t : __int_as_float(0x7f800000);
    if (dirs) { dirs[3 * (size_t)i] = d.x; dirs[3 * (size_t)i + 1] = d.y; dirs[3 * (size_t)i + 2] = d.z; }
}

// Parity hook for the secondary-ray scan (PT_TRAVERSAL_BRUTE_FAST): closest_hit_fast applied to primary rays.
__global__ void __launch_bounds__(kBlock) k_trace_primary_fast(SceneView s, const __grid_constant__ Frame f, int jitter, float ox, float oy, uint32_t sample,
                                                               int32_t* __restrict__ obj, float* __restrict__ tt, float* __restrict__ dirs,
                                                               const __grid_constant__ FastParam fp, uint32_t n_fast_pairs, uint32_t n_fast_spheres) {
    extern __shared__ float4 smem[];
    Geo<true> g = stage_scene<true>(s, smem, false);
    uint32_t i = blockIdx.x * kBlock + threadIdx.x;
    if (i >= f.npix) return;
    uint32_t x = i % f.cam.width, y = i / f.cam.width;
    if (jitter) {
        uint4 r = rng_block(i, sample, 0xFFFFFFFFu, 0u, f.k0, f.k1);
        ox = u01(r.x); oy = u01(r.y);
    }
    float3 d = camera_dir(f.cam, x, f.cam.height - 1 - y, ox, oy);
    float t; int id;
    closest_hit_fast(fp, n_fast_pairs, n_fast_spheres, g, f.cam.origin, d, f.t_min, t, id);
    obj[i] = id;
    tt[i] = id >= 0 ? t : __int_as_float(0x7f800000);
    if (dirs) { dirs[3 * (size_t)i] = d.x; dirs[3 * (size_t)i + 1] = d.y; dirs[3 * (size_t)i + 2] = d.z; }
}

// pt_render_moments: total += batch, moment2 += batch^2, batch = 0
__global__ void k_fold_moments(float4* __restrict__ batch, float4* __restrict__ total, float4* __restrict__ m2, uint32_t npix) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    const float4 b = batch[i];
    float4 t = total[i], q = m2[i];
    t.x += b.x; t.y += b.y; t.z += b.z;
    q.x = fmaf(b.x, b.x, q.x); q.y = fmaf(b.y, b.y, q.y); q.z = fmaf(b.z, b.z, q.z);
    total[i] = t; m2[i] = q;
    batch[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}
__global__ void k_moments_out(const float4* __restrict__ m2, uint32_t npix, float* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    const float4 q = m2[i];
    out[3 * (size_t)i] = q.x; out[3 * (size_t)i + 1] = q.y; out[3 * (size_t)i + 2] = q.z;
}

struct Mode { bool smem, brute; size_t smem_bytes; };
Mode pick_mode(const SceneDev& s, uint32_t traversal) {
    Mode m;
    // A tree over fewer than two objects has no inner node to walk (lbvh.cu: k_trivial_tree exists for pt_bvh_download only):
    // such scenes are always scanned.
    m.brute = traversal >= PT_TRAVERSAL_BRUTE || (traversal == PT_TRAVERSAL_AUTO && s.n_objects <= 16) || s.n_objects < 2;
    m.smem = s.n_objects <= kMaxSmemPrims;
    m.smem_bytes = !m.smem ? 0 : (m.brute && s.brute) ? sizeof(float4) * 5 * (size_t)s.n_brute
                                                     : sizeof(float4) * ((m.brute ? 0 : (size_t)s.n_nodes * 4) + (size_t)s.n_objects * 3);
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

bool primary_cull_rects_host(const pt_shape* shapes, uint32_t n, const pt_camera& cam, int32_t* rects, int32_t* entry_of_object, uint32_t* n_entries) {
    if (n == 0 || n > 1024) return false;
    SceneDev sc;                                          // host fields only: what make_cull_param reads
    sc.n_objects = n;
    sc.shapes_host.assign(shapes, shapes + n);
    uint32_t pairs = 0, entries = 0;
    for (uint32_t i = 0; i < n; ++entries) {              // the pairing rule of pt_scene_create / k_brute_table
        const bool pair = i + 1 < n && shapes[i].kind == PT_SHAPE_TRIANGLE && shapes[i + 1].kind == PT_SHAPE_TRIANGLE;
        pairs += pair; i += pair ? 2 : 1;
    }
    if (pairs > (uint32_t)kParamPairs) return false;
    sc.n_pairs = pairs; sc.n_brute = entries;
    const CullParam cp = make_cull_param(sc, cam, true);
    auto put = [&](uint32_t e, int4 r) {
        const bool empty = r.x == 0x7fffffff;
        const bool full = r.z == 0x7fffffff;
        rects[4 * e] = empty ? 1 : r.x; rects[4 * e + 1] = empty ? 1 : r.y;
        rects[4 * e + 2] = empty ? 0 : full ? (int32_t)cam.width - 1 : r.x + r.z; rects[4 * e + 3] = empty ? 0 : full ? (int32_t)cam.height - 1 : r.y + r.w;
    };
    uint32_t ep = 0, es = 0;
    for (uint32_t i = 0; i < n;) {
        const bool pair = i + 1 < n && shapes[i].kind == PT_SHAPE_TRIANGLE && shapes[i + 1].kind == PT_SHAPE_TRIANGLE;
        if (pair) {
            put(ep, cp.pair[(size_t)kParamPairs - pairs + ep]);
            entry_of_object[i] = entry_of_object[i + 1] = (int32_t)ep; ++ep; i += 2;
        } else {
            const uint32_t e = pairs + es;
            put(e, es < (uint32_t)kCullSingles ? cp.single[es] : make_int4(0, 0, 0x7fffffff, 0x7fffffff));
            entry_of_object[i] = (int32_t)e; ++es; i += 1;
        }
    }
    *n_entries = entries;
    return true;
}

int resolve_launch(pt_context* ctx, const float4* d_accum, uint32_t w, uint32_t h, uint32_t spp, float* d_rad, uint8_t* d_rgba) {
    uint32_t npix = w * h;
    k_resolve<<<(npix + 255) / 256, 256, 0, ctx->stream>>>(d_accum, npix, (float)spp, d_rad, reinterpret_cast<uchar4*>(d_rgba));
    PT_CUDA_CHECK(ctx, cudaGetLastError());
    return PT_OK;
}

int fold_moments_launch(pt_context* ctx, float4* d_batch, float4* d_total, float4* d_moment2, uint32_t npix) {
    k_fold_moments<<<(npix + 255) / 256, 256, 0, ctx->stream>>>(d_batch, d_total, d_moment2, npix);
    PT_CUDA_CHECK(ctx, cudaGetLastError());
    return PT_OK;
}
int moments_out_launch(pt_context* ctx, const float4* d_moment2, uint32_t npix, float* d_out_rgb) {
    k_moments_out<<<(npix + 255) / 256, 256, 0, ctx->stream>>>(d_moment2, npix, d_out_rgb);
    PT_CUDA_CHECK(ctx, cudaGetLastError());
    return PT_OK;
}

int wavefront_trace_primary(pt_context* ctx, const SceneDev& scene, const pt_camera& cam, const pt_render_params& prm, int jitter,
                            float ox, float oy, uint32_t sample, int32_t* d_obj, float* d_t, float* d_dir, uint64_t* launches) {
    SceneView sv = make_scene_view(scene);
    Frame f = make_frame(cam, prm, nullptr);
    Mode mode = pick_mode(scene, prm.traversal);
    uint32_t grid = (f.npix + kBlock - 1) / kBlock;
    const int scan = (mode.brute && mode.smem && scene.brute) ? pick_scan_mode(scene, cam, prm.traversal) : 0;
    if (prm.traversal == PT_TRAVERSAL_BRUTE_FAST) {
        if (scan != 4) { ctx->err = "PT_TRAVERSAL_BRUTE_FAST: this scene has no planar-entry table (too many triangles, or outside the FP32 coordinate domain)"; return PT_ERR_INVALID; }
        const FastParam fp = make_fast_param(scene);
        int rc = set_smem(ctx, k_trace_primary_fast, mode.smem_bytes); if (rc) return rc;
        k_trace_primary_fast<<<grid, kBlock, mode.smem_bytes, ctx->stream>>>(sv, f, jitter, ox, oy, sample, d_obj, d_t, d_dir, fp, scene.n_fast_pairs, scene.n_fast_spheres);
        PT_CUDA_CHECK(ctx, cudaGetLastError());
        if (launches) *launches += 1;
        return PT_OK;
    }
    if (scan >= 2) {   // the constant-bank scan the shared-memory wavefront runs on this scene
        const BruteParam bp = make_brute_param(scene);
        // the pixel-rectangle cull holds for ray offsets inside the pixel cell (the jitter is in (0, 1))
        const CullParam cull = make_cull_param(scene, cam, jitter || (ox >= 0.0f && ox <= 1.0f && oy >= 0.0f && oy <= 1.0f));
        if (scan == 2) {   // (scan 3 and 4 trace primary rays with the same in-domain pair test)
            int rc = set_smem(ctx, k_trace_primary_param<true>, mode.smem_bytes); if (rc) return rc;
            k_trace_primary_param<true><<<grid, kBlock, mode.smem_bytes, ctx->stream>>>(sv, f, jitter, ox, oy, sample, d_obj, d_t, d_dir, bp, cull);
        } else {
            int rc = set_smem(ctx, k_trace_primary_param<false>, mode.smem_bytes); if (rc) return rc;
            k_trace_primary_param<false><<<grid, kBlock, mode.smem_bytes, ctx->stream>>>(sv, f, jitter, ox, oy, sample, d_obj, d_t, d_dir, bp, cull);
        }
        PT_CUDA_CHECK(ctx, cudaGetLastError());
        if (launches) *launches += 1;
        return PT_OK;
    }
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
    SceneView sv = make_scene_view(scene);
    Frame f = make_frame(rl.cam, rl.prm, rl.d_accum);
    Mode mode = pick_mode(scene, rl.prm.traversal);
    uint64_t launches = 0, iters = 0;
    if (f.total_work == 0) { if (stats) { stats->samples = 0; } return PT_OK; }
    // small scenes: queues in shared memory, one persistent launch (smem_wavefront.cu)
    if (!(rl.prm.flags & PT_FLAG_HBM_QUEUES) && smem_wavefront_applicable(ctx, scene, mode.brute))
        return smem_wavefront_render(ctx, rl, mode.brute, stats);

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
    // LBVH scenes: a resident grid of warps streams the queues (k_trace_*_stream)
    const bool stream_bvh = !mode.brute;
    const bool count = stream_bvh && (rl.prm.flags & PT_FLAG_COUNTERS) != 0;
    uint32_t stream_grid = 0, stream_chunk = 0;
    const int stream_block = mode.smem ? kStreamBlockSmem : kBlock;
    if (stream_bvh) {
        int occ_c = 0, occ_s = 0;
        auto setup = [&](auto kc, auto ks, size_t bytes) -> int {
            int rc2 = set_smem(ctx, kc, bytes); if (rc2) return rc2;
            rc2 = set_smem(ctx, ks, bytes); if (rc2) return rc2;
            PT_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_c, kc, stream_block, bytes));
            PT_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_s, ks, stream_block, bytes));
            return PT_OK;
        };
        if (mode.smem) rc = count ? setup(k_trace_closest_stream<true, true>, k_trace_shadow_stream<true, true>, mode.smem_bytes)
                                  : setup(k_trace_closest_stream<true, false>, k_trace_shadow_stream<true, false>, mode.smem_bytes);
        else rc = count ? setup(k_trace_closest_stream<false, true>, k_trace_shadow_stream<false, true>, 0)
                        : setup(k_trace_closest_stream<false, false>, k_trace_shadow_stream<false, false>, 0);
        if (rc) return rc;
        const int occ = occ_c < occ_s ? occ_c : occ_s;
        if (occ < 1) { ctx->err = "stream traversal kernels do not fit on an SM"; return PT_ERR_CUDA; }
        stream_grid = (uint32_t)ctx->sm_count * (uint32_t)occ;        // one resident wave; the queue is handed out dynamically
        stream_chunk = mode.smem ? kStreamChunkSmem : kStreamChunkL2;
    }
    const uint32_t sgrid = P / kShadeSlots;
    f.chunk = (f.total_work + sgrid - 1) / sgrid;
    const bool profile = (rl.prm.flags & PT_FLAG_PROFILE) != 0;
    float ms_k[PT_K_COUNT] = {0, 0, 0, 0, 0};
    uint32_t n_k[PT_K_COUNT] = {0, 0, 0, 0, 0};
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
    // The shadow rays of a bounce and the closest hits of the next one are independent once the bounce is shaded: the shadow
    // launch goes to a second stream and the two resident grids fill each other's tails (each launch ends with a few long
    // walks on an otherwise idle GPU).  Off under PT_FLAG_PROFILE, whose per-class event times assume serial launches.
    const bool overlap = mis && stream_bvh && !profile;
    cudaStream_t st_sh = st;
    if (overlap) {
        if (!ctx->stream2) {
            PT_CUDA_CHECK(ctx, cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
            PT_CUDA_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_shaded, cudaEventDisableTiming));
            PT_CUDA_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_shadowed, cudaEventDisableTiming));
        }
        st_sh = ctx->stream2;
    }
    bool shadow_pending = false;
    bool done = false;
    uint64_t nodes_visited = 0, prim_tests = 0;
    unsigned long long* walk = reinterpret_cast<unsigned long long*>(ctx->d_counters + C_WALK);
    while (!done) {
        PT_CUDA_CHECK(ctx, cudaMemsetAsync(ctx->d_counters + C_SHADOW_BASE, 0, 8 * (C_TOTAL - C_SHADOW_BASE), st));
        for (int it = 0; it < kItersPerBatch; ++it) {
            // chunk counters of this iteration's two stream launches: words 0 and 1 of its (otherwise unused) shadow line
            unsigned long long* cur_c = reinterpret_cast<unsigned long long*>(ctx->d_counters + C_SHADOW_BASE + it * C_STRIDE);
            unsigned long long* cur_s = cur_c + 1;
#define TRACE(S, B) k_trace_closest<S, B><<<grid, kBlock, mode.smem_bytes, st>>>(pool, sv, f.t_min)
            prof_begin(PT_K_TRACE_CLOSEST);
            if (stream_bvh) {
                if (mode.smem && count) k_trace_closest_stream<true, true><<<stream_grid, kStreamBlockSmem, mode.smem_bytes, st>>>(pool, sv, f.t_min, stream_chunk, cur_c, walk);
                else if (mode.smem) k_trace_closest_stream<true, false><<<stream_grid, kStreamBlockSmem, mode.smem_bytes, st>>>(pool, sv, f.t_min, stream_chunk, cur_c, walk);
                else if (count) k_trace_closest_stream<false, true><<<stream_grid, kBlock, 0, st>>>(pool, sv, f.t_min, stream_chunk, cur_c, walk);
                else k_trace_closest_stream<false, false><<<stream_grid, kBlock, 0, st>>>(pool, sv, f.t_min, stream_chunk, cur_c, walk);
            } else PT_DISPATCH(mode, TRACE);
            prof_end();
#undef TRACE
            prof_begin(PT_K_SHADE);
            if (overlap && shadow_pending) PT_CUDA_CHECK(ctx, cudaStreamWaitEvent(st, ctx->ev_shadowed, 0));   // shading rewrites the shadow queue
            k_shade<<<sgrid, kShadeBlock, 0, st>>>(pool, sv, f, ctx->d_counters, it);
            prof_end();
            launches += 2;
            if (mis) {
                if (overlap) {
                    PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_shaded, st));
                    PT_CUDA_CHECK(ctx, cudaStreamWaitEvent(st_sh, ctx->ev_shaded, 0));
                }
                prof_begin(PT_K_TRACE_SHADOW);
#define SHADOW(S, B) k_trace_shadow<S, B><<<grid, kBlock, mode.smem_bytes, st>>>(pool, sv, f.t_min, f.accum)
                if (stream_bvh) {
                    if (mode.smem && count) k_trace_shadow_stream<true, true><<<stream_grid, kStreamBlockSmem, mode.smem_bytes, st_sh>>>(pool, sv, f.t_min, f.accum, stream_chunk, cur_s, walk);
                    else if (mode.smem) k_trace_shadow_stream<true, false><<<stream_grid, kStreamBlockSmem, mode.smem_bytes, st_sh>>>(pool, sv, f.t_min, f.accum, stream_chunk, cur_s, walk);
                    else if (count) k_trace_shadow_stream<false, true><<<stream_grid, kBlock, 0, st_sh>>>(pool, sv, f.t_min, f.accum, stream_chunk, cur_s, walk);
                    else k_trace_shadow_stream<false, false><<<stream_grid, kBlock, 0, st_sh>>>(pool, sv, f.t_min, f.accum, stream_chunk, cur_s, walk);
                } else PT_DISPATCH(mode, SHADOW);
#undef SHADOW
                prof_end();
                launches++;
                if (overlap) { PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_shadowed, st_sh)); shadow_pending = true; }
            }
        }
        iters += kItersPerBatch;
        if (overlap && shadow_pending) { PT_CUDA_CHECK(ctx, cudaStreamWaitEvent(st, ctx->ev_shadowed, 0)); shadow_pending = false; }
        PT_CUDA_CHECK(ctx, cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 8 * C_TOTAL, cudaMemcpyDeviceToHost, st));
        PT_CUDA_CHECK(ctx, cudaStreamSynchronize(st));
        prof_collect();
        done = ctx->h_counters[C_ALIVE_BASE + (kItersPerBatch - 1) * C_STRIDE] == 0;
        nodes_visited += ctx->h_counters[C_WALK]; prim_tests += ctx->h_counters[C_WALK + 1];
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
        stats->nodes_visited = nodes_visited; stats->prim_tests = prim_tests; stats->scan_mode = 0xffffffffu;
        for (int k = 0; k < PT_K_COUNT; ++k) { stats->ms_kernel[k] = ms_k[k]; stats->n_kernel[k] = n_k[k]; }
    }
    return PT_OK;
}
