// smem_wavefront.cu — the wavefront sample loop with its queues in SHARED memory (sm_100a, 227 KB per SM).
//
// Same stages as wavefront.cu (K1 generate, K3 closest hit, K4/K6/K7 shade, K5 shadow, K8 accumulate), same device
// functions (pt_wave.cuh), same counter-based RNG — but when the whole scene (primitives, materials, LBVH or scan table) fits
// in a few KB of shared memory, nothing about a path needs to leave the SM between bounces.  Each CTA owns kSwSlots path
// slots whose SoA float4 state (ray origin|eta, direction|pdf, throughput, hit, meta) lives in shared memory, owns a
// contiguous range of the (pixel, sample) work list, and runs the wavefront to completion on its own:
//
//   loop:  T  work list = [slots whose path continues ... | ... slots to re-seed]: the re-seeded ones get a camera ray
//             (K1, camera.rs:139-147) from the CTA's work range, then every entry finds its closest hit (world.rs:270-290)
//             and is appended to the list of its hit class (mirror / diffuse / emissive / miss): match + popc and one shared
//             atomic per class present in the warp
//          A  warps pull 32-slot tiles of the class lists, expensive classes first: shade_vertex() (rendering.rs:34-142),
//             the vertex's shadow ray traced in line (rendering.rs:62-80), finished paths accumulated (world.rs:296-313);
//             each slot is appended to the front (continues) or the back (re-seed) of the next work list, so camera-ray
//             generation is dense too
//
// Two barriers per iteration and nothing else: every counter exists twice and the copies alternate, so none is reset while
// a straggler may still read it.  All warps run the same phase at the same time ON PURPOSE: the instruction cache is 32 KB
// (phase T is 20 KB of code, the hot part of phase A 27 KB).  A barrier-free variant (six ring queues, warps taking whichever
// full tile exists) ran at a third of the speed with a 57 % instruction-cache hit rate, and at half the speed when the warps
// were steered to take the same kind of item (trace / shade rounds without barriers): per-item queue bookkeeping by one lane,
// fences and sentinel polling cost more than the barriers they remove.  Two CTAs per SM lose 14 %.  Two populations walking
// a fixed segment sequence T0 T1 A0 A1 with completion counters instead of barriers (3.7 % waiting) lose 15 % as well.
//
// HBM traffic is what the algorithm needs: one float4 atomic add per finished path / unoccluded shadow ray.
// CTAs never communicate, so one launch renders the frame: no per-bounce launches, no host round trips.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "pt_wave.cuh"

using namespace ptw;

namespace {

// One CTA of 768 threads per SM owning 3072 slots: the class sort sees a window wide enough that almost every 32-slot tile
// is a single class (measured: 512-slot windows lose ~12 % to mixed tiles).  Shapes measured in round 2 (threads x slots per
// thread, Gsamples/s at 128 spp): 896 x 3 (72 registers) 6.65, 768 x 4 (80) 6.75, 640 x 5 (96) 6.76, 704 x 4 6.61, 640 x 4 6.58,
// 576 x 5 6.44, 800 x 4 6.25, 1024 x 3 (64 registers, spills) 6.23, 960 x 3 6.05.  Phase T with dynamic 32-entry tiles (like
// phase A) instead of a static entry per thread and group: 6.69 against 6.75.
#ifndef PT_SW_BLOCK
#define PT_SW_BLOCK 768
#endif
#ifndef PT_SW_SPT
#define PT_SW_SPT 4
#endif
#ifndef PT_SW_MIN_BLOCKS
#define PT_SW_MIN_BLOCKS 1
#endif
constexpr int kSwBlock = PT_SW_BLOCK, kSwSpt = PT_SW_SPT, kSwSlots = kSwBlock * kSwSpt, kSwMinBlocks = PT_SW_MIN_BLOCKS;
static_assert(kSwSlots < 65536, "slot ids are 16-bit and the two list lengths share one 32-bit counter");

// Shared-memory atomic add issued as is.  (For atomicAdd() on shared memory the compiler emits a warp-aggregation prologue —
// vote, popc, leader election — even where one elected lane calls it; here the aggregation is done by hand already.)
PT_DEV uint32_t atom_add_shared(uint32_t* p, uint32_t v) {
    uint32_t old;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
    return old;
}

// A 64-bit add issued by ONE elected lane of a converged warp, result broadcast to all lanes.  Written with elect.sync because
// ptxas wraps every shared-memory atomic it cannot prove single-lane in a vote / find-leader / popc / shuffle sequence (17
// instructions around `if (lane == 0) atom`); after ELECT it emits the bare ATOMS.
PT_DEV unsigned long long warp_atom_add_shared64(unsigned long long* p, unsigned long long v) {
    unsigned long long old = 0;
    uint32_t leader;
    asm volatile(
        "{\n .reg .pred q;\n elect.sync %1|q, 0xffffffff;\n @q atom.shared.add.u64 %0, [%2], %3;\n}"
        : "+l"(old), "=r"(leader) : "r"((uint32_t)__cvta_generic_to_shared(p)), "l"(v) : "memory");
    return __shfl_sync(0xffffffffu, old, leader);
}

struct SwOut {
    uint4* block_stats;       // per CTA: primary, shadow, scattered, vertices
    uint32_t* block_iters;    // per CTA: wavefront iterations it ran
    unsigned long long* block_clk;   // per CTA (PROF only): cycles in phases T+S, A, H
    unsigned long long* work_next;   // next unclaimed (pixel, sample) work item of the frame (zeroed before the launch)
};

// Bytes of dynamic shared memory for a scene (must match the carve-up in the kernel).
size_t sw_smem_bytes(uint32_t n_objects, uint32_t n_nodes, uint32_t n_lights, bool brute) {
    size_t b = 0;
    b += 16 * (size_t)kSwSlots * 3;                 // ro, rd, beta
    b += 8 * (size_t)kSwSlots * 2;                  // hit, meta
    b += 16 * ((brute ? 0 : (size_t)n_nodes * 4) + (size_t)n_objects * 4 + (size_t)n_objects * 2);
    b += 4 * (((size_t)n_lights + 3) & ~size_t(3));
    b += 2 * 3 * (((size_t)kSwSlots + 7) & ~size_t(7));   // two class-list arrays, next work list
    b += brute ? 16 * 5 * (size_t)n_objects : 0;    // brute-force scan table (at most one entry per object)
    return b;
}

// MODE 0: LBVH from shared memory; 1: scan table from shared memory; 2: its triangle pairs from the kernel parameter `bp`
// (constant bank, unrolled) and only the unpaired primitives from shared memory; 3: as 2 for scenes and cameras inside the
// FP32 coordinate domain (SceneDev::fp32_domain): the pair test without the reciprocal's range check and without `u <= 1`;
// 4: as 3 for primary rays, the planar-entry scan from the kernel parameter `fp` for shadow and scattered rays.
template <int MODE, bool PROF>
__global__ void __launch_bounds__(kSwBlock, kSwMinBlocks) k_smem_wavefront(SceneView s, const __grid_constant__ Frame f, SwOut out,
                                                                           const __grid_constant__ BruteParam bp,
                                                                           const __grid_constant__ FastParam fp, uint32_t n_fast_pairs, uint32_t n_fast_spheres,
                                                                           const __grid_constant__ CullParam cull) {
    constexpr bool BRUTE = MODE != 0;
    constexpr int SPT = kSwSpt;
    extern __shared__ float4 smem[];
    float4* s_ro = smem;
    float4* s_rd = s_ro + kSwSlots;
    float4* s_beta = s_rd + kSwSlots;
    uint2* s_hit = reinterpret_cast<uint2*>(s_beta + kSwSlots);
    uint2* s_meta = s_hit + kSwSlots;
    // the lists come before the scene so that their addresses are compile-time offsets (with the scene first the compiler
    // rebuilt `base + n_objects * 96 + ...` at every use: 2.8 % of all instructions)
    constexpr uint32_t LST = (kSwSlots + 7) & ~7;
    uint16_t* s_ord_a = reinterpret_cast<uint16_t*>(s_meta + kSwSlots);                     // class lists: mirror ->  <- diffuse
    uint16_t* s_ord_b = s_ord_a + LST;                                                      //              emissive ->  <- miss
    uint16_t* s_list = s_ord_b + LST;                                                       // next work list (phase A writes)
    float4* s_nodes = reinterpret_cast<float4*>(s_list + LST);
    float4* s_prims = s_nodes + (BRUTE ? 0 : s.n_nodes * 4);
    float4* s_mats = s_prims + s.n_objects * 4;
    float4* s_brute = s_mats + s.n_objects * 2;
    uint32_t* s_lights = reinterpret_cast<uint32_t*>(s_brute + (BRUTE ? s.n_brute * 5 : 0));
    // Counters alternate between two copies so that none is reset while a straggler may still read it: iteration i counts the
    // classes in s_cls[i & 1] and reads its work list sizes from s_len[i & 1]; phase A of iteration i fills s_len[(i + 1) & 1].
    __shared__ uint32_t s_cls[2][4];           // entries per class list
    // next work list: continuing entries (front) | entries to re-seed (back) << 16, and in the upper half the tile counter of the
    // phase A that fills it — one 64-bit atomic per tile appends the finished tile's slots AND fetches the next tile
    __shared__ unsigned long long s_len[2];
    __shared__ unsigned long long s_tot[3];    // statistics of this CTA: vertices, scattered rays, shadow rays
    __shared__ unsigned long long s_clk[3];    // PROF: t_prev, T+S, A

    const uint32_t tid = threadIdx.x, lane = tid & 31, lt_mask = (1u << lane) - 1u;
    const bool mis = f.strategy == PT_STRATEGY_MIS;
    // Work is handed to the CTAs in chunks from one global counter (guided self-scheduling: a quarter of an even share of what
    // is left, never less than one window), so no CTA is left finishing a static range while the others idle.
    // A CTA holds TWO ranges, the one it is consuming and the next, so that an iteration never runs short of work items
    // before the frame does (a slot that finds no work item leaves the window for good).
    __shared__ unsigned long long s_rng[2][4];       // [parity][lo, hi, lo2, hi2): current and next range of the work list
    __shared__ uint32_t s_org[2][4];                 // [parity][sample, pixel of lo; sample, pixel of lo2]: work item w = sample * npix + pixel
    auto grab = [&](unsigned long long* dst, uint32_t* org) {       // one thread
        const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(out.work_next);
        const unsigned long long left = f.total_work > cur ? f.total_work - cur : 0ull;
        unsigned long long size = left / (4ull * gridDim.x);       // measured at 128 spp: /2 5.7, /4 6.40, one window 6.39, static 6.27 Gsamples/s
        if (size < (unsigned long long)kSwSlots) size = kSwSlots;
        if (size > (1ull << 30)) size = 1ull << 30;                // offsets inside a range stay 32-bit
        const unsigned long long c = atomicAdd(out.work_next, size);
        dst[0] = c < f.total_work ? c : f.total_work;
        dst[1] = c + size < f.total_work ? c + size : f.total_work;
        org[0] = (uint32_t)(dst[0] / f.npix); org[1] = (uint32_t)(dst[0] % f.npix);   // the only 64-bit divisions: once per range
    };

    // ---- stage the scene once
    if (!BRUTE) for (uint32_t i = tid; i < s.n_nodes * 4; i += kSwBlock) s_nodes[i] = __ldg(s.nodes + i);
    for (uint32_t i = tid; i < s.n_objects * 4; i += kSwBlock) s_prims[i] = __ldg(s.prims + i);
    for (uint32_t i = tid; i < s.n_objects * 2; i += kSwBlock) s_mats[i] = __ldg(s.mats + i);
    for (uint32_t i = tid; i < s.n_lights; i += kSwBlock) s_lights[i] = __ldg(s.light_ids + i);
    Geo<true, 4> g; g.nodes = s_nodes; g.prims = s_prims;
    if (BRUTE && s.brute) {
        for (uint32_t i = tid; i < s.n_brute * 5; i += kSwBlock) s_brute[i] = __ldg(s.brute + i);
        g.brute = s_brute; g.n_brute = s.n_brute; g.n_pairs = s.n_pairs;
    }
    // MODE 4: primary rays keep the bit-exact scan (ids and t are pinned to the FP32 oracle); shadow and scattered rays take
    // the planar-entry scan (pt_wave.cuh: fast_pair)
    // A camera ray that misses the padded scene bounds misses every primitive (the slab test is conservative, the one the LBVH
    // walk relies on): at 1920x1080 ~40 % of the default scene's primary rays end here, warp-uniformly (re-seeding is dense).
    // MODE >= 2: the per-entry pixel rectangles of `cull` do the same per pair of triangles (closest_hit_param_culled).
    auto trace_closest = [&](float3 ro, float3 rd, bool primary, uint32_t px, uint32_t py, float& t, int& id) {
        if (MODE >= 2 && primary) { closest_hit_param_culled<MODE == 2>(bp, cull, g, px, py, ro, rd, f.t_min, t, id); return; }
        if (MODE == 1 && primary) {
            const float3 inv = v3(__frcp_rn(rd.x), __frcp_rn(rd.y), __frcp_rn(rd.z));
            float tn;
            if (!slab(f.blo, f.bhi, ro, inv, f.t_min, FLT_MAX, tn)) { t = FLT_MAX; id = -1; return; }
        }
        if (MODE == 4 && !primary) closest_hit_fast(fp, n_fast_pairs, n_fast_spheres, g, ro, rd, f.t_min, t, id);
        else if (MODE >= 2) closest_hit_param<MODE == 2>(bp, g, ro, rd, f.t_min, t, id);
        else closest_hit<true, BRUTE>(g, s.n_objects, ro, rd, f.t_min, t, id);
    };
    auto trace_any = [&](float3 ro, float3 rd, float t_max) -> bool {
        if (MODE == 4) return any_hit_fast(fp, n_fast_pairs, n_fast_spheres, g, ro, rd, f.t_min, t_max);
        if (MODE >= 2) return any_hit_param<MODE == 2>(bp, g, ro, rd, f.t_min, t_max);
        return any_hit<true, BRUTE>(g, s.n_objects, ro, rd, f.t_min, t_max);
    };
    const SceneShared sc{s_prims, s_mats, s_lights, s.n_lights};

    // ---- the first work list re-seeds every slot
    for (uint32_t i = tid; i < (uint32_t)kSwSlots; i += kSwBlock) s_list[i] = (uint16_t)i;
    if (tid == 0) {
        s_len[0] = (unsigned long long)kSwSlots << 16; s_len[1] = 0ull;
        for (int c = 0; c < 4; ++c) s_cls[0][c] = s_cls[1][c] = 0;
        s_tot[0] = s_tot[1] = s_tot[2] = 0ull;
        if (PROF) { s_clk[0] = clock64(); s_clk[1] = s_clk[2] = 0; }
        grab(s_rng[0], s_org[0]); grab(s_rng[0] + 2, s_org[0] + 2);
    }
    // Statistics: ONE packed per-lane register (vertices | scattered rays << 10 | shadow rays << 20; a lane adds at most kSwSpt of
    // each per iteration), flushed into 64-bit shared totals every 64 iterations.  Three separate counters were spilled to local
    // memory by ptxas at the 80-register cap and every tile waited on their reloads.
    static_assert(64 * kSwSpt < 1024, "the packed per-lane statistics hold 10 bits per field between flushes");
    uint32_t c_pack = 0, iters = 0;
    auto flush_stats = [&]() {
        const uint32_t v = __reduce_add_sync(0xffffffffu, c_pack & 1023u), sc2 = __reduce_add_sync(0xffffffffu, (c_pack >> 10) & 1023u),
                       sh = __reduce_add_sync(0xffffffffu, c_pack >> 20);
        if (lane == 0) { atomicAdd(&s_tot[0], (unsigned long long)v); atomicAdd(&s_tot[1], (unsigned long long)sc2); atomicAdd(&s_tot[2], (unsigned long long)sh); }
        c_pack = 0;
    };
    uint64_t n_primary = 0;                                          // work items this CTA started (uniform)
    __syncthreads();

    while (true) {
        // ---- phase T: entries [0, n_cont) continue a path; the last n_regen entries take new work items base, base+1, ...
        // from the back (entry kSwSlots-1 first) while the CTA's range lasts.  Entries in between, and re-seed entries past
        // the end of the range, are dead: they join no class list and phase A never touches them.
        const uint32_t par = iters & 1;
        const uint32_t n_cont = (uint32_t)s_len[par] & 0xffffu, n_regen = ((uint32_t)s_len[par] >> 16) & 0xffffu;
        const uint64_t base = s_rng[par][0], base2 = s_rng[par][2];
        const uint32_t avail1 = (uint32_t)(s_rng[par][1] - base), avail2 = (uint32_t)(s_rng[par][3] - base2);   // each range is at most 2^30 items
        const uint64_t avail = (uint64_t)avail1 + avail2;
        const uint32_t n_new = avail < n_regen ? (uint32_t)avail : n_regen;
        const uint32_t smp1 = s_org[par][0], pix1 = s_org[par][1], smp2 = s_org[par][2], pix2 = s_org[par][3];
        if (n_cont + n_new == 0) break;                              // uniform: both ranges are empty only when the frame's work list is
        if (tid == 0) {
            s_len[par ^ 1] = 0ull;
            unsigned long long* nx = s_rng[par ^ 1];
            uint32_t* no = s_org[par ^ 1];
            auto advance = [&](uint32_t smp, uint32_t pix, uint32_t by, uint32_t* o) {   // by <= kSwSlots
                const uint32_t p = pix + by, q = p / f.npix;
                o[0] = smp + q; o[1] = p - q * f.npix;
            };
            if (n_new >= avail1) {                                   // current range used up: the next one becomes current, a new next is fetched
                nx[0] = base2 + (n_new - avail1); nx[1] = s_rng[par][3];
                advance(smp2, pix2, n_new - avail1, no);
                grab(nx + 2, no + 2);                                // (empty once the frame is handed out)
            } else {
                nx[0] = base + n_new; nx[1] = s_rng[par][1]; nx[2] = base2; nx[3] = s_rng[par][3];
                advance(smp1, pix1, n_new, no); no[2] = smp2; no[3] = pix2;
            }
        }
#pragma unroll 1                                                     // one copy of the scan: the instruction cache is 32 KB
        for (int k = 0; k < SPT; ++k) {
            const uint32_t e = k * kSwBlock + tid;
            uint32_t c = CLS_DEAD, j = 0;
            const bool cont = e < n_cont, seed = e >= kSwSlots - n_new;
            if (cont || seed) {
                j = s_list[e];
                PT_CHECK(j < (uint32_t)kSwSlots);
                float4 ro, rd;
                uint32_t px = 0, py = 0;                              // pixel column and flipped row of a camera ray
                if (seed) {
                    uint2 m;
                    const uint32_t k_new = kSwSlots - 1 - e;
                    const bool first = k_new < avail1;
                    make_primary_at(f, first ? smp1 : smp2, first ? pix1 + k_new : pix2 + (k_new - avail1), ro, rd, m, px, py);
                    s_ro[j] = ro; s_rd[j] = rd; s_meta[j] = m; s_beta[j] = make_float4(1.f, 1.f, 1.f, 0.f);
                } else {
                    ro = s_ro[j]; rd = s_rd[j];
                }
                float t; int id;
                trace_closest(v3(ro), v3(rd), seed, px, py, t, id);
                PT_CHECK(id < (int)s.n_objects);
                s_hit[j] = make_uint2(__float_as_uint(t), (uint32_t)id);
                if (id < 0) c = CLS_MISS;
                else {
                    uint32_t kind = __float_as_uint(s_mats[2 * id].w);
                    c = kind == PT_MATK_MIRROR ? CLS_MIRROR : kind == PT_MATK_EMISSIVE ? CLS_EMISSIVE : CLS_DIFFUSE;
                }
            }
            // append to the class list: one shared atomic per class present in the warp (four ballots + two packed atomics by an
            // elected lane were measured instead: 6.51 against 6.66 Gsamples/s)
            const uint32_t same = __match_any_sync(0xffffffffu, c);
            const uint32_t rank = __popc(same & lt_mask);
            uint32_t pos = 0;
            if (rank == 0 && c != CLS_DEAD) pos = atom_add_shared(&s_cls[par][c], (uint32_t)__popc(same));
            pos = __shfl_sync(0xffffffffu, pos, __ffs(same) - 1) + rank;
            if (c != CLS_DEAD) {
                PT_CHECK(pos < (uint32_t)kSwSlots);
                uint16_t* ord = (c == CLS_MIRROR || c == CLS_DIFFUSE) ? s_ord_a : s_ord_b;
                ord[(c == CLS_MIRROR || c == CLS_EMISSIVE) ? pos : kSwSlots - 1 - pos] = (uint16_t)j;
            }
        }
        __syncthreads();
        if (PROF && tid == 0) { unsigned long long t = clock64(); s_clk[1] += t - s_clk[0]; s_clk[0] = t; }
        if (tid < 4) s_cls[par ^ 1][tid] = 0;
        const uint32_t n_mir = s_cls[par][CLS_MIRROR], n_dif = s_cls[par][CLS_DIFFUSE], n_emi = s_cls[par][CLS_EMISSIVE], n_mis = s_cls[par][CLS_MISS];
        const uint32_t i0 = (n_mir + 31) / 32, i1 = i0 + (n_dif + 31) / 32, i2 = i1 + (n_emi + 31) / 32, i3 = i2 + (n_mis + 31) / 32;

        // ---- phase A: dynamic tiles of the class-sorted list
        unsigned long long got = warp_atom_add_shared64(&s_len[par ^ 1], 1ull << 32);
        while (true) {
            const uint32_t tile = (uint32_t)(got >> 32);
            if (tile >= i3) break;
            uint32_t idx, cnt; const uint16_t* ord; bool fwd;         // expensive tiles first: mirror, diffuse, emissive, miss
            if (tile < i0)      { idx = tile * 32 + lane;        cnt = n_mir; ord = s_ord_a; fwd = true; }
            else if (tile < i1) { idx = (tile - i0) * 32 + lane; cnt = n_dif; ord = s_ord_a; fwd = false; }
            else if (tile < i2) { idx = (tile - i1) * 32 + lane; cnt = n_emi; ord = s_ord_b; fwd = true; }
            else                { idx = (tile - i2) * 32 + lane; cnt = n_mis; ord = s_ord_b; fwd = false; }
            const bool alive = idx < cnt;
            const uint32_t j = alive ? ord[fwd ? idx : kSwSlots - 1 - idx] : 0u;
            PT_CHECK(j < (uint32_t)kSwSlots && n_mir + n_dif <= (uint32_t)kSwSlots && n_emi + n_mis <= (uint32_t)kSwSlots);
            const uint2 meta = s_meta[j];
            bool finished = false, emit_shadow = false;
            float3 sh_d = v3(0.f, 0.f, 0.f), sh_c = v3(0.f, 0.f, 0.f), p_hit = v3(0.f, 0.f, 0.f);
            float sh_tmax = 0.f;
            if (alive) {
                VertexOut vo = shade_vertex(sc, f, mis, s_ro[j], s_rd[j], v3(s_beta[j]), s_hit[j], meta);
                finished = vo.finished;
                c_pack += (vo.vertex ? 1u : 0u) + (vo.finished ? 0u : 1u << 10) + (vo.emit_shadow ? 1u << 20 : 0u);
                p_hit = v3(vo.new_ro);                                // a finished path may still owe its last vertex's shadow ray
                emit_shadow = vo.emit_shadow; sh_d = vo.sh_d; sh_c = vo.sh_c; sh_tmax = vo.sh_tmax;
                if (vo.finished) {
                    // K8 accumulate (world.rs:296-313)
                    if (vo.add_L) atomicAdd(f.accum + META_PIXEL(meta), make_float4(vo.Ladd.x, vo.Ladd.y, vo.Ladd.z, 0.f));
                } else {
                    s_ro[j] = vo.new_ro; s_rd[j] = vo.new_rd; s_meta[j] = vo.new_meta;
                    s_beta[j] = make_float4(vo.new_beta.x, vo.new_beta.y, vo.new_beta.z, 0.f);
                }
            }
            // K5 in line: the shadow ray starts at the hit point (the new ray origin) and is traced by the lane that shaded
            // the vertex; unoccluded contributions go straight to the pixel (rendering.rs:62-80).
            if (emit_shadow) {
                if (!trace_any(p_hit, sh_d, sh_tmax)) atomicAdd(f.accum + META_PIXEL(meta), make_float4(sh_c.x, sh_c.y, sh_c.z, 0.f));
            }
            // next work list: continuing slots from the front, slots to re-seed from the back
            {
                const uint32_t fin = __ballot_sync(0xffffffffu, alive && finished), con = __ballot_sync(0xffffffffu, alive && !finished);
                // one atomic for both ends of the list: continuing count in the low half, re-seed count in the high half
                got = warp_atom_add_shared64(&s_len[par ^ 1], (unsigned long long)((uint32_t)__popc(con) | ((uint32_t)__popc(fin) << 16)) | (1ull << 32));
                const uint32_t b = (uint32_t)got, bfin = b >> 16, bcon = b & 0xffffu;
                PT_CHECK(bfin + bcon + __popc(fin) + __popc(con) <= (uint32_t)kSwSlots);
                if (alive) s_list[finished ? kSwSlots - 1 - (bfin + __popc(fin & lt_mask)) : bcon + __popc(con & lt_mask)] = (uint16_t)j;
            }
        }
        ++iters;
        n_primary += n_new;
        if ((iters & 63u) == 0u) flush_stats();
        __syncthreads();
        if (PROF && tid == 0) { unsigned long long t = clock64(); s_clk[2] += t - s_clk[0]; s_clk[0] = t; }
    }

    // ---- bookkeeping
    flush_stats();
    __syncthreads();
    if (tid == 0) {
        out.block_stats[blockIdx.x] = make_uint4((uint32_t)n_primary, (uint32_t)s_tot[2], (uint32_t)s_tot[1], (uint32_t)s_tot[0]);
        out.block_iters[blockIdx.x] = iters;
        if (PROF) { out.block_clk[3 * blockIdx.x] = s_clk[1]; out.block_clk[3 * blockIdx.x + 1] = s_clk[2]; out.block_clk[3 * blockIdx.x + 2] = 0; }
    }
}

}  // namespace

// Whether the shared-memory wavefront applies to this scene: the queues plus the staged scene must leave room for at
// least three CTAs per SM (occupancy is what hides the barriers between phases).
bool smem_wavefront_applicable(const pt_context* ctx, const SceneDev& scene, bool brute) {
    (void)ctx;
    size_t bytes = sw_smem_bytes(scene.n_objects, scene.n_nodes, scene.n_lights, brute) + 1024;
    return scene.n_objects > 0 && scene.n_objects <= kMaxSmemPrims && bytes * (kSwMinBlocks < 3 ? kSwMinBlocks : 3) <= 227 * 1024;
}

int smem_wavefront_render(pt_context* ctx, const RenderLaunch& rl, bool brute, pt_stats* stats) {
    cudaStream_t st = ctx->stream;
    const SceneDev& scene = *rl.scene;
    SceneView sv = make_scene_view(scene);
    Frame f = make_frame(rl.cam, rl.prm, rl.d_accum);
    if (f.total_work == 0) { if (stats) stats->samples = 0; return PT_OK; }
    f.blo = make_float3(scene.bounds[0], scene.bounds[1], scene.bounds[2]);
    f.bhi = make_float3(scene.bounds[3], scene.bounds[4], scene.bounds[5]);
    const bool profile = (rl.prm.flags & (PT_FLAG_PROFILE | PT_FLAG_PHASE_CLOCKS)) != 0;   // CUDA events around the launch
    const bool phases = (rl.prm.flags & PT_FLAG_PHASE_CLOCKS) != 0;                           // the kernel variant that reads clock64 per phase
    const size_t smem = sw_smem_bytes(scene.n_objects, scene.n_nodes, scene.n_lights, brute);

    auto pick = [&](auto kernel, int& occ) -> int {
        PT_CUDA_CHECK(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PT_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kSwBlock, smem));
        return PT_OK;
    };
    const int mode = !brute ? 0 : pick_scan_mode(scene, rl.cam, rl.prm.traversal);
    const BruteParam bp = make_brute_param(scene);
    const FastParam fp = make_fast_param(scene);
    const CullParam cull = make_cull_param(scene, rl.cam, mode >= 2);
    const uint32_t n_fast_pairs = scene.n_fast_pairs;
    int occ = 0, rc;
    auto with_kernel = [&](auto&& fn) -> int {
        switch (mode * 2 + (phases ? 1 : 0)) {
            case 0: return fn(k_smem_wavefront<0, false>);
            case 1: return fn(k_smem_wavefront<0, true>);
            case 2: return fn(k_smem_wavefront<1, false>);
            case 3: return fn(k_smem_wavefront<1, true>);
            case 4: return fn(k_smem_wavefront<2, false>);
            case 5: return fn(k_smem_wavefront<2, true>);
            case 6: return fn(k_smem_wavefront<3, false>);
            case 7: return fn(k_smem_wavefront<3, true>);
            case 8: return fn(k_smem_wavefront<4, false>);
            default: return fn(k_smem_wavefront<4, true>);
        }
    };
    rc = with_kernel([&](auto kernel) { return pick(kernel, occ); });
    if (rc) return rc;
    if (occ < 1) { ctx->err = "smem wavefront: kernel does not fit on an SM"; return PT_ERR_CUDA; }

    // one resident wave of CTAs by default (persistent: each CTA runs its share of the work list to completion);
    // path_pool overrides the number of slots, i.e. the number of CTAs
    uint64_t want = rl.prm.path_pool ? ((uint64_t)rl.prm.path_pool + kSwSlots - 1) / kSwSlots : (uint64_t)ctx->sm_count * occ;
    uint64_t need = (f.total_work + kSwSlots - 1) / kSwSlots;
    uint32_t grid = (uint32_t)(want < need ? want : need);
    if (grid == 0) grid = 1;
    f.chunk = 0;                                                     // unused here: work is handed out dynamically (SwOut::work_next)

    size_t out_bytes = (16 + 4 + 24) * (size_t)grid + 1024 + 256;
    if (out_bytes > ctx->pool_bytes) {
        if (ctx->pool_mem) cudaFree(ctx->pool_mem);
        ctx->pool_mem = nullptr; ctx->pool_bytes = 0;
        PT_CUDA_CHECK(ctx, cudaMalloc(&ctx->pool_mem, out_bytes));
        ctx->pool_bytes = out_bytes;
    }
    SwOut out;
    char* p = static_cast<char*>(ctx->pool_mem);
    out.block_stats = reinterpret_cast<uint4*>(p);
    out.block_clk = reinterpret_cast<unsigned long long*>(p + 16 * (size_t)grid);
    out.block_iters = reinterpret_cast<uint32_t*>(p + (16 + 24) * (size_t)grid);
    out.work_next = reinterpret_cast<unsigned long long*>(p + (((16 + 24 + 4) * (size_t)grid + 255) & ~size_t(255)));
    PT_CUDA_CHECK(ctx, cudaMemsetAsync(out.work_next, 0, 8, st));

    if (profile) {
        while (ctx->prof_events.size() < 2) { cudaEvent_t e; PT_CUDA_CHECK(ctx, cudaEventCreate(&e)); ctx->prof_events.push_back(e); }
        PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->prof_events[0], st));
    }
    with_kernel([&](auto kernel) { kernel<<<grid, kSwBlock, smem, st>>>(sv, f, out, bp, fp, n_fast_pairs, scene.n_fast_spheres, cull); return PT_OK; });
    PT_CUDA_CHECK(ctx, cudaGetLastError());
    if (profile) PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->prof_events[1], st));
    if (stats) {
        std::vector<uint4> bs(grid);
        std::vector<uint32_t> it(grid);
        PT_CUDA_CHECK(ctx, cudaMemcpyAsync(bs.data(), out.block_stats, 16 * (size_t)grid, cudaMemcpyDeviceToHost, st));
        PT_CUDA_CHECK(ctx, cudaMemcpyAsync(it.data(), out.block_iters, 4 * (size_t)grid, cudaMemcpyDeviceToHost, st));
        std::vector<unsigned long long> clk(phases ? 3 * (size_t)grid : 0);
        if (phases) PT_CUDA_CHECK(ctx, cudaMemcpyAsync(clk.data(), out.block_clk, 24 * (size_t)grid, cudaMemcpyDeviceToHost, st));
        PT_CUDA_CHECK(ctx, cudaStreamSynchronize(st));
        uint64_t a = 0, b = 0, c = 0, d = 0, itmax = 0;
        for (auto& v : bs) { a += v.x; b += v.y; c += v.z; d += v.w; }
        for (auto v : it) itmax = v > itmax ? v : itmax;
        stats->samples = f.total_work;
        stats->rays_primary = a; stats->rays_shadow = b; stats->rays_scattered = c; stats->path_vertices = d;
        stats->wavefront_iters = itmax; stats->kernel_launches = 1; stats->path_pool = grid * kSwSlots;
        stats->n_kernel[PT_K_SMEM_WAVEFRONT] = 1;
        stats->scan_mode = (uint32_t)mode;
        if (profile) {
            float ms = 0;
            cudaEventElapsedTime(&ms, ctx->prof_events[0], ctx->prof_events[1]);
            stats->ms_kernel[PT_K_SMEM_WAVEFRONT] = ms;
            double t = 0, sh = 0, h = 0;
            for (size_t k = 0; phases && k < grid; ++k) { t += (double)clk[3 * k]; sh += (double)clk[3 * k + 1]; h += (double)clk[3 * k + 2]; }
            double tot = t + sh + h;
            if (tot > 0) { stats->phase_share[0] = (float)(t / tot); stats->phase_share[1] = (float)(sh / tot); stats->phase_share[2] = (float)(h / tot); }
        }
    }
    return PT_OK;
}

// ---------------------------------------------------------------------------------------------------------------------------
// pt_render_device_noalloc: the same kernel with its scratch memory in a file-scope device array instead of cudaMalloc'd
// buffers (no counterpart in the reference; added for callers that cannot allocate:
// code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// Scratch of one call = the float4 accumulation buffer of the frame (what pt_render keeps in pt_context::d_accum) followed by
// the kernel's SwOut block (per-CTA statistics and the work hand-out counter, pt_context::pool_mem in pt_render).  The array is
// part of the module image: every GPU that loads the library holds kNoAllocScratchBytes of it until the process exits.  All
// calls on a device share it and it keeps the last call's contents, so the accumulation buffer and the work counter are
// cleared at the start of every call; two calls that overlap on the GPU (two contexts on one device, or one context used from
// two threads) race on it.
constexpr uint32_t kNoAllocMaxPixels = PT_NOALLOC_MAX_PIXELS;                 // 65536 pixels -> 1 MiB of float4 sums
constexpr uint32_t kNoAllocMaxGrid = 1024;                                    // CTAs (148 x occupancy on a B200; path_pool may ask for more)
constexpr size_t kNoAllocAccumBytes = sizeof(float4) * (size_t)kNoAllocMaxPixels;
constexpr size_t kNoAllocOutBytes = (((16 + 4 + 24) * (size_t)kNoAllocMaxGrid + 255) & ~size_t(255)) + 256;
constexpr size_t kNoAllocScratchBytes = kNoAllocAccumBytes + kNoAllocOutBytes;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

namespace {
constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[kNoAllocMaxDevices] = {};   // device address of g_noalloc_scratch per device ordinal (the name itself is a host-side stand-in)

// Device address of the scratch array on ctx->device, looked up once per device with that device current.
int noalloc_scratch(pt_context* ctx, unsigned char** out) {
    if (ctx->device < 0 || ctx->device >= kNoAllocMaxDevices) { ctx->err = "pt_render_device_noalloc: device ordinal beyond 63"; return PT_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[ctx->device]) {
        PT_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
        void* p = nullptr;
        PT_CUDA_CHECK(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) { ctx->err = "pt_render_device_noalloc: cudaGetSymbolAddress returned a null address"; return PT_ERR_CUDA; }
        g_noalloc_ptr[ctx->device] = p;
    }
    *out = static_cast<unsigned char*>(g_noalloc_ptr[ctx->device]);
    return PT_OK;
}
}  // namespace

// Enqueues clear -> k_smem_wavefront -> k_resolve on the context's stream; no cudaMalloc / cudaFree / event creation on this path.
int smem_wavefront_render_noalloc(pt_context* ctx, const SceneDev& scene, const pt_camera& cam, const pt_render_params& prm,
                                  float* d_radiance, uint8_t* d_rgba) {
    cudaStream_t st = ctx->stream;
    const uint64_t npix = (uint64_t)cam.width * cam.height;
    if (npix > kNoAllocMaxPixels) {
        ctx->err = "pt_render_device_noalloc: frame of " + std::to_string(npix) + " pixels exceeds the static scratch (PT_NOALLOC_MAX_PIXELS = " +
                   std::to_string(kNoAllocMaxPixels) + "); use pt_render";
        return PT_ERR_INVALID;
    }
    if (prm.flags != 0) { ctx->err = "pt_render_device_noalloc: flags must be 0 (profiling, counters and the HBM-queue formulation allocate)"; return PT_ERR_INVALID; }
    // the traversal rule of pt_render (wavefront.cu: pick_mode)
    const bool brute = prm.traversal >= PT_TRAVERSAL_BRUTE || (prm.traversal == PT_TRAVERSAL_AUTO && scene.n_objects <= 16) || scene.n_objects < 2;
    if (!smem_wavefront_applicable(ctx, scene, brute)) {
        ctx->err = "pt_render_device_noalloc: the scene does not fit the shared-memory wavefront (its HBM path queues need cudaMalloc); use pt_render";
        return PT_ERR_INVALID;
    }
    unsigned char* scratch = nullptr;
    int rc = noalloc_scratch(ctx, &scratch);
    if (rc) return rc;
    float4* d_accum = reinterpret_cast<float4*>(scratch);

    SceneView sv = make_scene_view(scene);
    Frame f = make_frame(cam, prm, d_accum);
    if (f.total_work == 0) { ctx->err = "pt_render_device_noalloc: spp must be > 0"; return PT_ERR_INVALID; }
    f.blo = make_float3(scene.bounds[0], scene.bounds[1], scene.bounds[2]);
    f.bhi = make_float3(scene.bounds[3], scene.bounds[4], scene.bounds[5]);
    f.chunk = 0;
    const size_t smem = sw_smem_bytes(scene.n_objects, scene.n_nodes, scene.n_lights, brute);
    const int mode = !brute ? 0 : pick_scan_mode(scene, cam, prm.traversal);
    auto with_kernel = [&](auto&& fn) -> int {
        switch (mode) {
            case 0: return fn(k_smem_wavefront<0, false>);
            case 1: return fn(k_smem_wavefront<1, false>);
            case 2: return fn(k_smem_wavefront<2, false>);
            case 3: return fn(k_smem_wavefront<3, false>);
            default: return fn(k_smem_wavefront<4, false>);
        }
    };
    int occ = 0;
    rc = with_kernel([&](auto kernel) -> int {
        PT_CUDA_CHECK(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PT_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kSwBlock, smem));
        return PT_OK;
    });
    if (rc) return rc;
    if (occ < 1) { ctx->err = "pt_render_device_noalloc: kernel does not fit on an SM"; return PT_ERR_CUDA; }
    const uint64_t want = prm.path_pool ? ((uint64_t)prm.path_pool + kSwSlots - 1) / kSwSlots : (uint64_t)ctx->sm_count * occ;
    const uint64_t need = (f.total_work + kSwSlots - 1) / kSwSlots;
    uint64_t grid = want < need ? want : need;
    if (grid == 0) grid = 1;
    if (grid > kNoAllocMaxGrid) {
        ctx->err = "pt_render_device_noalloc: " + std::to_string(grid) + " CTAs exceed the static scratch (" + std::to_string(kNoAllocMaxGrid) + "); lower path_pool";
        return PT_ERR_INVALID;
    }
    SwOut out;
    unsigned char* p = scratch + kNoAllocAccumBytes;
    out.block_stats = reinterpret_cast<uint4*>(p);
    out.block_clk = reinterpret_cast<unsigned long long*>(p + 16 * (size_t)grid);
    out.block_iters = reinterpret_cast<uint32_t*>(p + (16 + 24) * (size_t)grid);
    out.work_next = reinterpret_cast<unsigned long long*>(p + (((16 + 24 + 4) * (size_t)grid + 255) & ~size_t(255)));
    // the array keeps the previous call's sums and counter: both must start at zero
    PT_CUDA_CHECK(ctx, cudaMemsetAsync(d_accum, 0, sizeof(float4) * (size_t)npix, st));
    PT_CUDA_CHECK(ctx, cudaMemsetAsync(out.work_next, 0, 8, st));
    const FastParam fp = make_fast_param(scene);
    const BruteParam bp = make_brute_param(scene);
    const CullParam cull = make_cull_param(scene, cam, mode >= 2);
    with_kernel([&](auto kernel) { kernel<<<(uint32_t)grid, kSwBlock, smem, st>>>(sv, f, out, bp, fp, scene.n_fast_pairs, scene.n_fast_spheres, cull); return PT_OK; });
    PT_CUDA_CHECK(ctx, cudaGetLastError());
    return resolve_launch(ctx, d_accum, cam.width, cam.height, prm.spp, d_radiance, d_rgba);
}
