// pt_wave.cuh — device code shared by the two wavefront formulations (HBM queues: wavefront.cu; shared-memory queues:
// smem_wavefront.cu): frame / scene views, primary-ray generation, LBVH traversal, and the per-vertex shading step.
// Citations are into /root/reference/src.
#pragma once
#include <cfloat>
#include <cmath>
#include <cstring>

#include "pt_internal.h"
#include "pt_shade.cuh"

namespace ptw {
using namespace ptd;

constexpr int kStackDepth = 64;
constexpr uint32_t kMaxSmemPrims = 1024;   // scene staged in shared memory up to this many objects
constexpr uint32_t kMaxPathDepth = 254;    // the path state keeps the depth in 8 bits (pack_meta): a path that survives roulette this
                                           // long is cut (rendering.rs:91-98 has no hard cap; with the reference's MAX_DEPTH = 50 the
                                           // survival probability at depth 254 is below 2^-250)

struct Frame {
    CameraDev cam;
    uint32_t npix;
    uint64_t total_work;     // npix * spp of this call
    uint64_t chunk;          // work items owned by each shade-CTA (contiguous range)
    uint32_t sample_offset, min_depth, max_depth, strategy;
    uint32_t k0, k1;         // Philox key
    float t_min;
    float4* accum;           // per pixel: sum r, g, b, (reserved)
    float3 blo, bhi;         // padded scene bounds (SceneDev::bounds); blo.x > bhi.x: none
    uint64_t magic_npix, magic_width;   // div_magic constants of npix and cam.width
};

struct SceneView {
    const float4* prims;
    const float4* mats;
    const uint32_t* light_ids;
    const float4* nodes;
    uint32_t n_objects, n_lights, n_nodes;
    const float4* brute;     // brute-force scan table (may be null)
    uint32_t n_brute, n_pairs;
    float box_pad, ro_near;  // one-FMA box test from shared memory (slab_fma): boxes are padded by box_pad when staged, which covers ray origins with |ro| <= ro_near
};

// n / d for any 32-bit n with the host's m = floor(2^64 / d) + 1 (d >= 2): exact, because n * (m * d - 2^64) < 2^64.
PT_DEV uint32_t div_magic(uint32_t n, uint64_t m) { return (uint32_t)__umul64hi((uint64_t)n, m); }
inline uint64_t magic_of(uint32_t d) { return d >= 2 ? ~0ull / d + 1ull : 0ull; }   // (for a power of two this is 2^64 / d itself: exact as well)
// Camera ray of work item (sample_rel, p): p = pixel index plus a small offset, p / npix further samples (the shared-memory
// wavefront hands out contiguous ranges and decodes them without 64-bit division).
PT_DEV void make_primary_at(const Frame& f, uint32_t sample_rel, uint32_t p, float4& ro, float4& rd, uint2& meta, uint32_t& px, uint32_t& py_flipped) {
    const uint32_t q = div_magic(p, f.magic_npix);
    const uint32_t pixel = p - q * f.npix, sample = f.sample_offset + sample_rel + q;
    const uint32_t y = div_magic(pixel, f.magic_width), x = pixel - y * f.cam.width;
    uint4 r = rng_block(pixel, sample, 0xFFFFFFFFu, 0u, f.k0, f.k1);
    px = x; py_flipped = f.cam.height - 1 - y;
    float3 d = camera_dir(f.cam, x, py_flipped, u01(r.x), u01(r.y));             // world.rs:299
    ro = make_float4(f.cam.origin.x, f.cam.origin.y, f.cam.origin.z, 1.0f);      // eta_ratio = 1, camera.rs:14
    rd = make_float4(d.x, d.y, d.z, 0.0f);
    meta = make_uint2(pixel | 0x80000000u, sample & 0x00ffffffu);
}
PT_DEV void make_primary(const Frame& f, uint64_t w, float4& ro, float4& rd, uint2& meta, uint32_t& px, uint32_t& py_flipped) {
    uint32_t sample = f.sample_offset + (uint32_t)(w / f.npix);
    uint32_t pixel = (uint32_t)(w % f.npix);
    uint32_t x = pixel % f.cam.width, y = pixel / f.cam.width;
    uint4 r = rng_block(pixel, sample, 0xFFFFFFFFu, 0u, f.k0, f.k1);
    px = x; py_flipped = f.cam.height - 1 - y;
    float3 d = camera_dir(f.cam, x, py_flipped, u01(r.x), u01(r.y));             // world.rs:299
    ro = make_float4(f.cam.origin.x, f.cam.origin.y, f.cam.origin.z, 1.0f);      // eta_ratio = 1, camera.rs:14
    rd = make_float4(d.x, d.y, d.z, 0.0f);
    meta = make_uint2(pixel | 0x80000000u, sample & 0x00ffffffu);
}
PT_DEV void make_primary(const Frame& f, uint64_t w, float4& ro, float4& rd, uint2& meta) {
    uint32_t px, py;
    make_primary(f, w, ro, rd, meta, px, py);
}

// ------------------------------------------------------------------------------------ traversal
// Scene staged in shared memory: [nodes: 4 float4 each][prims: 3 float4 each (q0,q1,q2)]
template <bool SMEM, int PS = 3>
struct Geo {
    const float4* nodes;
    const float4* brute = nullptr;   // smem copy of the brute-force scan table, n_brute entries of 5 float4 (null: scan prims):
    uint32_t n_brute = 0, n_pairs = 0;   // n_pairs triangle pairs first, then the unpaired primitives
    const float4* prims;   // stride PS (smem: 3 = q0,q1,q2 only; 4 = with the normal) or 4 (global)
    float ro_near = 0.f;   // see SceneView
    PT_DEV void node(int i, float4& a, float4& b, float4& c, float4& d) const {
        if (SMEM) { a = nodes[4 * i]; b = nodes[4 * i + 1]; c = nodes[4 * i + 2]; d = nodes[4 * i + 3]; }
        else { a = __ldg(nodes + 4 * i); b = __ldg(nodes + 4 * i + 1); c = __ldg(nodes + 4 * i + 2); d = __ldg(nodes + 4 * i + 3); }
    }
    PT_DEV void prim(int i, float4& q0, float4& q1, float4& q2) const {
        if (SMEM) { q0 = prims[PS * i]; q1 = prims[PS * i + 1]; q2 = prims[PS * i + 2]; }
        else { q0 = __ldg(prims + 4 * (size_t)i); q1 = __ldg(prims + 4 * (size_t)i + 1); q2 = __ldg(prims + 4 * (size_t)i + 2); }
    }
};
template <bool SMEM>
PT_DEV Geo<SMEM> stage_scene(const SceneView& s, float4* smem, bool want_nodes) {
    Geo<SMEM> g;
    if (SMEM) {
        if (!want_nodes && s.brute) {        // brute force: only the scan table is needed
            for (uint32_t i = threadIdx.x; i < s.n_brute * 5; i += blockDim.x) smem[i] = __ldg(s.brute + i);
            __syncthreads();
            g.nodes = smem; g.prims = smem; g.brute = smem; g.n_brute = s.n_brute; g.n_pairs = s.n_pairs;
            return g;
        }
        uint32_t nn = want_nodes ? s.n_nodes * 4 : 0;
        for (uint32_t i = threadIdx.x; i < nn; i += blockDim.x) {       // [Llo,left][Lhi,right][Rlo,0][Rhi,0]: boxes padded outward by box_pad
            float4 v = __ldg(s.nodes + i);
            const float p = (i & 1u) ? s.box_pad : -s.box_pad;
            v.x += p; v.y += p; v.z += p;
            smem[i] = v;
        }
        for (uint32_t i = threadIdx.x; i < s.n_objects * 3; i += blockDim.x) smem[nn + i] = __ldg(s.prims + 4 * (i / 3) + (i % 3));
        __syncthreads();
        g.nodes = smem; g.prims = smem + nn; g.ro_near = s.ro_near;
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

// The same test with one FMA per plane: t = lo * inv + c, c = -(ro * inv) rounded once per ray.  Against the form above the
// plane distances of axis k carry an extra absolute error of u |c_k| (u = 2^-24: the rounding of c), i.e. the planes of that
// axis are displaced by at most u |ro_k| in space.  The boxes staged in shared memory are padded by box_pad = 8 u max|scene
// coordinate| (stage_scene), which covers every origin with |ro_k| <= ro_near = 4 max|scene coordinate| whatever the direction;
// a ray that starts farther away (a distant camera) gets slack = 3 u max|c| added to the comparison instead.  Either way the
// test enters a superset of the boxes the exact form enters, so ids and t of the walk are unchanged.  (A slack for every ray
// was the first form: a near-axis-parallel ray has |c_k| ~ 1e6, lost all culling and walked the whole tree — a handful of
// such rays per launch set its duration on large trees.)
PT_DEV bool slab_fma(float3 lo, float3 hi, float3 inv, float3 c, float t_min, float t_max, float slack, float& tnear) {
    float tx0 = fmaf(lo.x, inv.x, c.x), tx1 = fmaf(hi.x, inv.x, c.x);
    float ty0 = fmaf(lo.y, inv.y, c.y), ty1 = fmaf(hi.y, inv.y, c.y);
    float tz0 = fmaf(lo.z, inv.z, c.z), tz1 = fmaf(hi.z, inv.z, c.z);
    float tn = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), fmaxf(fminf(tz0, tz1), t_min));
    float tf = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), fminf(fmaxf(tz0, tz1), t_max));
    tnear = tn;
    return tn <= fmaf(tf, 1.0000004f, slack);
}
PT_DEV void slab_fma_setup(float3 ro, float3 inv, float ro_near, float3& c, float& slack) {
    c = v3(-(ro.x * inv.x), -(ro.y * inv.y), -(ro.z * inv.z));
    const bool near = fmaxf(fmaxf(fabsf(ro.x), fabsf(ro.y)), fabsf(ro.z)) <= ro_near;
    slack = near ? 0.0f : 1.8e-7f * fmaxf(fmaxf(fabsf(c.x), fabsf(c.y)), fabsf(c.z));    // 3 u max|c| = 1.79e-7 max|c|
}

// ---- brute-force scan table (lbvh.cu: k_brute_table): triangle pairs in ascending object order, then the unpaired primitives.
// Both tests of a pair see the t_max in force before the pair and the updates re-check the inclusive bound in index order, so
// the pairs alone reproduce the sequential scan of world.rs:281-287 (ties go to the later object).  The unpaired primitives
// come after every pair, out of index order, and therefore break ties on the object index explicitly.
template <bool CHECKED = true>
PT_DEV void pair_closest(float4 A, float4 B, float4 C, float4 D, float4 E, float3 ro, float3 rd, float t_min, float& tmax, int& best_id) {
    bool oka, okb;
    float2 t = hit_triangle2<CHECKED, false>(pk(A.x, A.y), pk(A.z, A.w), pk(B.x, B.y), pk(B.z, B.w), pk(C.x, C.y), pk(C.z, C.w), pk(D.x, D.y),
                                             pk(D.z, D.w), pk(E.x, E.y), ro, rd, t_min, tmax, oka, okb);
    if (oka) { tmax = t.x; best_id = __float_as_int(E.z); }
    if (okb & (t.y <= tmax)) { tmax = t.y; best_id = __float_as_int(E.w); }   // also lane b's upper test: tmax <= the bound before the pair
}
PT_DEV void single_closest(float4 q0, float4 q1, float4 q2, float4 E, float3 ro, float3 rd, float t_min, float& tmax, int& best_id) {
    const int id = __float_as_int(E.z);
    float t = hit_prim(q0, q1, q2, ro, rd, t_min, tmax);
    if (t >= 0.0f && (t < tmax || id > best_id)) { tmax = t; best_id = id; }
}
template <bool CHECKED = true>
PT_DEV bool pair_any(float4 A, float4 B, float4 C, float4 D, float4 E, float3 ro, float3 rd, float t_min, float t_max) {
    bool oka, okb;
    hit_triangle2<CHECKED, true>(pk(A.x, A.y), pk(A.z, A.w), pk(B.x, B.y), pk(B.z, B.w), pk(C.x, C.y), pk(C.z, C.w), pk(D.x, D.y), pk(D.z, D.w),
                                 pk(E.x, E.y), ro, rd, t_min, t_max, oka, okb);
    return oka | okb;
}
// Up to kParamPairs pairs of the table can also travel as a kernel parameter: the scan then takes its triangle operands
// from the constant bank through uniform registers (no shared-memory loads, no vector registers) — the small-scene form the
// shared-memory wavefront uses (only when ALL pairs fit).  The pairs are RIGHT-aligned in the parameter (pair j of n sits in
// slot kParamPairs - n + j), so the unrolled scan is entered once, at the slot of the first pair, and runs to the end in
// ascending object order without a per-pair exit test.
constexpr int kParamPairs = kBruteParamPairs;
static_assert(kParamPairs == 8, "the unrolled scans below are written for 8 slots");
struct BruteParam { float4 e[5 * kParamPairs]; };
#define PT_PARAM_SCAN(n_pairs, STEP)                                                                                          \
    switch (n_pairs) {                                                                                                        \
        case 8: STEP(0) [[fallthrough]];                                                                                      \
        case 7: STEP(1) [[fallthrough]];                                                                                      \
        case 6: STEP(2) [[fallthrough]];                                                                                      \
        case 5: STEP(3) [[fallthrough]];                                                                                      \
        case 4: STEP(4) [[fallthrough]];                                                                                      \
        case 3: STEP(5) [[fallthrough]];                                                                                      \
        case 2: STEP(6) [[fallthrough]];                                                                                      \
        case 1: STEP(7) [[fallthrough]];                                                                                      \
        default: break;                                                                                                       \
    }
template <bool CHECKED, class G>
PT_DEV void closest_hit_param(const BruteParam& bp, const G& g, float3 ro, float3 rd, float t_min, float& best_t, int& best_id) {
    best_t = FLT_MAX; best_id = -1;
    float tmax = __int_as_float(0x7f800000);
#define PT_STEP(k) pair_closest<CHECKED>(bp.e[5 * (k)], bp.e[5 * (k) + 1], bp.e[5 * (k) + 2], bp.e[5 * (k) + 3], bp.e[5 * (k) + 4], ro, rd, t_min, tmax, best_id);
    PT_PARAM_SCAN(g.n_pairs, PT_STEP)
#undef PT_STEP
    for (uint32_t e = g.n_pairs; e < g.n_brute; ++e)
        single_closest(g.brute[5 * e], g.brute[5 * e + 1], g.brute[5 * e + 2], g.brute[5 * e + 4], ro, rd, t_min, tmax, best_id);
    if (best_id >= 0) best_t = tmax;
}
// Screen-space cull of the primary-ray scan.  Every camera ray of pixel (x, y_flipped) passes through the image-plane cell
// [x, x+1] x [y_flipped, y_flipped+1] (camera.rs:139-147 with the jitter in (0, 1)), so a pair of triangles (or a sphere) whose
// projection lies outside that cell cannot be hit: the host projects each entry in f64 (make_cull_param below) and stores a
// padded pixel rectangle; a warp skips an entry when none of its lanes' pixels is inside.  Lanes that run the test anyway get
// the exact answer, so ids and t are those of the full scan bit for bit (tested through pt_trace_primary).  Entries the host
// cannot bound (a vertex behind the camera, a plane seen edge-on, the camera inside a sphere) keep the full-screen rectangle.
constexpr int kCullSingles = 8;
struct CullParam {
    int4 pair[kParamPairs];        // (x0, y0, x1 - x0, y1 - y0) in the slot order of BruteParam; empty: x0 = INT_MAX
    int4 single[kCullSingles];     // the first unpaired entries of the scan table, in table order
    uint32_t enabled;              // host-side note; 0: every rectangle is the full screen (e.g. the caller's ray offsets are not in [0, 1])
};
PT_DEV bool cull_inside(int4 c, uint32_t x, uint32_t y) { return ((x - (uint32_t)c.x) <= (uint32_t)c.z) & ((y - (uint32_t)c.y) <= (uint32_t)c.w); }
template <bool CHECKED, class G>
PT_DEV void closest_hit_param_culled(const BruteParam& bp, const CullParam& cull, const G& g, uint32_t x, uint32_t y, float3 ro, float3 rd, float t_min,
                                     float& best_t, int& best_id) {
    best_t = FLT_MAX; best_id = -1;
    float tmax = __int_as_float(0x7f800000);
#define PT_STEP(k)                                                                                                            \
    if (__any_sync(__activemask(), cull_inside(cull.pair[k], x, y)))                                                          \
        pair_closest<CHECKED>(bp.e[5 * (k)], bp.e[5 * (k) + 1], bp.e[5 * (k) + 2], bp.e[5 * (k) + 3], bp.e[5 * (k) + 4], ro, rd, t_min, tmax, best_id);
    PT_PARAM_SCAN(g.n_pairs, PT_STEP)
#undef PT_STEP
    for (uint32_t e = g.n_pairs; e < g.n_brute; ++e) {
        const uint32_t k = e - g.n_pairs;
        if (k < (uint32_t)kCullSingles && !__any_sync(__activemask(), cull_inside(cull.single[k], x, y))) continue;
        single_closest(g.brute[5 * e], g.brute[5 * e + 1], g.brute[5 * e + 2], g.brute[5 * e + 4], ro, rd, t_min, tmax, best_id);
    }
    if (best_id >= 0) best_t = tmax;
}
template <bool CHECKED, class G>
PT_DEV bool any_hit_param(const BruteParam& bp, const G& g, float3 ro, float3 rd, float t_min, float t_max) {
    bool occluded = false;
#define PT_STEP(k) occluded |= pair_any<CHECKED>(bp.e[5 * (k)], bp.e[5 * (k) + 1], bp.e[5 * (k) + 2], bp.e[5 * (k) + 3], bp.e[5 * (k) + 4], ro, rd, t_min, t_max);
    PT_PARAM_SCAN(g.n_pairs, PT_STEP)
#undef PT_STEP
    for (uint32_t e = g.n_pairs; e < g.n_brute; ++e)
        occluded |= hit_prim(g.brute[5 * e], g.brute[5 * e + 1], g.brute[5 * e + 2], ro, rd, t_min, t_max) >= 0.0f;
    return occluded;
}

// ---- secondary-ray scan.  The north star pins PRIMARY hits bit for bit (ids) and to 1e-5 (t); shadow rays need a boolean and
// scattered rays a hit that is right to rounding.  For them every triangle is a planar entry (SceneDev::fast_host) tested with
// its precomputed world -> (u, v, w) affine map: 18 packed FMAs give origin and direction in the entry's frame for TWO entries
// at once, t = -o_w / d_w, (u, v) = o_uv + t d_uv, and a parallelogram made of two triangles is ONE entry — a third of the
// arithmetic of two Moller-Trumbore tests per pair of triangles.  Tested against the f64 oracle through pt_trace_primary
// (PT_TRAVERSAL_BRUTE_FAST): ids equal except on edges, t within 1e-5.
constexpr int kFastPairs = kFastParamPairs;
static_assert(kFastPairs == 4, "PT_FAST_SCAN below is written for 4 slots");
constexpr int kFastSpheres = kFastParamSpheres;
static_assert(kFastSpheres == 4, "PT_FAST_SPHERES below is written for 4 slots");
struct FastParam { float4 e[8 * kFastPairs]; float4 sph[kFastSpheres]; int sph_id[kFastSpheres]; };   // spheres right-aligned like the pairs
#define PT_FAST_SPHERES(n, STEP)                                                                                              \
    switch (n) {                                                                                                              \
        case 4: STEP(0) [[fallthrough]];                                                                                      \
        case 3: STEP(1) [[fallthrough]];                                                                                      \
        case 2: STEP(2) [[fallthrough]];                                                                                      \
        case 1: STEP(3) [[fallthrough]];                                                                                      \
        default: break;                                                                                                       \
    }
#define PT_FAST_SCAN(n_pairs, STEP)                                                                                           \
    switch (n_pairs) {                                                                                                        \
        case 4: STEP(0) [[fallthrough]];                                                                                      \
        case 3: STEP(1) [[fallthrough]];                                                                                      \
        case 2: STEP(2) [[fallthrough]];                                                                                      \
        case 1: STEP(3) [[fallthrough]];                                                                                      \
        default: break;                                                                                                       \
    }
// t of both entries and whether each is hit in [t_min, t_max]; uv = u + v tells the two triangles of a parallelogram apart
template <bool SKIP>
PT_DEV float2 fast_pair(const float4* E, float3 ro, float3 rd, float t_min, float t_max, bool& ok_a, bool& ok_b, float2& uv) {
    const float2 ox = pk(ro.x), oy = pk(ro.y), oz = pk(ro.z), dx = pk(rd.x), dy = pk(rd.y), dz = pk(rd.z);
    auto row_o = [&](float4 A, float4 B) { return __ffma2_rn(oz, pk(B.x, B.y), __ffma2_rn(oy, pk(A.z, A.w), __ffma2_rn(ox, pk(A.x, A.y), pk(B.z, B.w)))); };
    auto row_d = [&](float4 A, float4 B) { return __ffma2_rn(dz, pk(B.x, B.y), __ffma2_rn(dy, pk(A.z, A.w), __fmul2_rn(dx, pk(A.x, A.y)))); };
    const float2 ow = row_o(E[4], E[5]), dw = row_d(E[4], E[5]);
    float ia, ib;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ia) : "f"(dw.x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ib) : "f"(dw.y));
    const float2 t = __fmul2_rn(pneg(ow), pk(ia, ib));
    // The plane is met inside [t_min, t_max] for few (entry, ray) combinations — a shadow ray inside a convex room meets none:
    // when no lane of the warp needs the (u, v) rows of this pair they are skipped (12 of the 21 packed operations).
    const bool in_a = (t.x >= t_min) & (t.x <= t_max), in_b = (t.y >= t_min) & (t.y <= t_max);
    if (SKIP && !__any_sync(__activemask(), in_a | in_b)) { ok_a = ok_b = false; uv = pk(0.f); return t; }
    const float2 u = __ffma2_rn(t, row_d(E[0], E[1]), row_o(E[0], E[1]));
    const float2 v = __ffma2_rn(t, row_d(E[2], E[3]), row_o(E[2], E[3]));
    uv = __fadd2_rn(u, v);
    // a ray parallel to the plane gives t = +-inf or NaN: refused by the range test (t_max <= FLT_MAX)
    // (u, v) are centred on the entry: the unit square is |u|, |v| <= 1/2 and `u + v <= 1` of a triangle is uv <= 0 (E[7] = lim - 1)
    ok_a = (fmaxf(fabsf(u.x), fabsf(v.x)) <= 0.5f) & (uv.x <= E[7].x) & in_a;
    ok_b = (fmaxf(fabsf(u.y), fabsf(v.y)) <= 0.5f) & (uv.y <= E[7].y) & in_b;
    return t;
}
// shape.rs:53-89 without the correctly rounded divisions and square root of hit_sphere (same cancellation-free quadratic)
PT_DEV float hit_sphere_fast(float3 c, float radius, float3 ro, float3 rd, float t_min, float t_max) {
    const float3 oc = ro - c;
    const float a = dot(rd, rd), bp = -dot(oc, rd);
    const float ra = __frcp_rn(a) , k = bp * ra;
    const float3 l = oc + rd * k;
    const float disc = radius * radius - dot(l, l);
    const float cc = dot(oc, oc) - radius * radius;
    const float q = bp + copysignf(sqrtf(fmaxf(a * disc, 0.0f)), bp);
    const float r0 = __fdividef(cc, q), r1 = q * ra;
    const float t_near = bp >= 0.0f ? r0 : r1, t_far = bp >= 0.0f ? r1 : r0;
    const bool near_ok = !(t_near < t_min || t_max < t_near), far_ok = !(t_far < t_min || t_max < t_far);
    return (disc < 0.0f || !(near_ok || far_ok)) ? -1.0f : (near_ok ? t_near : t_far);
}
template <class G>
PT_DEV void closest_hit_fast(const FastParam& fp, uint32_t n_fast_pairs, uint32_t n_fast_spheres, const G& g, float3 ro, float3 rd, float t_min, float& best_t, int& best_id) {
    best_t = FLT_MAX; best_id = -1;
    float tmax = FLT_MAX;
#define PT_STEP(k)                                                                                                            \
    {                                                                                                                         \
        bool oka, okb; float2 uv;                                                                                             \
        const float2 t = fast_pair<false>(fp.e + 8 * (k), ro, rd, t_min, tmax, oka, okb, uv);                                        \
        const float4 ids = fp.e[8 * (k) + 6];                                                                                 \
        if (oka) { tmax = t.x; best_id = __float_as_int(uv.x <= 0.0f ? ids.x : ids.z); }                                      \
        if (okb & (t.y <= tmax)) { tmax = t.y; best_id = __float_as_int(uv.y <= 0.0f ? ids.y : ids.w); }                      \
    }
    PT_FAST_SCAN(n_fast_pairs, PT_STEP)
#undef PT_STEP
    (void)g;
#define PT_STEP(k)                                                                                                            \
    {                                                                                                                         \
        const float t = hit_sphere_fast(v3(fp.sph[k]), fp.sph[k].w, ro, rd, t_min, tmax);                                     \
        if (t >= 0.0f) { tmax = t; best_id = fp.sph_id[k]; }                                                                  \
    }
    PT_FAST_SPHERES(n_fast_spheres, PT_STEP)
#undef PT_STEP
    if (best_id >= 0) best_t = tmax;
}
template <class G>
PT_DEV bool any_hit_fast(const FastParam& fp, uint32_t n_fast_pairs, uint32_t n_fast_spheres, const G& g, float3 ro, float3 rd, float t_min, float t_max) {
    bool occluded = false;
#define PT_STEP(k)                                                                                                            \
    {                                                                                                                         \
        bool oka, okb; float2 uv;                                                                                             \
        fast_pair<true>(fp.e + 8 * (k), ro, rd, t_min, t_max, oka, okb, uv);                                                        \
        occluded |= oka | okb;                                                                                                \
    }
    PT_FAST_SCAN(n_fast_pairs, PT_STEP)
#undef PT_STEP
    (void)g;
#define PT_STEP(k) occluded |= hit_sphere_fast(v3(fp.sph[k]), fp.sph[k].w, ro, rd, t_min, t_max) >= 0.0f;
    PT_FAST_SPHERES(n_fast_spheres, PT_STEP)
#undef PT_STEP
    return occluded;
}

// Closest hit.  Winner = smallest t in [t_min, inf); equal t -> higher object index (world.rs:281-287 scans
// in index order with an inclusive t_max, so a later object replaces an earlier one on a tie).
template <bool SMEM, bool BRUTE, class G>
PT_DEV void closest_hit(const G& g, uint32_t n_objects, float3 ro, float3 rd, float t_min, float& best_t, int& best_id) {
    best_t = FLT_MAX; best_id = -1;
    float tmax = __int_as_float(0x7f800000);
    auto test = [&](int id) {
        float4 q0, q1, q2;
        g.prim(id, q0, q1, q2);
        float t = hit_prim(q0, q1, q2, ro, rd, t_min, tmax);
        if (t >= 0.0f && (t < tmax || id > best_id)) { tmax = t; best_id = id; }
    };
    if (BRUTE && SMEM && g.brute) {
        for (uint32_t e = 0; e < g.n_pairs; ++e)
            pair_closest(g.brute[5 * e], g.brute[5 * e + 1], g.brute[5 * e + 2], g.brute[5 * e + 3], g.brute[5 * e + 4], ro, rd, t_min, tmax, best_id);
        for (uint32_t e = g.n_pairs; e < g.n_brute; ++e)
            single_closest(g.brute[5 * e], g.brute[5 * e + 1], g.brute[5 * e + 2], g.brute[5 * e + 4], ro, rd, t_min, tmax, best_id);
    } else if (BRUTE) {
        for (int id = 0; id < (int)n_objects; ++id) test(id);
    } else {
        float3 inv = v3(__frcp_rn(rd.x), __frcp_rn(rd.y), __frcp_rn(rd.z));   // exact: the slab margins assume IEEE reciprocals
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
                PT_CHECK(sp < kStackDepth);
                stack[sp++] = cr; node = cl;
            } else if (hl) node = cl;
            else if (hr) node = cr;
            else { if (sp == 0) break; node = stack[--sp]; }
        }
    }
    if (best_id >= 0) best_t = tmax;
}
// Any hit in [t_min, t_max] (rendering.rs:62-65)
template <bool SMEM, bool BRUTE, class G>
PT_DEV bool any_hit(const G& g, uint32_t n_objects, float3 ro, float3 rd, float t_min, float t_max) {
    if (BRUTE && SMEM && g.brute) {
        bool occluded = false;
        for (uint32_t e = 0; e < g.n_pairs; ++e)
            occluded |= pair_any(g.brute[5 * e], g.brute[5 * e + 1], g.brute[5 * e + 2], g.brute[5 * e + 3], g.brute[5 * e + 4], ro, rd, t_min, t_max);
        for (uint32_t e = g.n_pairs; e < g.n_brute; ++e)
            occluded |= hit_prim(g.brute[5 * e], g.brute[5 * e + 1], g.brute[5 * e + 2], ro, rd, t_min, t_max) >= 0.0f;
        return occluded;
    }
    if (BRUTE) {
        bool occluded = false;
        for (int id = 0; id < (int)n_objects; ++id) {
            float4 q0, q1, q2;
            g.prim(id, q0, q1, q2);
            occluded |= hit_prim(q0, q1, q2, ro, rd, t_min, t_max) >= 0.0f;
        }
        return occluded;
    }
    float3 inv = v3(__frcp_rn(rd.x), __frcp_rn(rd.y), __frcp_rn(rd.z));   // exact: the slab margins assume IEEE reciprocals
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
        if (hl && hr) { PT_CHECK(sp < kStackDepth); stack[sp++] = cr; node = cl; }
        else if (hl) node = cl;
        else if (hr) node = cr;
        else { if (sp == 0) return false; node = stack[--sp]; }
    }
}

// ---- warp-persistent LBVH traversal with dynamic ray fetch (scenes whose BVH is walked through L1/L2: C3, C4).
// One thread per ray with a fixed ray per thread leaves most lanes idle on incoherent rays: the warp runs until its longest
// walk ends and the primitive tests sit inside the node loop (measured on the 1 M-triangle heightfield: 4 of 32 lanes active
// per instruction).  Here a warp streams chunks of the ray queue; lanes descend inner nodes until each holds a leaf, then all
// leaves are tested together (the "while-while" loop of Aila & Laine, HPG 2009), and whenever PT_BVH_REFILL lanes have
// finished their ray the idle lanes fetch the next rays of the queue.  Same tests, same tie rule (equal t -> higher object
// index) as closest_hit<>: the result does not depend on the visiting order.
//   next(want, idle_mask, ro, rd, t_max, slot, exhausted) -> this lane got a ray   (warp-collective, called by all lanes)
//   finish(slot, t, id)                                     per lane, when its walk ends (ANY: id >= 0 means occluded)
#ifndef PT_BVH_MIN_INNER
#define PT_BVH_MIN_INNER 8
#endif
#ifndef PT_BVH_MIN_INNER_SMEM
#define PT_BVH_MIN_INNER_SMEM 8
#endif
#ifndef PT_BVH_REFILL
#define PT_BVH_REFILL 16   // measured on C3 / C4: 1 -> 880 / 249 Msamples/s, 8 -> 902 / 275, 16 -> 923 / 283, 32 (no refill) -> 899 / 258
#endif
#ifndef PT_BVH_REFILL_L2
#define PT_BVH_REFILL_L2 16   // the L1/L2 walk; C4 with static ranges: 8 -> 371, 12 -> 383, 16 -> 389, 24 -> 397; with chunks: 16 -> 468, 24 -> 461
#endif
constexpr int kWalkDone = -2147483647 - 1;
// COUNT: node fetches and primitive tests are counted per lane and added to counters[0] / counters[1] when the warp retires
// (PT_FLAG_COUNTERS: the numerator of the HBM roofline).
template <bool SMEM, bool ANY, bool COUNT, class G, class Next, class Finish>
PT_DEV void bvh_stream(const G& g, float t_min, Next next, Finish finish, unsigned long long* __restrict__ counters = nullptr) {
    uint32_t n_node_fetch = 0, n_prim_test = 0;
    int node = kWalkDone, sp = 0, best_id = -1;
    int stack[kStackDepth];
    float3 ro = v3(0.f, 0.f, 0.f), rd = ro, inv = ro, cneg = ro;
    float tmax = 0.f, slack = 0.f;
    uint32_t slot = 0;
    bool exhausted = false;
    constexpr int kMinInner = SMEM ? PT_BVH_MIN_INNER_SMEM : PT_BVH_MIN_INNER;
    while (true) {
        const uint32_t idle = __ballot_sync(0xffffffffu, node == kWalkDone);
        if (exhausted ? idle == 0xffffffffu : __popc(idle) >= (SMEM ? PT_BVH_REFILL : PT_BVH_REFILL_L2)) {
            if (exhausted) break;
            float3 o, d; float tm; uint32_t sl;
            if (next(node == kWalkDone, idle, o, d, tm, sl, exhausted)) {
                ro = o; rd = d; tmax = tm; slot = sl; best_id = -1; sp = 0; node = 0;
                inv = v3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));   // exact: the slab margins assume IEEE reciprocals
                if (SMEM) slab_fma_setup(o, inv, g.ro_near, cneg, slack);
            }
        }
        // inner nodes: descend until this lane holds a leaf (negative: ~object id) or its stack runs empty.  The warp leaves
        // the loop as soon as fewer than kMinInner lanes are still descending while others hold a leaf: those test their
        // leaves and rejoin the descent instead of idling until the longest descent ends; the lanes cut short keep their
        // node and resume in the next round.  Measured with static queue ranges: C4 (nodes through L1/L2) 336 -> 380
        // Msamples/s, C3 (nodes in shared memory, where the vote per step is a tenth of the step) 1005 -> 967; with the
        // queue handed out in chunks (wavefront.cu: kStreamChunk*) both gain: C3 1279 -> 1320 (4: 1272, 12: 1326, 16: 1302).
        while (true) {
            if (kMinInner > 0) {
                const uint32_t inner = __ballot_sync(0xffffffffu, node >= 0);
                if (inner == 0u) break;
                if (__popc(inner) < kMinInner && __any_sync(0xffffffffu, node < 0 && node > kWalkDone + 1)) break;
                if (node < 0) continue;
            } else if (node < 0) break;
            float4 a, b, c, d;
            g.node(node, a, b, c, d);
            if (COUNT) ++n_node_fetch;
            const int cl = __float_as_int(a.w), cr = __float_as_int(b.w);
            float tl, tr;
            // nodes in shared memory: the walk is instruction-bound, so the one-FMA-per-plane form (12 instructions fewer per node);
            // nodes through L1/L2: latency-bound at 48 registers, where the extra per-ray values would spill
            const bool hl = SMEM ? slab_fma(v3(a), v3(b), inv, cneg, t_min, tmax, slack, tl) : slab(v3(a), v3(b), ro, inv, t_min, tmax, tl);
            const bool hr = SMEM ? slab_fma(v3(c), v3(d), inv, cneg, t_min, tmax, slack, tr) : slab(v3(c), v3(d), ro, inv, t_min, tmax, tr);
            if (hl && hr) {
                const bool swap = tr < tl;
                PT_CHECK(sp < kStackDepth);
                stack[sp++] = swap ? cl : cr;
                node = swap ? cr : cl;
            } else if (hl) node = cl;
            else if (hr) node = cr;
            else node = sp ? stack[--sp] : kWalkDone + 1;             // kWalkDone + 1: walk over, result not yet written
        }
        // leaves, all lanes together
        if (node < 0 && node > kWalkDone + 1) {
            const int id = ~node;
            float4 q0, q1, q2;
            g.prim(id, q0, q1, q2);
            if (COUNT) ++n_prim_test;
            const float t = hit_prim(q0, q1, q2, ro, rd, t_min, tmax);
            if (ANY) {
                if (t >= 0.0f) { best_id = id; sp = 0; }              // occluded: stop this walk
            } else if (t >= 0.0f && (t < tmax || id > best_id)) { tmax = t; best_id = id; }
            node = (ANY && best_id >= 0) ? kWalkDone + 1 : sp ? stack[--sp] : kWalkDone + 1;
        }
        if (node == kWalkDone + 1) {
            finish(slot, best_id >= 0 ? tmax : FLT_MAX, best_id);
            node = kWalkDone;
        }
    }
    if (COUNT) {
        n_node_fetch = __reduce_add_sync(0xffffffffu, n_node_fetch);
        n_prim_test = __reduce_add_sync(0xffffffffu, n_prim_test);
        if ((threadIdx.x & 31) == 0) { atomicAdd(counters, (unsigned long long)n_node_fetch); atomicAdd(counters + 1, (unsigned long long)n_prim_test); }
    }
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


// ------------------------------------------------------------------------------------ scene access for shading
// Global-memory view (HBM-queue wavefront): 4 float4 per primitive, 2 float4 per material, through the read-only path.
struct SceneGlobal {
    const float4* prims; const float4* mats; const uint32_t* light_ids; uint32_t n_lights;
    PT_DEV void prim(int i, float4& q0, float4& q1, float4& q2, float4& q3) const {
        q0 = __ldg(prims + 4 * (size_t)i); q1 = __ldg(prims + 4 * (size_t)i + 1); q2 = __ldg(prims + 4 * (size_t)i + 2); q3 = __ldg(prims + 4 * (size_t)i + 3);
    }
    PT_DEV Mat mat(int i) const { return load_mat(mats, i); }
    PT_DEV float3 emission(int i) const { return v3(__ldg(mats + 2 * (size_t)i)); }
    PT_DEV uint32_t light(uint32_t k) const { return __ldg(light_ids + k); }
};
// Shared-memory view (smem wavefront): same records staged once per CTA.
struct SceneShared {
    const float4* prims; const float4* mats; const uint32_t* light_ids; uint32_t n_lights;
    PT_DEV void prim(int i, float4& q0, float4& q1, float4& q2, float4& q3) const {
        q0 = prims[4 * i]; q1 = prims[4 * i + 1]; q2 = prims[4 * i + 2]; q3 = prims[4 * i + 3];
    }
    PT_DEV Mat mat(int i) const {
        float4 m0 = mats[2 * i], m1 = mats[2 * i + 1];
        Mat m; m.color = v3(m0); m.kind = __float_as_uint(m0.w); m.roughness = m1.x; m.metallic = m1.y; m.ior = m1.z;
        return m;
    }
    PT_DEV float3 emission(int i) const { return v3(mats[2 * i]); }
    PT_DEV uint32_t light(uint32_t k) const { return light_ids[k]; }
};

// ------------------------------------------------------------------------------------ one path vertex
// rendering.rs:34-142 in the iterative form of SURVEY A.6, for ONE live slot whose closest hit is known.
// Outcome: either the path ends (finished; add_L says whether Ladd goes to the pixel), or it continues with a new
// ray / throughput / meta and possibly a shadow ray (emit_shadow) whose contribution sh_c is added when unoccluded.
struct VertexOut {
    bool finished, add_L, emit_shadow, vertex;
    float3 Ladd;
    float3 sh_d, sh_c; float sh_tmax;          // shadow ray from p (== new_ro.xyz) toward the light sample
    float4 new_ro, new_rd; float3 new_beta; uint2 new_meta;
};
template <class Scene>
PT_DEV VertexOut shade_vertex(const Scene& sc, const Frame& f, bool mis, float4 ro4, float4 rd4, float3 beta, uint2 h, uint2 meta) {
    VertexOut o;
    o.finished = false; o.add_L = false; o.emit_shadow = false; o.vertex = false;
    o.Ladd = v3(0.f, 0.f, 0.f); o.sh_d = v3(0.f, 0.f, 0.f); o.sh_c = v3(0.f, 0.f, 0.f); o.sh_tmax = 0.f;
    o.new_ro = ro4; o.new_rd = rd4; o.new_beta = beta; o.new_meta = meta;
    const uint32_t depth = META_DEPTH(meta);
    if (h.y == 0xffffffffu) { o.finished = true; return o; }   // background is black, rendering.rs:140-141
    const float3 ro = v3(ro4), rd = v3(rd4);
    const int obj = (int)h.y;
    PT_CHECK(obj >= 0);
    const float t = __uint_as_float(h.x);
    float4 q0, q1, q2, q3;
    sc.prim(obj, q0, q1, q2, q3);
    Mat m = sc.mat(obj);
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
            if (is_tri) { float3 ld; float dd; ps = triangle_pdf_omega(q3.w, v3(q3), ro, p, ld, dd); }
            else ps = sphere_pdf_omega(v3(q0), q1.x, ro);
            float pb = rd4.w;
            L = beta * (m.color * (pb / (pb + ps)));
        }
        o.finished = true; o.add_L = true; o.Ladd = L;
        return o;
    }
    o.vertex = true;
    const uint32_t pixel = META_PIXEL(meta), sample = META_SAMPLE(meta);
    // block 0 = [light r1, light r2, bsdf r1, bsdf r2] (+ the light pick from the low byte of each word); block 1 =
    // [lobe u, roulette u]: only GGX surfaces choose a lobe and roulette starts at min_depth (q == 1 before), so most vertices
    // never generate it.
    uint4 b0 = rng_block(pixel, sample, depth, 0u, f.k0, f.k1);
    uint4 b1 = make_uint4(0u, 0u, 0u, 0u);
    if (m.kind == PT_MATK_MIRROR || depth >= f.min_depth) b1 = rng_block(pixel, sample, depth, 1u, f.k0, f.k1);
    // ---- NEE (rendering.rs:53-81, world.rs:251-267); visibility resolved by the shadow stage
    bool want_shadow = false;
    if (mis && sc.n_lights > 0) {
        uint32_t lw = (b0.x & 0xffu) | ((b0.y & 0xffu) << 8) | ((b0.z & 0xffu) << 16) | ((b0.w & 0xffu) << 24);
        uint32_t li = sc.light(urange(lw, sc.n_lights));
        PT_CHECK(urange(lw, sc.n_lights) < sc.n_lights);
        float4 l0, l1, l2, l3;
        sc.prim((int)li, l0, l1, l2, l3);
        float3 le = sc.emission((int)li);
        float3 ldir; float dist, pdf_shape;
        sample_light_shape(l0, l1, l2, l3, p, u01(b0.x), u01(b0.y), ldir, dist, pdf_shape);
        float pl = pdf_shape / (float)sc.n_lights;        // world.rs:260
        float cos_theta = fabsf(dot(n, ldir));            // rendering.rs:68
        float3 fe; float pb;
        bsdf_eval(m, rd, ro4.w, ldir, n, fe, pb);         // inherited eta_ratio (Q3)
        float w_nee = pl / (pl + pb);
        float3 c = beta * fe * le * (w_nee * cos_theta / pl);
        if (finite3(c) && (c.x != 0.f || c.y != 0.f || c.z != 0.f)) {
            want_shadow = true;
            o.sh_d = ldir; o.sh_c = c; o.sh_tmax = dist - 0.001f;   // rendering.rs:64
        }
    }
    // ---- BSDF sample + Russian roulette (rendering.rs:83-102)
    float eta_new = front ? 1.0f / mat_eta(m) : mat_eta(m);   // rendering.rs:20-25
    float3 wo, fs; float pdf, cos_s;
    const bool sampled = bsdf_sample(m, rd, eta_new, n, u01(b0.z), u01(b0.w), u01(b1.x), wo, fs, pdf, cos_s);
    float3 tn = beta * fs * (cos_s / pdf);
    const bool ok = sampled && finite3(tn);
    if (!ok) {
        // Mirror's failure tuple (normal, 0, pdf = 1, cos = 0), mirror.rs:214-217 / :258-265 / :293-300: next_throughput is zero,
        // so before the roulette depth (q == 1, rendering.rs:91-92) the reference keeps this vertex's direct light and nothing
        // after it; from the roulette depth on q == 0 kills the path together with its direct light (Q1).  A non-positive pdf of
        // the cosine-weighted samplers ends the path without its direct light (documented deviation).
        o.finished = true;
        if (!sampled && m.kind == PT_MATK_MIRROR && depth < f.min_depth && want_shadow) { o.emit_shadow = true; o.new_ro = make_float4(p.x, p.y, p.z, eta_new); }
        return o;
    }
    float q = rr_prob(f, tn, depth);
    if (u01(b1.y) > q || depth >= kMaxPathDepth) {
        o.finished = true;                                  // D of this vertex is dropped too (Q1)
        return o;
    }
    o.emit_shadow = want_shadow;
    o.new_beta = tn / q;
    o.new_ro = make_float4(p.x, p.y, p.z, eta_new);         // rendering.rs:86-87
    o.new_rd = make_float4(wo.x, wo.y, wo.z, pdf);
    o.new_meta = pack_meta(pixel, sample, depth + 1, true);
    return o;
}

// ------------------------------------------------------------------------------------ host helpers
inline SceneView make_scene_view(const SceneDev& d) {
    SceneView v;
    v.prims = d.prims; v.mats = d.mats; v.light_ids = d.light_ids; v.nodes = d.nodes;
    v.n_objects = d.n_objects; v.n_lights = d.n_lights; v.n_nodes = d.n_nodes;
    v.brute = d.brute; v.n_brute = d.n_brute; v.n_pairs = d.n_pairs;
    float mx = 0.f;
    for (int k = 0; k < 6; ++k) if (std::isfinite(d.bounds[k]) && std::fabs(d.bounds[k]) < 1e30f) mx = std::fmax(mx, std::fabs(d.bounds[k]));
    v.box_pad = 8.0f * 5.9604645e-8f * mx; v.ro_near = 4.0f * mx;
    return v;
}
// Which form of the brute-force scan a scene / camera / PT_TRAVERSAL_* value gets (pt_stats.scan_mode): 1 = scan table from
// shared memory, checked pair test; 2 = triangle pairs from the constant bank, checked; 3 = constant bank, in-domain (no
// reciprocal range check, no `u <= 1`: pt_math.cuh).  Shared by the shared-memory wavefront and pt_trace_primary so that the
// parity hook runs exactly the scan the render runs.
// 4 = 3 for primary rays + the planar-entry scan (fast_pair) for shadow and scattered rays: the default of small scenes.
inline int pick_scan_mode(const SceneDev& scene, const pt_camera& cam, uint32_t traversal) {
    const bool param_ok = scene.brute && scene.n_pairs > 0 && scene.n_pairs <= (uint32_t)kParamPairs && !scene.brute_host.empty();
    if (!param_ok || traversal == PT_TRAVERSAL_BRUTE_TABLE) return 1;
    bool in_domain = scene.fp32_domain;
    for (int k = 0; k < 3; ++k) if (!(fabsf(cam.origin[k]) <= kFp32DomainMax)) in_domain = false;
    if (!in_domain || traversal == PT_TRAVERSAL_BRUTE_PARAM_CHECKED) return 2;
    return (scene.n_fast_pairs > 0 && traversal != PT_TRAVERSAL_BRUTE_EXACT) ? 4 : 3;
}
inline FastParam make_fast_param(const SceneDev& scene) {   // right-aligned like BruteParam (PT_FAST_SCAN)
    FastParam fp;
    memset(&fp, 0, sizeof(fp));
    if (scene.n_fast_pairs && scene.n_fast_pairs <= (uint32_t)kFastPairs)
        memcpy(fp.e + 8 * ((size_t)kFastPairs - scene.n_fast_pairs), scene.fast_host.data(), sizeof(float4) * 8 * (size_t)scene.n_fast_pairs);
    for (uint32_t k = 0; k < scene.n_fast_spheres; ++k) {
        fp.sph[kFastSpheres - scene.n_fast_spheres + k] = scene.fast_spheres[k];
        fp.sph_id[kFastSpheres - scene.n_fast_spheres + k] = scene.fast_sphere_ids[k];
    }
    return fp;
}
inline BruteParam make_brute_param(const SceneDev& scene) {   // right-aligned: the scan is entered at slot kParamPairs - n_pairs (PT_PARAM_SCAN)
    BruteParam bp;
    memset(&bp, 0, sizeof(bp));
    if (scene.n_pairs <= (uint32_t)kParamPairs && !scene.brute_host.empty())
        memcpy(bp.e + 5 * ((size_t)kParamPairs - scene.n_pairs), scene.brute_host.data(), sizeof(float4) * 5 * (size_t)scene.n_pairs);
    return bp;
}
// Pixel rectangles of the primary-ray cull (closest_hit_param_culled).  A point P is seen through image-plane coordinates
// (u, v) with P - O = s (A + u H + v V), A = lower_left_corner - O, s > 0; pixel x covers u (W - 1) in (x, x + 1).  The bounds
// are computed in f64 from the same f32 camera and vertex values the device uses and padded by 2 pixels: the FP32 test can
// accept a ray that misses the exact triangle by rounding, which is far below a pixel unless the plane is seen edge-on
// (|cos| < 1e-2 between the plane and the line of sight: such entries are not culled).
inline CullParam make_cull_param(const SceneDev& scene, const pt_camera& cam, bool enable) {
    CullParam cp;
    const int4 full = make_int4(0, 0, 0x7fffffff, 0x7fffffff), none = make_int4(0x7fffffff, 0x7fffffff, 0, 0);
    for (auto& r : cp.pair) r = full;
    for (auto& r : cp.single) r = full;
    cp.enabled = 0;
    const uint32_t n = scene.n_objects;
    if (!enable || scene.shapes_host.size() != n || scene.n_pairs > (uint32_t)kParamPairs) return cp;
    const double O[3] = {cam.origin[0], cam.origin[1], cam.origin[2]};
    double M[3][3];   // columns H, V, A
    for (int r = 0; r < 3; ++r) { M[r][0] = cam.horizontal[r]; M[r][1] = cam.vertical[r]; M[r][2] = (double)cam.lower_left_corner[r] - O[r]; }
    const double det = M[0][0] * (M[1][1] * M[2][2] - M[1][2] * M[2][1]) - M[0][1] * (M[1][0] * M[2][2] - M[1][2] * M[2][0]) +
                       M[0][2] * (M[1][0] * M[2][1] - M[1][1] * M[2][0]);
    if (!(std::fabs(det) > 1e-300) || !std::isfinite(det)) return cp;
    double inv[3][3];
    inv[0][0] = (M[1][1] * M[2][2] - M[1][2] * M[2][1]) / det; inv[0][1] = (M[0][2] * M[2][1] - M[0][1] * M[2][2]) / det; inv[0][2] = (M[0][1] * M[1][2] - M[0][2] * M[1][1]) / det;
    inv[1][0] = (M[1][2] * M[2][0] - M[1][0] * M[2][2]) / det; inv[1][1] = (M[0][0] * M[2][2] - M[0][2] * M[2][0]) / det; inv[1][2] = (M[0][2] * M[1][0] - M[0][0] * M[1][2]) / det;
    inv[2][0] = (M[1][0] * M[2][1] - M[1][1] * M[2][0]) / det; inv[2][1] = (M[0][1] * M[2][0] - M[0][0] * M[2][1]) / det; inv[2][2] = (M[0][0] * M[1][1] - M[0][1] * M[1][0]) / det;
    struct Rect { double x0 = 1e300, x1 = -1e300, y0 = 1e300, y1 = -1e300; bool ok = true; };
    auto add_point = [&](Rect& rc, const double P[3]) {
        const double d[3] = {P[0] - O[0], P[1] - O[1], P[2] - O[2]};
        const double su = inv[0][0] * d[0] + inv[0][1] * d[1] + inv[0][2] * d[2], sv = inv[1][0] * d[0] + inv[1][1] * d[1] + inv[1][2] * d[2],
                     sw = inv[2][0] * d[0] + inv[2][1] * d[1] + inv[2][2] * d[2];
        if (!(sw > 1e-4) || !std::isfinite(su) || !std::isfinite(sv) || !std::isfinite(sw)) { rc.ok = false; return; }   // behind (or near) the camera plane
        const double xf = su / sw * ((double)cam.width - 1.0), yf = sv / sw * ((double)cam.height - 1.0);
        if (!(std::fabs(xf) < 1e9) || !(std::fabs(yf) < 1e9)) { rc.ok = false; return; }
        rc.x0 = std::fmin(rc.x0, xf); rc.x1 = std::fmax(rc.x1, xf); rc.y0 = std::fmin(rc.y0, yf); rc.y1 = std::fmax(rc.y1, yf);
    };
    auto add_triangle = [&](Rect& rc, const pt_shape& sh) {
        double Pv[3][3];
        for (int v = 0; v < 3; ++v) for (int c = 0; c < 3; ++c) Pv[v][c] = sh.v[3 * v + c];
        const double e1[3] = {Pv[1][0] - Pv[0][0], Pv[1][1] - Pv[0][1], Pv[1][2] - Pv[0][2]}, e2[3] = {Pv[2][0] - Pv[0][0], Pv[2][1] - Pv[0][1], Pv[2][2] - Pv[0][2]};
        const double nn[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
        const double nl = std::sqrt(nn[0] * nn[0] + nn[1] * nn[1] + nn[2] * nn[2]);
        if (!(nl > 0.0) || !std::isfinite(nl)) { rc.ok = false; return; }
        double far2 = 0.0;
        for (int v = 0; v < 3; ++v) { double q = 0; for (int c = 0; c < 3; ++c) q += (Pv[v][c] - O[c]) * (Pv[v][c] - O[c]); far2 = std::fmax(far2, q); }
        const double dist = std::fabs(nn[0] * (O[0] - Pv[0][0]) + nn[1] * (O[1] - Pv[0][1]) + nn[2] * (O[2] - Pv[0][2])) / nl;
        if (!(dist >= 1e-2 * std::sqrt(far2))) { rc.ok = false; return; }                                        // seen edge-on
        for (int v = 0; v < 3; ++v) add_point(rc, Pv[v]);
    };
    auto add_sphere = [&](Rect& rc, const pt_shape& sh) {
        const double c[3] = {sh.v[0], sh.v[1], sh.v[2]}, r = std::fabs((double)sh.v[3]) * (1.0 + 1e-6);
        double d2 = 0; for (int k = 0; k < 3; ++k) d2 += (O[k] - c[k]) * (O[k] - c[k]);
        if (!(std::sqrt(d2) > r * 1.001 + 1e-30) || !std::isfinite(r)) { rc.ok = false; return; }                 // camera inside or on the sphere
        for (int m = 0; m < 8; ++m) { const double P[3] = {c[0] + ((m & 1) ? r : -r), c[1] + ((m & 2) ? r : -r), c[2] + ((m & 4) ? r : -r)}; add_point(rc, P); }
    };
    auto to_int4 = [&](const Rect& rc) {
        if (!rc.ok) return full;
        const double pad = 2.0;
        const double x0 = std::floor(rc.x0) - pad, x1 = std::floor(rc.x1) + pad, y0 = std::floor(rc.y0) - pad, y1 = std::floor(rc.y1) + pad;
        const double W1 = (double)cam.width - 1.0, H1 = (double)cam.height - 1.0;
        if (x1 < 0.0 || y1 < 0.0 || x0 > W1 || y0 > H1) return none;
        const int ix0 = (int)std::fmax(x0, 0.0), ix1 = (int)std::fmin(x1, W1), iy0 = (int)std::fmax(y0, 0.0), iy1 = (int)std::fmin(y1, H1);
        return make_int4(ix0, iy0, ix1 - ix0, iy1 - iy0);
    };
    uint32_t ep = 0, es = 0;
    for (uint32_t i = 0; i < n;) {                       // the pairing rule of k_brute_table (lbvh.cu)
        const bool pair = i + 1 < n && scene.shapes_host[i].kind == PT_SHAPE_TRIANGLE && scene.shapes_host[i + 1].kind == PT_SHAPE_TRIANGLE;
        Rect rc;
        if (pair) {
            add_triangle(rc, scene.shapes_host[i]); add_triangle(rc, scene.shapes_host[i + 1]);
            if (ep < scene.n_pairs) cp.pair[(size_t)kParamPairs - scene.n_pairs + ep] = to_int4(rc);
            ++ep; i += 2;
        } else {
            if (scene.shapes_host[i].kind == PT_SHAPE_TRIANGLE) add_triangle(rc, scene.shapes_host[i]); else add_sphere(rc, scene.shapes_host[i]);
            if (es < (uint32_t)kCullSingles) cp.single[es] = to_int4(rc);
            ++es; i += 1;
        }
    }
    cp.enabled = 1;
    return cp;
}
inline Frame make_frame(const pt_camera& c, const pt_render_params& p, float4* accum) {
    Frame f;
    f.cam.origin = make_float3(c.origin[0], c.origin[1], c.origin[2]);
    f.cam.llc = make_float3(c.lower_left_corner[0], c.lower_left_corner[1], c.lower_left_corner[2]);
    f.cam.horizontal = make_float3(c.horizontal[0], c.horizontal[1], c.horizontal[2]);
    f.cam.vertical = make_float3(c.vertical[0], c.vertical[1], c.vertical[2]);
    f.cam.width = c.width; f.cam.height = c.height;
    f.npix = c.width * c.height;
    f.magic_npix = magic_of(f.npix); f.magic_width = magic_of(c.width);
    f.total_work = (uint64_t)f.npix * p.spp;
    f.sample_offset = p.sample_offset; f.min_depth = p.min_depth; f.max_depth = p.max_depth; f.strategy = p.strategy;
    f.k0 = (uint32_t)p.seed; f.k1 = (uint32_t)(p.seed >> 32);
    f.t_min = p.t_min;
    f.accum = accum;
    f.blo = make_float3(1.f, 1.f, 1.f); f.bhi = make_float3(-1.f, -1.f, -1.f);
    return f;
}

}  // namespace ptw
