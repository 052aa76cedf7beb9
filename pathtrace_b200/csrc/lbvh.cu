// lbvh.cu — GPU LBVH build (K2a-e) for sm_100a: padded primitive AABBs + scene bounds, 30-bit Morton
// codes, hand-written LSD radix sort, Karras hierarchy, atomic bottom-up refit.
//
// The reference has no acceleration structure (src/world.rs:270-290 scans all objects); this replaces
// that scan.  The build SPEC is in DESIGN.md §LBVH and is restated sequentially by the CPU twin in
// oracle/pt_lbvh_ref.hpp; tests/test_gpu_parity.py (test_lbvh_bit_exact) requires codes, order, topology and boxes to be
// bit-exact between the two.  All box arithmetic is min/max (order independent) so the atomics below
// cannot perturb the result.
#include <cfloat>

#include "pt_internal.h"
#include "pt_math.cuh"

namespace {

constexpr int kEmptyChild = INT32_MIN;

__device__ __forceinline__ uint32_t f2ord(float f) {   // monotone float -> uint
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}
__device__ __forceinline__ float pad_amount(float v) { return __fmaf_rn(fabsf(v), 9.5367431640625e-07f, 1e-6f); }

// K2a: padded AABB per primitive + scene bounds (ordered-uint atomics, one per warp per component)
__global__ void k_prim_aabb(const pt_shape* __restrict__ shapes, uint32_t n, float4* __restrict__ lo, float4* __restrict__ hi,
                            uint32_t* __restrict__ bounds_ord) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float l[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, h[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < n) {
        const pt_shape s = shapes[i];
        for (int k = 0; k < 3; ++k) {
            if (s.kind == PT_SHAPE_SPHERE) {
                l[k] = __fsub_rn(s.v[k], s.v[3]);
                h[k] = __fadd_rn(s.v[k], s.v[3]);
            } else {
                l[k] = fminf(fminf(s.v[k], s.v[3 + k]), s.v[6 + k]);
                h[k] = fmaxf(fmaxf(s.v[k], s.v[3 + k]), s.v[6 + k]);
            }
            l[k] = __fsub_rn(l[k], pad_amount(l[k]));
            h[k] = __fadd_rn(h[k], pad_amount(h[k]));
        }
        lo[i] = make_float4(l[0], l[1], l[2], 0.f);
        hi[i] = make_float4(h[0], h[1], h[2], 0.f);
    }
    for (int k = 0; k < 3; ++k) {
        float a = l[k], b = h[k];
        for (int o = 16; o > 0; o >>= 1) {
            a = fminf(a, __shfl_xor_sync(0xffffffffu, a, o));
            b = fmaxf(b, __shfl_xor_sync(0xffffffffu, b, o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(bounds_ord + k, f2ord(a));
            atomicMax(bounds_ord + 3 + k, f2ord(b));
        }
    }
}
__global__ void k_bounds_decode(const uint32_t* __restrict__ ord, float* __restrict__ bounds) {
    if (threadIdx.x < 6) bounds[threadIdx.x] = ord2f(ord[threadIdx.x]);
}

__device__ __forceinline__ uint32_t spread10(uint32_t v) {
    v &= 0x3ffu;
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
// K2b: 30-bit Morton code of the padded-box centroid
__global__ void k_morton(const float4* __restrict__ lo, const float4* __restrict__ hi, const float* __restrict__ bounds, uint32_t n,
                         uint32_t* __restrict__ codes, uint32_t* __restrict__ ids) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 l = lo[i], h = hi[i];
    float lv[3] = {l.x, l.y, l.z}, hv[3] = {h.x, h.y, h.z};
    uint32_t q[3];
    for (int k = 0; k < 3; ++k) {
        float c = __fmul_rn(__fadd_rn(lv[k], hv[k]), 0.5f);
        float ext = __fsub_rn(bounds[3 + k], bounds[k]);
        float nrm = ext > 0.0f ? __fdiv_rn(__fsub_rn(c, bounds[k]), ext) : 0.0f;
        float s = fminf(fmaxf(__fmul_rn(nrm, 1024.0f), 0.0f), 1023.0f);
        q[k] = (uint32_t)s;
    }
    codes[i] = (spread10(q[0]) << 2) | (spread10(q[1]) << 1) | spread10(q[2]);
    ids[i] = i;
}

// K2c: LSD radix sort, 4 bits per pass, 1024 keys per CTA (one per thread), stable.
constexpr int kSortThreads = 1024, kRadix = 16;
__global__ void __launch_bounds__(kSortThreads) k_sort_hist(const uint32_t* __restrict__ keys, uint32_t n, int shift,
                                                             uint32_t* __restrict__ hist, uint32_t nblocks) {
    __shared__ uint32_t sh[kRadix];
    if (threadIdx.x < kRadix) sh[threadIdx.x] = 0;
    __syncthreads();
    uint32_t i = blockIdx.x * kSortThreads + threadIdx.x;
    bool valid = i < n;
    uint32_t d = valid ? (keys[i] >> shift) & (kRadix - 1) : 0xffffffffu;
    uint32_t peers = __match_any_sync(0xffffffffu, d);
    if (valid && (uint32_t)(__ffs(peers) - 1) == (threadIdx.x & 31)) atomicAdd(&sh[d], __popc(peers));
    __syncthreads();
    if (threadIdx.x < kRadix) hist[threadIdx.x * nblocks + blockIdx.x] = sh[threadIdx.x];
}
// exclusive scan over the digit-major histogram (single CTA; total = 16 * nblocks entries)
__global__ void __launch_bounds__(1024) k_sort_scan(uint32_t* __restrict__ hist, uint32_t total) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < total; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint32_t v = i < total ? hist[i] : 0, x = v;
        for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((threadIdx.x & 31) >= o) x += y; }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_sums[threadIdx.x], s = w;
            for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, s, o); if (threadIdx.x >= o) s += y; }
            warp_sums[threadIdx.x] = s - w;
        }
        __syncthreads();
        uint32_t excl = carry + warp_sums[threadIdx.x >> 5] + x - v;
        if (i < total) hist[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
}
__global__ void __launch_bounds__(kSortThreads) k_sort_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                                uint32_t n, int shift, const uint32_t* __restrict__ hist, uint32_t nblocks,
                                                                uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out) {
    __shared__ uint32_t wh[32][kRadix];   // per-warp digit counts -> exclusive prefix over warps
    uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane < kRadix) wh[warp][lane] = 0;
    __syncwarp();
    uint32_t i = blockIdx.x * kSortThreads + threadIdx.x;
    bool valid = i < n;
    uint32_t key = valid ? keys_in[i] : 0, val = valid ? vals_in[i] : 0;
    uint32_t d = valid ? (key >> shift) & (kRadix - 1) : 0xffffffffu;
    uint32_t peers = __match_any_sync(0xffffffffu, d);
    uint32_t rank = __popc(peers & ((1u << lane) - 1));
    if (valid && rank == 0) wh[warp][d] = __popc(peers);
    __syncthreads();
    if (threadIdx.x < kRadix) {
        uint32_t run = 0;
        for (int w = 0; w < 32; ++w) { uint32_t c = wh[w][threadIdx.x]; wh[w][threadIdx.x] = run; run += c; }
    }
    __syncthreads();
    if (valid) {
        uint32_t pos = hist[d * nblocks + blockIdx.x] + wh[warp][d] + rank;
        keys_out[pos] = key;
        vals_out[pos] = val;
    }
}

// K2d: Karras 2012 hierarchy.  delta(i,j) = clz64 of the XOR of (code<<32 | position) keys.
__device__ __forceinline__ int delta(const uint32_t* __restrict__ codes, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = ((uint64_t)codes[i] << 32) | (uint32_t)i, b = ((uint64_t)codes[j] << 32) | (uint32_t)j;
    return __clzll((long long)(a ^ b));
}
__global__ void k_hierarchy(const uint32_t* __restrict__ codes, int n, int* __restrict__ child_l, int* __restrict__ child_r,
                            uint32_t* __restrict__ parent_of_node, uint32_t* __restrict__ parent_of_leaf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(codes, n, i, i + 1) - delta(codes, n, i, i - 1)) > 0 ? 1 : -1;
    int dmin = delta(codes, n, i, i - d);
    int lmax = 2;
    while (delta(codes, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta(codes, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(codes, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) / 2;
        if (delta(codes, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo_i = min(i, j), hi_i = max(i, j);
    // children encoded by SORTED leaf position here (~pos); k_refit rewrites leaves to ~original id
    if (lo_i == gamma) { child_l[i] = ~gamma; parent_of_leaf[gamma] = (uint32_t)i << 1; }
    else { child_l[i] = gamma; parent_of_node[gamma] = (uint32_t)i << 1; }
    if (hi_i == gamma + 1) { child_r[i] = ~(gamma + 1); parent_of_leaf[gamma + 1] = ((uint32_t)i << 1) | 1u; }
    else { child_r[i] = gamma + 1; parent_of_node[gamma + 1] = ((uint32_t)i << 1) | 1u; }
}
// K2e: bottom-up refit.  One thread per leaf; the second thread to reach a node continues upward.
__global__ void k_refit(const float4* __restrict__ lo, const float4* __restrict__ hi, const uint32_t* __restrict__ prim_sorted, int n,
                        const int* __restrict__ child_l, const int* __restrict__ child_r, const uint32_t* __restrict__ parent_of_node,
                        const uint32_t* __restrict__ parent_of_leaf, uint32_t* __restrict__ flags, float4* __restrict__ nodes) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    uint32_t prim = prim_sorted[p];
    float4 bl = lo[prim], bh = hi[prim];
    uint32_t par = parent_of_leaf[p];
    int self_enc = ~(int)prim;   // what the parent stores for this child
    while (true) {
        uint32_t node = par >> 1, side = par & 1u;
        float4* nd = nodes + (size_t)node * 4;
        if (side == 0) {
            nd[0] = make_float4(bl.x, bl.y, bl.z, __int_as_float(self_enc));
            // word 7 (right child id) lives in nd[1].w: written by the right-side thread below
            volatile float* w = reinterpret_cast<volatile float*>(nd + 1);
            w[0] = bh.x; w[1] = bh.y; w[2] = bh.z;
        } else {
            nd[2] = make_float4(bl.x, bl.y, bl.z, 0.f);
            nd[3] = make_float4(bh.x, bh.y, bh.z, 0.f);
            reinterpret_cast<volatile float*>(nd + 1)[3] = __int_as_float(self_enc);
        }
        __threadfence();
        if (atomicAdd(flags + node, 1u) == 0) return;   // first arrival: the sibling finishes this node
        __threadfence();
        // second arrival: union with the sibling's box, already visible
        volatile float* v = reinterpret_cast<volatile float*>(nd);
        int o = side == 0 ? 8 : 0;   // sibling slot
        float slx = v[o + 0], sly = v[o + 1], slz = v[o + 2], shx = v[o + 4], shy = v[o + 5], shz = v[o + 6];
        bl = make_float4(fminf(bl.x, slx), fminf(bl.y, sly), fminf(bl.z, slz), 0.f);
        bh = make_float4(fmaxf(bh.x, shx), fmaxf(bh.y, shy), fmaxf(bh.z, shz), 0.f);
        if (node == 0) return;
        self_enc = (int)node;
        par = parent_of_node[node];
    }
}
__global__ void k_trivial_tree(const float4* lo, const float4* hi, int n, float4* nodes) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float4 el = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, 0.f), eh = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, 0.f);
    float4 l = n ? lo[0] : el, h = n ? hi[0] : eh;
    nodes[0] = make_float4(l.x, l.y, l.z, __int_as_float(n ? ~0 : kEmptyChild));
    nodes[1] = make_float4(h.x, h.y, h.z, __int_as_float(kEmptyChild));
    nodes[2] = el;
    nodes[3] = eh;
}

// Scene packing: pt_shape / pt_material rows -> 64 B primitive records + 32 B material records
__global__ void k_pack_scene(const pt_shape* __restrict__ shapes, const pt_material* __restrict__ mats, uint32_t n,
                             float4* __restrict__ prims, float4* __restrict__ omats) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const pt_shape s = shapes[i];
    float3 a = make_float3(s.v[0], s.v[1], s.v[2]);
    if (s.kind == PT_SHAPE_TRIANGLE) {
        float3 e1 = ptd::xsub(make_float3(s.v[3], s.v[4], s.v[5]), a), e2 = ptd::xsub(make_float3(s.v[6], s.v[7], s.v[8]), a);
        float3 cr = ptd::xcross(e1, e2);
        float3 nn = ptd::xnormalize(cr);   // shape.rs:195
        float area = __fmul_rn(__fsqrt_rn(ptd::xdot(cr, cr)), 0.5f);   // shape.rs:228, used by the light pdf only
        prims[4 * i + 0] = make_float4(a.x, a.y, a.z, __uint_as_float(PT_KIND_TRIANGLE));
        prims[4 * i + 1] = make_float4(e1.x, e1.y, e1.z, 0.f);
        prims[4 * i + 2] = make_float4(e2.x, e2.y, e2.z, 0.f);
        prims[4 * i + 3] = make_float4(nn.x, nn.y, nn.z, area);
    } else {
        prims[4 * i + 0] = make_float4(a.x, a.y, a.z, __uint_as_float(PT_KIND_SPHERE));
        prims[4 * i + 1] = make_float4(s.v[3], 0.f, 0.f, 0.f);
        prims[4 * i + 2] = make_float4(0.f, 0.f, 0.f, 0.f);
        prims[4 * i + 3] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const pt_material m = mats[i];
    float4 m0, m1;
    switch (m.kind) {
        case PT_MAT_MIRROR:
            m0 = make_float4(m.p[1], m.p[2], m.p[3], __uint_as_float(m.kind));
            m1 = make_float4(m.p[0], m.p[4], m.p[5], 0.f);
            break;
        case PT_MAT_OREN_NAYAR:
            m0 = make_float4(m.p[0], m.p[1], m.p[2], __uint_as_float(m.kind));
            m1 = make_float4(m.p[3], m.p[4], m.p[5], 0.f);   // roughness, a, b
            break;
        default:
            m0 = make_float4(m.p[0], m.p[1], m.p[2], __uint_as_float(m.kind));
            m1 = make_float4(0.f, 0.f, 1.f, 0.f);
            break;
    }
    omats[2 * i] = m0;
    omats[2 * i + 1] = m1;
}

}  // namespace

// Brute-force scan table (small scenes).  Two consecutive triangles share an entry so the trace kernels can test them with
// packed FP32 instructions; the pair entries come first (ascending object order), the unpaired primitives after them.
// 5 float4 per entry:
//   pair   : (v0x_a,v0x_b,v0y_a,v0y_b) (v0z_a,v0z_b,e1x_a,e1x_b) (e1y_a,e1y_b,e1z_a,e1z_b) (e2x_a,e2x_b,e2y_a,e2y_b) (e2z_a,e2z_b,id_a,id_b)
//   single : q0 q1 q2 of the primitive record, -, (-,-,id,-1)
// One thread: at most 1024 objects, once per scene.  The host computes the same entry counts from the shape kinds.
__global__ void k_brute_table(const float4* __restrict__ prims, uint32_t n, uint32_t n_pairs, float4* __restrict__ table) {
    if (blockIdx.x || threadIdx.x) return;
    uint32_t ep = 0, es = n_pairs;
    for (uint32_t i = 0; i < n;) {
        float4 a0 = prims[4 * i], a1 = prims[4 * i + 1], a2 = prims[4 * i + 2];
        bool pair = i + 1 < n && __float_as_uint(a0.w) == PT_KIND_TRIANGLE && __float_as_uint(prims[4 * (i + 1)].w) == PT_KIND_TRIANGLE;
        if (pair) {
            float4 b0 = prims[4 * (i + 1)], b1 = prims[4 * (i + 1) + 1], b2 = prims[4 * (i + 1) + 2];
            table[5 * ep + 0] = make_float4(a0.x, b0.x, a0.y, b0.y);
            table[5 * ep + 1] = make_float4(a0.z, b0.z, a1.x, b1.x);
            table[5 * ep + 2] = make_float4(a1.y, b1.y, a1.z, b1.z);
            table[5 * ep + 3] = make_float4(a2.x, b2.x, a2.y, b2.y);
            table[5 * ep + 4] = make_float4(a2.z, b2.z, __int_as_float((int)i), __int_as_float((int)i + 1));
            i += 2; ++ep;
        } else {
            table[5 * es + 0] = a0; table[5 * es + 1] = a1; table[5 * es + 2] = a2;
            table[5 * es + 3] = make_float4(0.f, 0.f, 0.f, 0.f);
            table[5 * es + 4] = make_float4(0.f, 0.f, __int_as_float((int)i), __int_as_float(-1));
            i += 1; ++es;
        }
    }
}

int scene_pack_launch(pt_context* ctx, const pt_shape* d_shapes, const pt_material* d_mats, uint32_t n, SceneDev& out) {
    PT_CUDA_CHECK(ctx, cudaMalloc(&out.prims, sizeof(float4) * 4 * (size_t)(n ? n : 1)));
    PT_CUDA_CHECK(ctx, cudaMalloc(&out.mats, sizeof(float4) * 2 * (size_t)(n ? n : 1)));
    if (n) {
        k_pack_scene<<<(n + 255) / 256, 256, 0, ctx->stream>>>(d_shapes, d_mats, n, out.prims, out.mats);
        PT_CUDA_CHECK(ctx, cudaGetLastError());
        if (out.n_brute) {      // set by pt_scene_create for scenes small enough to be scanned from shared memory
            PT_CUDA_CHECK(ctx, cudaMalloc(&out.brute, sizeof(float4) * 5 * (size_t)out.n_brute));
            k_brute_table<<<1, 32, 0, ctx->stream>>>(out.prims, n, out.n_pairs, out.brute);
            PT_CUDA_CHECK(ctx, cudaGetLastError());
            const uint32_t np = out.n_pairs < (uint32_t)kBruteParamPairs ? out.n_pairs : (uint32_t)kBruteParamPairs;
            out.brute_host.assign(5 * (size_t)np, make_float4(0.f, 0.f, 0.f, 0.f));
            if (np) {
                PT_CUDA_CHECK(ctx, cudaMemcpyAsync(out.brute_host.data(), out.brute, sizeof(float4) * 5 * np, cudaMemcpyDeviceToHost, ctx->stream));
                PT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
            }
        }
    }
    return PT_OK;
}

int lbvh_build_device(pt_context* ctx, const pt_shape* d_shapes, uint32_t n, SceneDev& out, uint64_t* launches) {
    cudaStream_t st = ctx->stream;
    uint64_t nl = 0;
    uint32_t n1 = n ? n : 1;
    out.n_nodes = n > 1 ? n - 1 : 1;
    // One arena for every temporary (a cudaMalloc per buffer cost more than the whole build: 14 allocations ~ 30 ms against
    // ~0.7 ms of kernels for 1 M primitives), released on every return path; the build is timed around the kernels only.
    DevTemps tmp;
    const size_t nn = n > 1 ? n : 0;
    const uint32_t nblocks = n > 1 ? (n + kSortThreads - 1) / kSortThreads : 0;
    size_t off = 0;
    auto carve = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~size_t(255); return o; };
    const size_t o_lo = carve(sizeof(float4) * n1), o_hi = carve(sizeof(float4) * n1), o_bord = carve(6 * 4), o_bnd = carve(6 * 4);
    const size_t o_ka = carve(4 * nn), o_kb = carve(4 * nn), o_va = carve(4 * nn), o_vb = carve(4 * nn), o_hist = carve(4 * (size_t)kRadix * nblocks);
    const size_t o_cl = carve(4 * nn), o_cr = carve(4 * nn), o_pn = carve(4 * nn), o_pl = carve(4 * nn), o_fl = carve(4 * nn);
    char* arena = nullptr;
    PT_CUDA_CHECK(ctx, tmp.alloc(&arena, off));
    float4 *lo = reinterpret_cast<float4*>(arena + o_lo), *hi = reinterpret_cast<float4*>(arena + o_hi);
    uint32_t *bounds_ord = reinterpret_cast<uint32_t*>(arena + o_bord), *keys_a = reinterpret_cast<uint32_t*>(arena + o_ka),
             *keys_b = reinterpret_cast<uint32_t*>(arena + o_kb), *vals_a = reinterpret_cast<uint32_t*>(arena + o_va),
             *vals_b = reinterpret_cast<uint32_t*>(arena + o_vb), *hist = reinterpret_cast<uint32_t*>(arena + o_hist);
    float* d_bounds = reinterpret_cast<float*>(arena + o_bnd);
    int *cl = reinterpret_cast<int*>(arena + o_cl), *cr = reinterpret_cast<int*>(arena + o_cr);
    uint32_t *pn = reinterpret_cast<uint32_t*>(arena + o_pn), *pl = reinterpret_cast<uint32_t*>(arena + o_pl), *flags = reinterpret_cast<uint32_t*>(arena + o_fl);
    PT_CUDA_CHECK(ctx, cudaMalloc(&out.nodes, sizeof(float4) * 4 * (size_t)out.n_nodes));   // owned by the scene: pt_scene_destroy
    PT_CUDA_CHECK(ctx, cudaMalloc(&out.codes_sorted, 4 * (size_t)n1));
    PT_CUDA_CHECK(ctx, cudaMalloc(&out.prim_sorted, 4 * (size_t)n1));
    PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev[2], st));
    const uint32_t init_ord[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    PT_CUDA_CHECK(ctx, cudaMemcpyAsync(bounds_ord, init_ord, sizeof(init_ord), cudaMemcpyHostToDevice, st));
    if (n) {
        k_prim_aabb<<<(n + 255) / 256, 256, 0, st>>>(d_shapes, n, lo, hi, bounds_ord); nl++;
        k_bounds_decode<<<1, 32, 0, st>>>(bounds_ord, d_bounds); nl++;
        PT_CUDA_CHECK(ctx, cudaMemcpyAsync(out.bounds, d_bounds, 24, cudaMemcpyDeviceToHost, st));
    } else {
        for (int k = 0; k < 3; ++k) { out.bounds[k] = FLT_MAX; out.bounds[3 + k] = -FLT_MAX; }
    }
    if (n <= 1) {
        k_trivial_tree<<<1, 32, 0, st>>>(lo, hi, (int)n, out.nodes); nl++;
        if (n == 1) { PT_CUDA_CHECK(ctx, cudaMemsetAsync(out.codes_sorted, 0, 4, st)); PT_CUDA_CHECK(ctx, cudaMemsetAsync(out.prim_sorted, 0, 4, st)); }
    } else {
        k_morton<<<(n + 255) / 256, 256, 0, st>>>(lo, hi, d_bounds, n, keys_a, vals_a); nl++;
        uint32_t *ki = keys_a, *ko = keys_b, *vi = vals_a, *vo = vals_b;
        for (int shift = 0; shift < 32; shift += 4) {   // 8 passes; the top 2 bits are zero (30-bit codes)
            k_sort_hist<<<nblocks, kSortThreads, 0, st>>>(ki, n, shift, hist, nblocks);
            k_sort_scan<<<1, 1024, 0, st>>>(hist, kRadix * nblocks);
            k_sort_scatter<<<nblocks, kSortThreads, 0, st>>>(ki, vi, n, shift, hist, nblocks, ko, vo);
            nl += 3;
            uint32_t* t = ki; ki = ko; ko = t; t = vi; vi = vo; vo = t;
        }
        PT_CUDA_CHECK(ctx, cudaMemcpyAsync(out.codes_sorted, ki, 4 * (size_t)n, cudaMemcpyDeviceToDevice, st));
        PT_CUDA_CHECK(ctx, cudaMemcpyAsync(out.prim_sorted, vi, 4 * (size_t)n, cudaMemcpyDeviceToDevice, st));
        PT_CUDA_CHECK(ctx, cudaMemsetAsync(flags, 0, 4 * (size_t)n, st));
        k_hierarchy<<<(n - 1 + 255) / 256, 256, 0, st>>>(out.codes_sorted, (int)n, cl, cr, pn, pl); nl++;
        k_refit<<<(n + 255) / 256, 256, 0, st>>>(lo, hi, out.prim_sorted, (int)n, cl, cr, pn, pl, flags, out.nodes); nl++;
    }
    PT_CUDA_CHECK(ctx, cudaGetLastError());
    PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev[3], st));
    PT_CUDA_CHECK(ctx, cudaStreamSynchronize(st));
    if (launches) *launches += nl;
    return PT_OK;
}
