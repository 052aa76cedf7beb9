// pt_lbvh_ref.hpp — CPU twin of the GPU LBVH build (ORACLE / test infrastructure, not product).
//
// The reference has no BVH: src/world.rs:270-290 scans every object.  The north star asks for a
// GPU-built LBVH whose build is bit-exact against a CPU oracle, so this file is that oracle: a
// deterministic sequential restatement of the build SPEC in DESIGN.md §LBVH (Karras 2012,
// "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d Trees").  It is written
// independently of pathtrace_b200/csrc/lbvh.cu (sequential std::stable_sort + recursion-free
// range walk here; radix sort + per-node binary search + atomic bottom-up refit there).
//
// SPEC (all arithmetic IEEE binary32, no contraction except where fmaf is written):
//  1. prim AABB: sphere lo=c-r, hi=c+r; triangle lo/hi = min/max of the vertices; then each
//     bound v is pushed outward by pad(v) = fmaf(|v|, 2^-20, 1e-6f).
//  2. scene bounds = min/max over the padded prim AABBs.
//  3. centroid c = (lo+hi)*0.5f; n = (c - smin)/(smax - smin) per axis (0 when the extent is 0);
//     q = (uint32) min(max(n*1024.0f, 0), 1023); code = spread(qx)<<2 | spread(qy)<<1 | spread(qz).
//  4. stable sort by code (== sort by the 64-bit key code<<32 | original index).
//  5. hierarchy over the n sorted leaves: internal nodes 0..n-2, root 0, Karras' determine-range /
//     find-split with delta(i,j) = clz64((code_i<<32|i) ^ (code_j<<32|j)).
//  6. node = 16 x 32-bit words: [Llo.xyz, left] [Lhi.xyz, right] [Rlo.xyz, 0] [Rhi.xyz, 0];
//     child >= 0: internal node index; child < 0: leaf, ~child = ORIGINAL object index.
//     n == 1: node 0 = (leaf, empty); n == 0: node 0 = (empty, empty); empty box = lo +FLT_MAX,
//     hi -FLT_MAX, child word = 0x80000000 ... see kEmptyChild.
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "pt_oracle.hpp"

namespace pto {

constexpr int32_t kEmptyChild = INT32_MIN;   // ~kEmptyChild = 0x7fffffff: never a valid object

struct Aabb { float lo[3], hi[3]; };

inline float pad_amount(float v) { return std::fmaf(std::fabs(v), 9.5367431640625e-07f, 1e-6f); }

inline Aabb prim_aabb(const ShapeDesc& s) {
    Aabb b;
    if (s.kind == SHAPE_SPHERE) {
        for (int k = 0; k < 3; ++k) { b.lo[k] = s.v[k] - s.v[3]; b.hi[k] = s.v[k] + s.v[3]; }
    } else {
        for (int k = 0; k < 3; ++k) {
            b.lo[k] = std::fmin(std::fmin(s.v[k], s.v[3 + k]), s.v[6 + k]);
            b.hi[k] = std::fmax(std::fmax(s.v[k], s.v[3 + k]), s.v[6 + k]);
        }
    }
    for (int k = 0; k < 3; ++k) { b.lo[k] = b.lo[k] - pad_amount(b.lo[k]); b.hi[k] = b.hi[k] + pad_amount(b.hi[k]); }
    return b;
}
inline uint32_t spread10(uint32_t v) {   // 10 bits -> every third bit
    v &= 0x3ffu;
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
inline uint32_t morton30(const Aabb& b, const float smin[3], const float smax[3]) {
    uint32_t q[3];
    for (int k = 0; k < 3; ++k) {
        float c = (b.lo[k] + b.hi[k]) * 0.5f;
        float ext = smax[k] - smin[k];
        float n = ext > 0.0f ? (c - smin[k]) / ext : 0.0f;
        float s = std::fmin(std::fmax(n * 1024.0f, 0.0f), 1023.0f);
        q[k] = (uint32_t)s;
    }
    return (spread10(q[0]) << 2) | (spread10(q[1]) << 1) | spread10(q[2]);
}

struct LbvhRef {
    uint32_t n = 0;
    float smin[3], smax[3];
    std::vector<uint32_t> codes_sorted, prim_sorted;
    std::vector<uint32_t> nodes;   // 16 words per node, max(n-1,1) nodes
};

inline void put_box(uint32_t* node, int side, const Aabb& b, int32_t child) {
    uint32_t* lo = node + side * 8;
    uint32_t* hi = lo + 4;
    std::memcpy(lo, b.lo, 12);
    std::memcpy(hi, b.hi, 12);
    if (side == 0) { std::memcpy(node + 3, &child, 4); }
    else { std::memcpy(node + 7, &child, 4); lo[3] = 0; hi[3] = 0; }
}

inline LbvhRef lbvh_build_ref(const ShapeDesc* shapes, uint32_t n) {
    LbvhRef r;
    r.n = n;
    Aabb empty;
    for (int k = 0; k < 3; ++k) { empty.lo[k] = FLT_MAX; empty.hi[k] = -FLT_MAX; r.smin[k] = FLT_MAX; r.smax[k] = -FLT_MAX; }
    std::vector<Aabb> boxes(n);
    for (uint32_t i = 0; i < n; ++i) {
        boxes[i] = prim_aabb(shapes[i]);
        for (int k = 0; k < 3; ++k) { r.smin[k] = std::fmin(r.smin[k], boxes[i].lo[k]); r.smax[k] = std::fmax(r.smax[k], boxes[i].hi[k]); }
    }
    std::vector<uint64_t> keys(n);
    for (uint32_t i = 0; i < n; ++i) keys[i] = ((uint64_t)morton30(boxes[i], r.smin, r.smax) << 32) | i;
    std::sort(keys.begin(), keys.end());
    r.codes_sorted.resize(n); r.prim_sorted.resize(n);
    for (uint32_t i = 0; i < n; ++i) { r.codes_sorted[i] = (uint32_t)(keys[i] >> 32); r.prim_sorted[i] = (uint32_t)keys[i]; }
    uint32_t n_nodes = n > 1 ? n - 1 : 1;
    r.nodes.assign((size_t)n_nodes * 16, 0);
    if (n == 0) { put_box(&r.nodes[0], 0, empty, kEmptyChild); put_box(&r.nodes[0], 1, empty, kEmptyChild); return r; }
    if (n == 1) { put_box(&r.nodes[0], 0, boxes[r.prim_sorted[0]], ~(int32_t)r.prim_sorted[0]); put_box(&r.nodes[0], 1, empty, kEmptyChild); return r; }

    auto key64 = [&](int i) { return ((uint64_t)r.codes_sorted[i] << 32) | (uint32_t)i; };
    auto delta = [&](int i, int j) -> int {
        if (j < 0 || j >= (int)n) return -1;
        return __builtin_clzll(key64(i) ^ key64(j));
    };
    std::vector<int32_t> left(n - 1), right(n - 1);
    for (int i = 0; i < (int)n - 1; ++i) {
        int d = (delta(i, i + 1) - delta(i, i - 1)) > 0 ? 1 : -1;
        int dmin = delta(i, i - d);
        int lmax = 2;
        while (delta(i, i + lmax * d) > dmin) lmax *= 2;
        int l = 0;
        for (int t = lmax / 2; t >= 1; t /= 2)
            if (delta(i, i + (l + t) * d) > dmin) l += t;
        int j = i + l * d;
        int dnode = delta(i, j);
        int s = 0;
        int t = l;
        do {
            t = (t + 1) / 2;
            if (delta(i, i + (s + t) * d) > dnode) s += t;
        } while (t > 1);
        int gamma = i + s * d + std::min(d, 0);
        int lo_i = std::min(i, j), hi_i = std::max(i, j);
        left[i] = (lo_i == gamma) ? ~(int32_t)gamma : gamma;           // leaf encoded by SORTED position for now
        right[i] = (hi_i == gamma + 1) ? ~(int32_t)(gamma + 1) : gamma + 1;
    }
    // bottom-up boxes by post-order walk from the root (explicit stack)
    std::vector<Aabb> nb(n - 1);
    std::vector<uint8_t> state(n - 1, 0);
    std::vector<int32_t> stack;
    stack.push_back(0);
    auto child_box = [&](int32_t c) -> Aabb { return c < 0 ? boxes[r.prim_sorted[~c]] : nb[c]; };
    while (!stack.empty()) {
        int32_t v = stack.back();
        if (state[v] == 0) {
            state[v] = 1;
            if (left[v] >= 0) stack.push_back(left[v]);
            if (right[v] >= 0) stack.push_back(right[v]);
        } else {
            stack.pop_back();
            Aabb a = child_box(left[v]), b = child_box(right[v]);
            for (int k = 0; k < 3; ++k) { nb[v].lo[k] = std::fmin(a.lo[k], b.lo[k]); nb[v].hi[k] = std::fmax(a.hi[k], b.hi[k]); }
            auto enc = [&](int32_t c) { return c < 0 ? ~(int32_t)r.prim_sorted[~c] : c; };
            put_box(&r.nodes[(size_t)v * 16], 0, a, enc(left[v]));
            put_box(&r.nodes[(size_t)v * 16], 1, b, enc(right[v]));
        }
    }
    return r;
}

}  // namespace pto
