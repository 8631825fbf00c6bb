// bvh_build.cu — GPU acceleration-structure builder: replaces the driver BLAS build
// (src/RaytracingPass.zig:316-361,451-456) and the TLAS build (src/Tlas.zig:25-53,79-162).
//
// Pipeline per mesh (one BLAS per glTF mesh, one geometry per primitive, RaytracingPass.zig:277-352)
// and once more over the instances (TLAS):
//   gather leaves -> 63-bit Morton codes -> radix sort (cub) -> binary hierarchy -> top-down collapse
//   into 8-wide nodes with greedy surface-area expansion, octant slot assignment and 8-bit quantised
//   child boxes (80-byte nodes, Ylitie et al. 2017).
// The binary hierarchy is the Karras 2012 radix tree + bottom-up refit (LBVH) by default, or, with
// PT_FLAG_BVH_PLOC, parallel locally-ordered clustering (PLOC, Meister & Bittner 2018): clusters
// along the Morton curve repeatedly merge with their mutually nearest neighbour (smallest merged
// surface area within a window).  On the regularly tessellated synthetic scenes both give the same
// traversal cost (profiles/r1 notes: 14.7 vs 14.9 nodes per ray on C2), so the cheaper LBVH is the default.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cstdio>
#include <cstdlib>

#include <algorithm>
#include <vector>

#include "bvh_build.h"

namespace pt {

namespace {

constexpr int kThreads = 256;
inline int blocks_for(size_t n) { return (int)((n + kThreads - 1) / kThreads); }

// ------------------------------------------------------------------ float atomics on bounds
// (v + 0.0f turns -0.0f into +0.0f: as an int, -0.0f is INT_MIN and would win every signed minimum)
__device__ __forceinline__ void atomic_min_f(float* a, float v) {
    v += 0.0f;
    if (v >= 0.0f) atomicMin(reinterpret_cast<int*>(a), __float_as_int(v));
    else atomicMax(reinterpret_cast<unsigned int*>(a), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f(float* a, float v) {
    v += 0.0f;
    if (v >= 0.0f) atomicMax(reinterpret_cast<int*>(a), __float_as_int(v));
    else atomicMin(reinterpret_cast<unsigned int*>(a), __float_as_uint(v));
}

}  // namespace
}  // namespace pt

#include "bvh_sah.cuh"   // opt-in stages: binned-SAH hierarchy, cost table of the optimal collapse

namespace pt {
namespace {

__global__ void k_init_bounds(float* b) {
    if (threadIdx.x < 3) b[threadIdx.x] = 3.0e38f;
    else if (threadIdx.x < 6) b[threadIdx.x] = -3.0e38f;
}

// ------------------------------------------------------------------ leaves of a BLAS: triangles
// One thread per triangle of the mesh, (geometry, triangle) order.  prim_prefix[k] = number of
// triangles in primitives [first_prim, first_prim + k).
__global__ void k_gather_tris(const uint8_t* __restrict__ blob, const PrimRec* __restrict__ prims, uint32_t first_prim,
                              uint32_t num_prims, const uint32_t* __restrict__ prim_prefix, uint32_t n,
                              Bvh8Tri* __restrict__ out, float* __restrict__ leaf_lo, float* __restrict__ leaf_hi,
                              float* __restrict__ bounds) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // binary search: last k with prim_prefix[k] <= i
    uint32_t lo = 0, hi = num_prims;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (prim_prefix[mid] <= i) lo = mid; else hi = mid;
    }
    const uint32_t prim = first_prim + lo;
    const uint32_t tri = i - prim_prefix[lo];
    const PrimRec p = prims[prim];
    uint32_t id[3];
    if ((p.info & PT_UINT32_INDICES_MASK) == 0) {
        const uint16_t* q = reinterpret_cast<const uint16_t*>(blob + p.indices_offset) + 3u * tri;
        id[0] = q[0]; id[1] = q[1]; id[2] = q[2];
    } else {
        const uint32_t* q = reinterpret_cast<const uint32_t*>(blob + p.indices_offset) + 3u * tri;
        id[0] = q[0]; id[1] = q[1]; id[2] = q[2];
    }
    float v[3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float* q = reinterpret_cast<const float*>(blob + p.positions_offset + 12u * id[k]);
        v[k][0] = q[0]; v[k][1] = q[1]; v[k][2] = q[2];
    }
    Bvh8Tri t;
    t.v0 = make_float4(v[0][0], v[0][1], v[0][2], __uint_as_float(prim));
    t.v1 = make_float4(v[1][0], v[1][1], v[1][2], __uint_as_float(tri));
    t.v2 = make_float4(v[2][0], v[2][1], v[2][2], 0.0f);
    out[i] = t;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float mn = fminf(v[0][a], fminf(v[1][a], v[2][a]));
        float mx = fmaxf(v[0][a], fmaxf(v[1][a], v[2][a]));
        leaf_lo[3 * (size_t)i + a] = mn;
        leaf_hi[3 * (size_t)i + a] = mx;
        atomic_min_f(bounds + a, mn);
        atomic_max_f(bounds + 3 + a, mx);
    }
}

// ------------------------------------------------------------------ leaves of the TLAS: instances
// World box of instance `ids[i]`: the 8 corners of the BLAS root box through the 3x4 transform,
// evaluated in double and padded (conservative).
__global__ void k_instance_boxes(const PtInstance* __restrict__ inst, const uint32_t* __restrict__ ids, uint32_t n,
                                 const float* __restrict__ mesh_bounds /* 6 per mesh */, float* __restrict__ leaf_lo,
                                 float* __restrict__ leaf_hi, float* __restrict__ bounds) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const PtInstance& in = inst[ids[i]];
    const float* mb = mesh_bounds + 6 * in.mesh_index;
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (int c = 0; c < 8; ++c) {
        double p[3] = {(c & 1) ? mb[3] : mb[0], (c & 2) ? mb[4] : mb[1], (c & 4) ? mb[5] : mb[2]};
        for (int r = 0; r < 3; ++r) {
            double w = (double)in.transform[4 * r] * p[0] + (double)in.transform[4 * r + 1] * p[1] +
                       (double)in.transform[4 * r + 2] * p[2] + (double)in.transform[4 * r + 3];
            lo[r] = fminf(lo[r], __double2float_rd(w));
            hi[r] = fmaxf(hi[r], __double2float_ru(w));
        }
    }
    for (int a = 0; a < 3; ++a) {
        float pad = 1e-5f * fmaxf(fmaxf(fabsf(lo[a]), fabsf(hi[a])), hi[a] - lo[a]) + 1e-30f;
        lo[a] -= pad; hi[a] += pad;
        leaf_lo[3 * (size_t)i + a] = lo[a];
        leaf_hi[3 * (size_t)i + a] = hi[a];
        atomic_min_f(bounds + a, lo[a]);
        atomic_max_f(bounds + 3 + a, hi[a]);
    }
}

// Exact world boxes: every vertex of the instance's mesh through the transform.  The box of a rotated box is up
// to sqrt(3) larger per axis than the box of the rotated mesh, and every false top-level hit costs a ray an
// instance entry (ray transform, new frame, BLAS root test), so for rotated instances this is where
// traversal time goes (C4: DESIGN.md).  grid = (chunks, instances); leaf_lo/hi pre-set to +-3e38.
__global__ void k_instance_boxes_init(uint32_t n, float* __restrict__ leaf_lo, float* __restrict__ leaf_hi) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 3 * n) { leaf_lo[i] = 3.0e38f; leaf_hi[i] = -3.0e38f; }
}
__global__ void __launch_bounds__(256) k_instance_boxes_exact(const PtInstance* __restrict__ inst, const uint32_t* __restrict__ ids,
                                                               const Bvh8Tri* __restrict__ tris, const uint2* __restrict__ mesh_range,
                                                               float* __restrict__ leaf_lo, float* __restrict__ leaf_hi) {
    const uint32_t i = blockIdx.y;
    const PtInstance& in = inst[ids[i]];
    const uint2 rg = mesh_range[in.mesh_index];          // (first triangle, count)
    float m[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) m[k] = in.transform[k];
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < rg.y; t += gridDim.x * blockDim.x) {
        const float4* q = reinterpret_cast<const float4*>(tris + rg.x + t);
#pragma unroll
        for (int v = 0; v < 3; ++v) {
            const float4 p = __ldg(q + v);
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const float w = m[4 * r] * p.x + m[4 * r + 1] * p.y + m[4 * r + 2] * p.z + m[4 * r + 3];
                lo[r] = fminf(lo[r], w); hi[r] = fmaxf(hi[r], w);
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[r] = fminf(lo[r], __shfl_xor_sync(0xffffffffu, lo[r], o));
            hi[r] = fmaxf(hi[r], __shfl_xor_sync(0xffffffffu, hi[r], o));
        }
    }
    if ((threadIdx.x & 31u) == 0u && lo[0] <= hi[0]) {
#pragma unroll
        for (int r = 0; r < 3; ++r) { atomic_min_f(leaf_lo + 3 * (size_t)i + r, lo[r]); atomic_max_f(leaf_hi + 3 * (size_t)i + r, hi[r]); }
    }
}
// pad (fp32 transform here, fp32 inverse transform in traversal: a few ulp apart; 1e-5 relative is ~100x that)
__global__ void k_instance_boxes_finish(uint32_t n, float* __restrict__ leaf_lo, float* __restrict__ leaf_hi, float* __restrict__ bounds) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int a = 0; a < 3; ++a) {
        float lo = leaf_lo[3 * (size_t)i + a], hi = leaf_hi[3 * (size_t)i + a];
        const float pad = 1e-5f * fmaxf(fmaxf(fabsf(lo), fabsf(hi)), hi - lo) + 1e-30f;
        lo -= pad; hi += pad;
        leaf_lo[3 * (size_t)i + a] = lo; leaf_hi[3 * (size_t)i + a] = hi;
        atomic_min_f(bounds + a, lo);
        atomic_max_f(bounds + 3 + a, hi);
    }
}

// ------------------------------------------------------------------ Morton codes
__device__ __forceinline__ uint64_t expand21(uint32_t v) {
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void k_morton(const float* __restrict__ leaf_lo, const float* __restrict__ leaf_hi, uint32_t n,
                         const float* __restrict__ bounds, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t q[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float lo = bounds[a], ext = bounds[3 + a] - bounds[a];
        float c = 0.5f * (leaf_lo[3 * (size_t)i + a] + leaf_hi[3 * (size_t)i + a]);
        float u = ext > 0.0f ? (c - lo) / ext : 0.0f;
        u = fminf(fmaxf(u, 0.0f), 1.0f);
        q[a] = min(2097151u, (uint32_t)(u * 2097152.0f));
    }
    keys[i] = (expand21(q[0]) << 2) | (expand21(q[1]) << 1) | expand21(q[2]);
    vals[i] = i;
}

// ------------------------------------------------------------------ Karras 2012 radix tree
__device__ __forceinline__ int delta(const uint64_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(a ^ b));
}

// Internal node i in [0, n-1); leaves are referenced as (index | 0x80000000).
__global__ void k_karras(const uint64_t* __restrict__ keys, int n, Bvh2Node* __restrict__ nodes,
                         uint32_t* __restrict__ parent /* n-1 internal + n leaves */, uint2* __restrict__ range) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int first = min(i, j), last = max(i, j);
    uint32_t left = (first == gamma) ? ((uint32_t)gamma | 0x80000000u) : (uint32_t)gamma;
    uint32_t right = (last == gamma + 1) ? ((uint32_t)(gamma + 1) | 0x80000000u) : (uint32_t)(gamma + 1);
    nodes[i].left = left;
    nodes[i].right = right;
    range[i] = make_uint2((uint32_t)first, (uint32_t)last);
    // parent links: internal j -> slot j, leaf j -> slot (n-1) + j
    parent[(left & 0x80000000u) ? (n - 1) + (left & 0x7fffffffu) : left] = (uint32_t)i;
    parent[(right & 0x80000000u) ? (n - 1) + (right & 0x7fffffffu) : right] = (uint32_t)i;
    if (i == 0) parent[0] = 0xFFFFFFFFu;
}

// Bottom-up refit: one thread per (sorted) leaf; the second thread to reach a node continues.
__global__ void k_refit(int n, const uint32_t* __restrict__ order, const float* __restrict__ leaf_lo,
                        const float* __restrict__ leaf_hi, Bvh2Node* __restrict__ nodes,
                        const uint32_t* __restrict__ parent, uint32_t* __restrict__ flags) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t cur = parent[(n - 1) + i];
    while (cur != 0xFFFFFFFFu) {
        __threadfence();
        if (atomicAdd(&flags[cur], 1u) == 0u) return;  // first arrival: sibling not ready
        __threadfence();
        float lo[3], hi[3];
        const uint32_t ch[2] = {nodes[cur].left, nodes[cur].right};
#pragma unroll
        for (int a = 0; a < 3; ++a) { lo[a] = 3.0e38f; hi[a] = -3.0e38f; }
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            if (ch[c] & 0x80000000u) {
                size_t src = order[ch[c] & 0x7fffffffu];
#pragma unroll
                for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], leaf_lo[3 * src + a]); hi[a] = fmaxf(hi[a], leaf_hi[3 * src + a]); }
            } else {
                // volatile: written by another thread of this launch
                const volatile Bvh2Node* cn = nodes + ch[c];
#pragma unroll
                for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], cn->lo[a]); hi[a] = fmaxf(hi[a], cn->hi[a]); }
            }
        }
#pragma unroll
        for (int a = 0; a < 3; ++a) { nodes[cur].lo[a] = lo[a]; nodes[cur].hi[a] = hi[a]; }
        cur = parent[cur];
    }
}

// Verification / repair of the refit: every internal node's box must be the union of its children's.
// One Jacobi pass recomputes each node from its children and counts the nodes that changed; the host
// repeats it until nothing changes.  A correct refit needs exactly one (verifying) pass; the pass
// exists because one GPU run of this round produced a tree whose binary-walker results differed from
// the oracle on 9 of 200 000 rays (never reproduced in 16 later builds) — the builder must not depend
// on luck, and the check costs ~10 us.
__global__ void k_refit_verify(int n, const uint32_t* __restrict__ order, const float* __restrict__ leaf_lo,
                               const float* __restrict__ leaf_hi, Bvh2Node* __restrict__ nodes, uint32_t* __restrict__ changed) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    const uint32_t ch[2] = {nodes[i].left, nodes[i].right};
    for (int c = 0; c < 2; ++c) {
        if (ch[c] & 0x80000000u) {
            size_t src = order[ch[c] & 0x7fffffffu];
            for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], leaf_lo[3 * src + a]); hi[a] = fmaxf(hi[a], leaf_hi[3 * src + a]); }
        } else {
            const volatile Bvh2Node* cn = nodes + ch[c];
            for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], cn->lo[a]); hi[a] = fmaxf(hi[a], cn->hi[a]); }
        }
    }
    bool diff = false;
    for (int a = 0; a < 3; ++a) diff = diff || !(nodes[i].lo[a] == lo[a]) || !(nodes[i].hi[a] == hi[a]);
    if (diff) {
        for (int a = 0; a < 3; ++a) { nodes[i].lo[a] = lo[a]; nodes[i].hi[a] = hi[a]; }
        atomicAdd(changed, 1u);
    }
}

// n == 1: a single internal node with one leaf child.
__global__ void k_single_leaf(const float* __restrict__ leaf_lo, const float* __restrict__ leaf_hi, Bvh2Node* nodes, uint2* range) {
    for (int a = 0; a < 3; ++a) { nodes[0].lo[a] = leaf_lo[a]; nodes[0].hi[a] = leaf_hi[a]; }
    nodes[0].left = 0x80000000u;
    nodes[0].right = 0xFFFFFFFFu;
    range[0] = make_uint2(0u, 0u);
}

// Morton-ordered triangle array for the BVH2 walker
__global__ void k_reorder_tris(const Bvh8Tri* __restrict__ in, const uint32_t* __restrict__ order, uint32_t n, Bvh8Tri* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[order[i]];
}

__global__ void k_gather_u32(const uint32_t* __restrict__ in, const uint32_t* __restrict__ order, uint32_t n, uint32_t* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[order[i]];
}

// ------------------------------------------------------------------ PLOC binary hierarchy
// Cluster i = (ref, box); ref = leaf (sorted index | 0x80000000) or internal node index.  Node
// indices are handed out downwards from n-2 so that the last merge (the root) is node 0, which is
// what the collapse and the BVH2 walker expect.  range[node] = (0, leaves-1).
constexpr int kPlocRadius = 16;

__global__ void k_ploc_init(uint32_t n, const uint32_t* __restrict__ order, const float* __restrict__ leaf_lo,
                            const float* __restrict__ leaf_hi, uint32_t* __restrict__ cid, float* __restrict__ clo, float* __restrict__ chi) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t src = order[i];
    cid[i] = i | 0x80000000u;
    for (int a = 0; a < 3; ++a) { clo[3 * (size_t)i + a] = leaf_lo[3 * src + a]; chi[3 * (size_t)i + a] = leaf_hi[3 * src + a]; }
}

__global__ void k_ploc_nearest(uint32_t c, const float* __restrict__ clo, const float* __restrict__ chi, uint32_t* __restrict__ nn) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int)c) return;
    const float lo0 = clo[3 * (size_t)i], lo1 = clo[3 * (size_t)i + 1], lo2 = clo[3 * (size_t)i + 2];
    const float hi0 = chi[3 * (size_t)i], hi1 = chi[3 * (size_t)i + 1], hi2 = chi[3 * (size_t)i + 2];
    float best = 3.0e38f; int bj = -1;
    const int j0 = max(0, i - kPlocRadius), j1 = min((int)c - 1, i + kPlocRadius);
    for (int j = j0; j <= j1; ++j) {
        if (j == i) continue;
        const float dx = fmaxf(hi0, chi[3 * (size_t)j]) - fminf(lo0, clo[3 * (size_t)j]);
        const float dy = fmaxf(hi1, chi[3 * (size_t)j + 1]) - fminf(lo1, clo[3 * (size_t)j + 1]);
        const float dz = fmaxf(hi2, chi[3 * (size_t)j + 2]) - fminf(lo2, clo[3 * (size_t)j + 2]);
        const float ar = dx * dy + dy * dz + dz * dx;     // symmetric in (i, j)
        // ties (duplicated geometry): nearer index first, then even i looks up and odd i looks down, so
        // that runs of identical boxes still pair up (a mutual pair always exists: follow the preferred
        // direction to the end of the run)
        bool better = ar < best;
        if (ar == best && bj >= 0) {
            const int dj = abs(j - i), db = abs(bj - i);
            better = dj < db || (dj == db && ((i & 1) == 0) == (j > i));
        }
        if (better) { best = ar; bj = j; }
    }
    nn[i] = (uint32_t)bj;
}

__global__ void k_ploc_merge(uint32_t c, uint32_t n, const uint32_t* __restrict__ cid, const float* __restrict__ clo,
                             const float* __restrict__ chi, const uint32_t* __restrict__ nn, Bvh2Node* __restrict__ nodes,
                             uint2* __restrict__ range, uint32_t* __restrict__ counter, uint32_t* __restrict__ cid_out,
                             float* __restrict__ lo_out, float* __restrict__ hi_out, uint32_t* __restrict__ keep) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c) return;
    const uint32_t j = nn[i];
    float lo[3], hi[3];
    for (int a = 0; a < 3; ++a) { lo[a] = clo[3 * (size_t)i + a]; hi[a] = chi[3 * (size_t)i + a]; }
    uint32_t ref = cid[i], k = 1u;
    if (nn[j] == i) {
        if (i < j) {
            const uint32_t idx = (n - 2u) - atomicAdd(counter, 1u);
            const uint32_t l = cid[i], r = cid[j];
            for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], clo[3 * (size_t)j + a]); hi[a] = fmaxf(hi[a], chi[3 * (size_t)j + a]); }
            Bvh2Node nd;
            for (int a = 0; a < 3; ++a) { nd.lo[a] = lo[a]; nd.hi[a] = hi[a]; }
            nd.left = l; nd.right = r;
            nodes[idx] = nd;
            const uint32_t cl = (l & 0x80000000u) ? 1u : range[l].y + 1u, cr = (r & 0x80000000u) ? 1u : range[r].y + 1u;
            range[idx] = make_uint2(0u, cl + cr - 1u);
            ref = idx;
        } else {
            k = 0u;   // absorbed by cluster j
        }
    }
    keep[i] = k;
    cid_out[i] = ref;
    for (int a = 0; a < 3; ++a) { lo_out[3 * (size_t)i + a] = lo[a]; hi_out[3 * (size_t)i + a] = hi[a]; }
}

__global__ void k_ploc_compact(uint32_t c, const uint32_t* __restrict__ keep, const uint32_t* __restrict__ pos,
                               const uint32_t* __restrict__ cid_in, const float* __restrict__ lo_in, const float* __restrict__ hi_in,
                               uint32_t* __restrict__ cid_out, float* __restrict__ lo_out, float* __restrict__ hi_out,
                               uint32_t* __restrict__ new_count) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c) return;
    if (keep[i]) {
        const size_t o = pos[i];
        cid_out[o] = cid_in[i];
        for (int a = 0; a < 3; ++a) { lo_out[3 * o + a] = lo_in[3 * (size_t)i + a]; hi_out[3 * o + a] = hi_in[3 * (size_t)i + a]; }
    }
    if (i == c - 1) *new_count = pos[i] + keep[i];
}

// ------------------------------------------------------------------ BVH2 -> BVH8 collapse
struct CollapseArgs {
    const Bvh2Node* nodes2;     // this tree's internal nodes (local indices)
    const uint2* range;         // leaf range per internal node
    const uint32_t* order;      // sorted leaf -> source leaf
    const float* leaf_lo;       // source leaf boxes
    const float* leaf_hi;
    uint32_t max_leaf;          // <= kLeafStride: 2 for a BLAS (triangles per leaf slot), 1 for the TLAS
    const uint8_t* dp;          // decisions of k_collapse_dp (8 per binary node), or nullptr: greedy surface-area expansion
    // outputs
    Bvh8Node* nodes8;           // global node array
    uint32_t* node_counter;     // global
    uint32_t* tri_counter;      // global (BLAS: tris8 entries; TLAS: leaf-instance entries)
    const Bvh8Tri* tris_sorted; // BLAS: Morton-ordered triangles of this mesh
    Bvh8Tri* tris8;             // BLAS leaf triangle output (global)
    const uint32_t* leaf_ids;   // TLAS: source leaf -> instance id
    uint32_t* leaf_out;         // TLAS leaf output
    // work queue
    uint32_t* wq2;
    uint32_t* wq8;
    uint32_t* tail;
    uint32_t capacity;
    uint32_t* overflow;
};

struct CChild { uint32_t ref, count; float lo[3], hi[3]; uint32_t first; bool leaf; };   // first: sorted index of a packed run, else ~0

__device__ __forceinline__ CChild load_child(const CollapseArgs& a, uint32_t ref) {
    CChild c;
    c.ref = ref;
    c.first = 0xFFFFFFFFu;
    if (ref & 0x80000000u) {
        uint32_t j = ref & 0x7fffffffu;
        size_t src = a.order[j];
        for (int k = 0; k < 3; ++k) { c.lo[k] = a.leaf_lo[3 * src + k]; c.hi[k] = a.leaf_hi[3 * src + k]; }
        c.count = 1;
    } else {
        const Bvh2Node& n = a.nodes2[ref];
        for (int k = 0; k < 3; ++k) { c.lo[k] = n.lo[k]; c.hi[k] = n.hi[k]; }
        uint2 r = a.range[ref];
        c.count = r.y - r.x + 1;
    }
    c.leaf = (ref & 0x80000000u) || c.count <= a.max_leaf;   // greedy rule; the optimal collapse overrides it
    return c;
}

__device__ __forceinline__ float half_area(const CChild& c) {
    float dx = c.hi[0] - c.lo[0], dy = c.hi[1] - c.lo[1], dz = c.hi[2] - c.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_collapse(CollapseArgs a, uint32_t begin, uint32_t end) {
    uint32_t w = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= end) return;
    const uint32_t b2 = a.wq2[w], b8 = a.wq8[w];
    CChild ch[8];
    int cnt = 0;
    const bool optimal = a.dp != nullptr && a.nodes2[b2].right != 0xFFFFFFFFu;
    if (optimal) {
        uint32_t refs[8]; bool is_leaf[8];
        cnt = dp_cover(a.dp, a.nodes2, b2, refs, is_leaf);      // bvh_sah.cuh
        for (int c = 0; c < cnt; ++c) { ch[c] = load_child(a, refs[c]); ch[c].leaf = is_leaf[c]; }
    } else {
        const Bvh2Node& n = a.nodes2[b2];
        ch[cnt++] = load_child(a, n.left);
        if (n.right != 0xFFFFFFFFu) ch[cnt++] = load_child(a, n.right);
    }
    // greedy surface-area expansion up to 8 children
    while (!optimal && cnt < 8) {
        int best = -1; float best_a = -1.0f;
        for (int c = 0; c < cnt; ++c) {
            if (ch[c].leaf) continue;
            float ar = half_area(ch[c]);
            if (ar > best_a) { best_a = ar; best = c; }
        }
        if (best < 0) break;
        const Bvh2Node& n = a.nodes2[ch[best].ref];
        const uint32_t l = n.left, r = n.right;
        ch[best] = load_child(a, l);
        ch[cnt++] = load_child(a, r);
    }
    // node frame
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (int c = 0; c < cnt; ++c)
        for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], ch[c].lo[k]); hi[k] = fmaxf(hi[k], ch[c].hi[k]); }
    float p[3], scale[3]; int ebits[3];
    for (int k = 0; k < 3; ++k) {
        float ext = hi[k] - lo[k];
        float mag = fmaxf(fabsf(lo[k]), fabsf(hi[k]));
        ext = fmaxf(fmaxf(ext, mag * 1.2e-4f), 1e-30f) * 1.01f;  // grid step >= 4 ulp(mag)
        int e;
        frexpf(ext / 250.0f, &e);           // 2^e >= ext/250
        e = max(-125, min(126, e));
        ebits[k] = e + 127;
        scale[k] = __int_as_float(ebits[k] << 23);
        p[k] = lo[k] - 0.5f * scale[k];     // origin strictly below every child lo (half a grid step)
    }
    // slot assignment: maximise sum of dot(child centroid - node centroid, octant direction of slot)
    int slot_of[8]; bool slot_used[8]; bool assigned[8];
    for (int s = 0; s < 8; ++s) { slot_used[s] = false; assigned[s] = false; slot_of[s] = -1; }
    float nc[3] = {0.5f * (lo[0] + hi[0]), 0.5f * (lo[1] + hi[1]), 0.5f * (lo[2] + hi[2])};
    for (int it = 0; it < cnt; ++it) {
        float best = -3.0e38f; int bc = -1, bs = -1;
        for (int c = 0; c < cnt; ++c) {
            if (assigned[c]) continue;
            float dx = 0.5f * (ch[c].lo[0] + ch[c].hi[0]) - nc[0];
            float dy = 0.5f * (ch[c].lo[1] + ch[c].hi[1]) - nc[1];
            float dz = 0.5f * (ch[c].lo[2] + ch[c].hi[2]) - nc[2];
            for (int s = 0; s < 8; ++s) {
                if (slot_used[s]) continue;
                float cost = ((s & 1) ? dx : -dx) + ((s & 2) ? dy : -dy) + ((s & 4) ? dz : -dz);
                if (cost > best) { best = cost; bc = c; bs = s; }
            }
        }
        assigned[bc] = true; slot_used[bs] = true; slot_of[bc] = bs;
    }
    int child_in_slot[8];
    for (int s = 0; s < 8; ++s) child_in_slot[s] = -1;
    for (int c = 0; c < cnt; ++c) child_in_slot[slot_of[c]] = c;

    // classify, allocate
    uint32_t n_inner = 0, n_leaf_items = 0;
    for (int s = 0; s < 8; ++s) {
        int c = child_in_slot[s];
        if (c < 0) continue;
        if (ch[c].leaf) n_leaf_items += ch[c].count; else n_inner++;
    }
    uint32_t child_base = n_inner ? atomicAdd(a.node_counter, n_inner) : 0u;
    uint32_t tri_base = n_leaf_items ? atomicAdd(a.tri_counter, n_leaf_items) : 0u;
    uint32_t q_base = n_inner ? atomicAdd(a.tail, n_inner) : 0u;
    if (n_inner && q_base + n_inner > a.capacity) { atomicExch(a.overflow, 1u); return; }

    Bvh8Node out;
    out.px = p[0]; out.py = p[1]; out.pz = p[2];
    out.ex = (uint8_t)ebits[0]; out.ey = (uint8_t)ebits[1]; out.ez = (uint8_t)ebits[2];
    out.child_base = child_base; out.tri_base = tri_base;
    uint32_t imask = 0, inner_rank = 0, leaf_off = 0, valid = 0;
    const float delta = 1.0f / 1024.0f;   // quantisation slack (fraction of a grid step)
    for (int s = 0; s < 8; ++s) {
        int c = child_in_slot[s];
        if (c < 0) {
            out.qlo_x[s] = out.qlo_y[s] = out.qlo_z[s] = 255;   // inverted box: never hit
            out.qhi_x[s] = out.qhi_y[s] = out.qhi_z[s] = 0;
            continue;
        }
        uint8_t ql[3], qh[3];
        for (int k = 0; k < 3; ++k) {
            float fl = floorf((ch[c].lo[k] - p[k]) / scale[k] - delta);
            float fh = ceilf((ch[c].hi[k] - p[k]) / scale[k] + delta);
            ql[k] = (uint8_t)fminf(fmaxf(fl, 0.0f), 255.0f);
            qh[k] = (uint8_t)fminf(fmaxf(fh, 0.0f), 255.0f);
        }
        out.qlo_x[s] = ql[0]; out.qlo_y[s] = ql[1]; out.qlo_z[s] = ql[2];
        out.qhi_x[s] = qh[0]; out.qhi_y[s] = qh[1]; out.qhi_z[s] = qh[2];
        if (!ch[c].leaf) {
            imask |= 1u << s;
            valid |= 1u << (24u + (uint32_t)s);
            a.wq2[q_base + inner_rank] = ch[c].ref;
            a.wq8[q_base + inner_rank] = child_base + inner_rank;
            inner_rank++;
        } else {
            const uint32_t k = ch[c].count;
            // unary count in the slot's two bits of `valid`; the items themselves are dense in slot order (leaf_off)
            valid |= ((k == 1) ? 1u : 3u) << (kLeafStride * (uint32_t)s);
            // the (at most max_leaf) leaves under this child, left to right; any binary topology
            uint32_t st[4]; int sp = 0; uint32_t t = 0;
            if (ch[c].first != 0xFFFFFFFFu) {
                for (; t < k; ++t) {      // packed run of the Morton order
                    const uint32_t j = ch[c].first + t;
                    if (a.tris8) a.tris8[tri_base + leaf_off + t] = a.tris_sorted[j];
                    else a.leaf_out[tri_base + leaf_off + t] = a.leaf_ids[a.order[j]];
                }
            } else st[sp++] = ch[c].ref;
            while (sp) {
                const uint32_t ref = st[--sp];
                if (ref & 0x80000000u) {
                    const uint32_t j = ref & 0x7fffffffu;
                    if (a.tris8) a.tris8[tri_base + leaf_off + t] = a.tris_sorted[j];
                    else a.leaf_out[tri_base + leaf_off + t] = a.leaf_ids[a.order[j]];
                    ++t;
                } else {
                    const Bvh2Node& n = a.nodes2[ref];
                    if (n.right != 0xFFFFFFFFu && sp < 4) st[sp++] = n.right;
                    if (sp < 4) st[sp++] = n.left;
                }
            }
            leaf_off += k;
        }
    }
    out.imask = (uint8_t)imask;
    out.valid = valid; out.spare = 0u;
    a.nodes8[b8] = out;
}

}  // namespace

// ------------------------------------------------------------------ host orchestration
void SceneBvh::release() {
    cudaFree(nodes8); cudaFree(tris8); cudaFree(tlas_leaf_instance); cudaFree(nodes2); cudaFree(tris2);
    cudaFree(tlas2); cudaFree(tlas2_leaf_instance); cudaFree(mesh_bounds); cudaFree(mesh_tri_range); mesh_tri_range = nullptr;
    nodes8 = nullptr; tris8 = nullptr; tlas_leaf_instance = nullptr; nodes2 = nullptr; tris2 = nullptr;
    tlas2 = nullptr; tlas2_leaf_instance = nullptr; mesh_bounds = nullptr;
}

namespace {

struct Scratch {
    Bvh8Tri* tmp_tris = nullptr;
    float *leaf_lo = nullptr, *leaf_hi = nullptr, *bounds = nullptr;
    uint64_t *keys_in = nullptr, *keys_out = nullptr;
    uint32_t *vals_in = nullptr, *vals_out = nullptr, *parent = nullptr, *flags = nullptr, *wq2 = nullptr, *wq8 = nullptr;
    uint32_t* counters = nullptr;  // [0] node counter, [1] tri counter, [2] tlas leaf counter, [3] queue tail, [4] overflow
    uint2* range = nullptr;
    uint32_t* prim_prefix = nullptr;
    void* cub_temp = nullptr;
    size_t cub_bytes = 0;
    // PLOC: double-buffered clusters, nearest-neighbour / keep / scan arrays
    uint32_t *cid[2] = {nullptr, nullptr}, *nn = nullptr, *keep = nullptr, *pos = nullptr;
    float *clo[2] = {nullptr, nullptr}, *chi[2] = {nullptr, nullptr};
    bool lbvh = false;
    // opt-in stages of bvh_sah.cuh
    bool sah = false;             // binary hierarchy by top-down binned SAH instead of the Morton radix tree / PLOC
    bool optimal = false;         // cost-optimal collapse (k_collapse_dp) instead of the greedy expansion
    SahScratch sahs{};            // allocated when sah
    float* dp_F = nullptr;        // 8 per binary node, allocated when optimal
    uint8_t* dp_D = nullptr;
    uint32_t* cur_parent = nullptr;   // parent links of the tree just built (Karras layout), for k_collapse_dp
    uint32_t sah_levels = 0;
    uint32_t capacity = 0;        // entries of wq2 / wq8
    uint32_t max_leaf = kLeafStride;   // triangles per BLAS leaf slot = the format's limit (two bits of Bvh8Node::valid per slot);
                                  // 1 / 2 / 3 measured on C2 in round 1: 1.438 / 1.406 / 1.429 ms, C4: 19.4 / 19.5 / 20.3 ms, and 3 again on
                                  // the round-2 kernel: no better (profiles/r2t_variants.txt) -- a triangle test costs ~0.7 node tests here
    uint32_t refit_repairs = 0;   // nodes the verification pass had to correct (expected 0)
    void release() {
        cudaFree(cid[0]); cudaFree(cid[1]); cudaFree(nn); cudaFree(keep); cudaFree(pos);
        cudaFree(clo[0]); cudaFree(clo[1]); cudaFree(chi[0]); cudaFree(chi[1]);
        cudaFree(tmp_tris); cudaFree(leaf_lo); cudaFree(leaf_hi); cudaFree(bounds); cudaFree(keys_in); cudaFree(keys_out);
        cudaFree(vals_in); cudaFree(vals_out); cudaFree(parent); cudaFree(flags); cudaFree(wq2); cudaFree(wq8);
        cudaFree(counters); cudaFree(range); cudaFree(prim_prefix); cudaFree(cub_temp);
        cudaFree(sahs.idx[0]); cudaFree(sahs.idx[1]); cudaFree(sahs.seg[0]); cudaFree(sahs.seg[1]); cudaFree(sahs.segs[0]); cudaFree(sahs.segs[1]);
        cudaFree(sahs.counters); cudaFree(sahs.cb); cudaFree(sahs.bins); cudaFree(sahs.flag); cudaFree(sahs.scan);
        cudaFree(dp_F); cudaFree(dp_D);
    }
};

// sort + hierarchy + refit for n leaves whose boxes are in s.leaf_lo/hi and whose bounds are in s.bounds
int build_tree(Scratch& s, uint32_t n, Bvh2Node* nodes2_out, cudaStream_t st) {
    k_morton<<<blocks_for(n), kThreads, 0, st>>>(s.leaf_lo, s.leaf_hi, n, s.bounds, s.keys_in, s.vals_in);
    PT_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(s.cub_temp, s.cub_bytes, s.keys_in, s.keys_out, s.vals_in, s.vals_out, (int)n, 0, 63, st));
    if (n == 1) {
        // vals_out[0] == 0
        k_single_leaf<<<1, 1, 0, st>>>(s.leaf_lo, s.leaf_hi, nodes2_out, s.range);
    } else if (!s.lbvh && !s.sah) {
        uint32_t* counter = s.counters + 6;     // [6] nodes created, [7] clusters left
        PT_CUDA_CHECK(cudaMemsetAsync(counter, 0, 8, st));
        k_ploc_init<<<blocks_for(n), kThreads, 0, st>>>(n, s.vals_out, s.leaf_lo, s.leaf_hi, s.cid[0], s.clo[0], s.chi[0]);
        uint32_t c = n;
        int cur = 0;
        for (int it = 0; c > 1; ++it) {
            if (it > (1 << 16)) { set_error("PLOC did not converge"); return PT_ERR_CUDA; }
            k_ploc_nearest<<<blocks_for(c), kThreads, 0, st>>>(c, s.clo[cur], s.chi[cur], s.nn);
            // merge writes (ref, box) in place of cluster i into the other buffer; compact brings them back
            k_ploc_merge<<<blocks_for(c), kThreads, 0, st>>>(c, n, s.cid[cur], s.clo[cur], s.chi[cur], s.nn, nodes2_out, s.range, counter,
                                                             s.cid[cur ^ 1], s.clo[cur ^ 1], s.chi[cur ^ 1], s.keep);
            PT_CUDA_CHECK(cub::DeviceScan::ExclusiveSum(s.cub_temp, s.cub_bytes, s.keep, s.pos, (int)c, st));
            k_ploc_compact<<<blocks_for(c), kThreads, 0, st>>>(c, s.keep, s.pos, s.cid[cur ^ 1], s.clo[cur ^ 1], s.chi[cur ^ 1], s.cid[cur],
                                                               s.clo[cur], s.chi[cur], counter + 1);
            uint32_t left = 0;
            PT_CUDA_CHECK(cudaMemcpyAsync(&left, counter + 1, 4, cudaMemcpyDeviceToHost, st));
            PT_CUDA_CHECK(cudaStreamSynchronize(st));
            if (left == 0 || left >= c) { set_error("PLOC made no progress"); return PT_ERR_CUDA; }
            c = left;
        }
    } else {
        PT_CUDA_CHECK(cudaMemsetAsync(s.flags, 0, sizeof(uint32_t) * n, st));
        if (s.sah) {
            // the Morton sort above is not needed by this stage (it orders the leaves itself); vals_out receives its order
            const uint32_t* order = nullptr;
            const int rc = build_tree_sah(s.sahs, n, s.leaf_lo, s.leaf_hi, nodes2_out, s.range, s.parent, &order, &s.sah_levels, st);
            if (rc == 2) { set_error("binned-SAH build exceeded its level budget"); return PT_ERR_CUDA; }
            if (rc != 0) { set_error(std::string("binned-SAH build: ") + cudaGetErrorString(cudaGetLastError())); return PT_ERR_CUDA; }
            PT_CUDA_CHECK(cudaMemcpyAsync(s.vals_out, order, 4 * (size_t)n, cudaMemcpyDeviceToDevice, st));
        } else {
            k_karras<<<blocks_for(n - 1), kThreads, 0, st>>>(s.keys_out, (int)n, nodes2_out, s.parent, s.range);
        }
        k_refit<<<blocks_for(n), kThreads, 0, st>>>((int)n, s.vals_out, s.leaf_lo, s.leaf_hi, nodes2_out, s.parent, s.flags);
        uint32_t* changed = s.counters + 8;
        for (int pass = 0;; ++pass) {
            PT_CUDA_CHECK(cudaMemsetAsync(changed, 0, 4, st));
            k_refit_verify<<<blocks_for(n - 1), kThreads, 0, st>>>((int)n, s.vals_out, s.leaf_lo, s.leaf_hi, nodes2_out, changed);
            uint32_t h = 0;
            PT_CUDA_CHECK(cudaMemcpyAsync(&h, changed, 4, cudaMemcpyDeviceToHost, st));
            PT_CUDA_CHECK(cudaStreamSynchronize(st));
            if (h == 0) break;
            s.refit_repairs += h;
            if (pass > 256) { set_error("BVH refit did not converge"); return PT_ERR_CUDA; }
        }
    }
    PT_CUDA_CHECK(cudaGetLastError());
    return PT_OK;
}

// level-synchronous collapse of the tree just built; returns the BVH8 root index
int collapse_tree(Scratch& s, CollapseArgs a, uint32_t n, uint32_t* root_out, cudaStream_t st) {
    a.dp = nullptr;
    // The cost table needs parent links, which the Morton radix tree and the SAH stage record (PLOC does not).
    if (s.optimal && n >= 2 && (s.lbvh || s.sah)) {
        DpArgs d;
        d.n = n; d.nodes = a.nodes2; d.range = a.range; d.parent = s.parent; d.order = a.order; d.leaf_lo = a.leaf_lo; d.leaf_hi = a.leaf_hi;
        d.visit = s.flags; d.F = s.dp_F; d.D = s.dp_D; d.max_leaf = a.max_leaf;
        // one node test costs what ~1.4 triangle tests cost in K-trace (profiles/r1_experiments.md: leaf-size sweeps)
        d.c_node = 1.0f; d.c_tri = 0.7f;
        PT_CUDA_CHECK(cudaMemsetAsync(s.flags, 0, sizeof(uint32_t) * n, st));
        k_collapse_dp<<<blocks_for(n), kThreads, 0, st>>>(d);
        PT_CUDA_CHECK(cudaGetLastError());
        a.dp = s.dp_D;
    }
    // allocate the root node
    uint32_t root;
    PT_CUDA_CHECK(cudaMemcpyAsync(&root, a.node_counter, 4, cudaMemcpyDeviceToHost, st));
    PT_CUDA_CHECK(cudaStreamSynchronize(st));
    uint32_t init[2] = {root + 1, 0};
    PT_CUDA_CHECK(cudaMemcpyAsync(a.node_counter, &init[0], 4, cudaMemcpyHostToDevice, st));
    uint32_t zero = 0, one = 1;
    PT_CUDA_CHECK(cudaMemcpyAsync(s.wq2, &zero, 4, cudaMemcpyHostToDevice, st));
    PT_CUDA_CHECK(cudaMemcpyAsync(s.wq8, &root, 4, cudaMemcpyHostToDevice, st));
    PT_CUDA_CHECK(cudaMemcpyAsync(a.tail, &one, 4, cudaMemcpyHostToDevice, st));
    uint32_t begin = 0, end = 1;
    while (begin < end) {
        k_collapse<<<(int)((end - begin + 63) / 64), 64, 0, st>>>(a, begin, end);
        uint32_t tail_ovf[2];
        PT_CUDA_CHECK(cudaMemcpyAsync(tail_ovf, a.tail, 8, cudaMemcpyDeviceToHost, st));
        PT_CUDA_CHECK(cudaStreamSynchronize(st));
        if (tail_ovf[1]) { set_error("BVH collapse work queue overflow"); return PT_ERR_CUDA; }
        begin = end; end = tail_ovf[0];
    }
    (void)n;
    *root_out = root;
    return PT_OK;
}

}  // namespace

namespace {

#define PT_CU(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { set_error(std::string(#expr) + ": " + cudaGetErrorString(_e)); return PT_ERR_CUDA; } } while (0)

// scratch for trees of up to L leaves (the caller releases it, also on failure)
int alloc_scratch(Scratch& s, uint32_t L, size_t max_prims, uint32_t builder_flags, cudaStream_t st) {
    const bool use_ploc = (builder_flags & PT_FLAG_BVH_PLOC) != 0;
    PT_CU(cudaMalloc(&s.tmp_tris, sizeof(Bvh8Tri) * L));
    PT_CU(cudaMalloc(&s.leaf_lo, sizeof(float) * 3 * L)); PT_CU(cudaMalloc(&s.leaf_hi, sizeof(float) * 3 * L));
    PT_CU(cudaMalloc(&s.bounds, sizeof(float) * 8));
    PT_CU(cudaMalloc(&s.keys_in, 8 * (size_t)L)); PT_CU(cudaMalloc(&s.keys_out, 8 * (size_t)L));
    PT_CU(cudaMalloc(&s.vals_in, 4 * (size_t)L)); PT_CU(cudaMalloc(&s.vals_out, 4 * (size_t)L));
    PT_CU(cudaMalloc(&s.parent, 4 * (size_t)(2 * L + 2))); PT_CU(cudaMalloc(&s.flags, 4 * (size_t)L));
    PT_CU(cudaMalloc(&s.wq2, 4 * (size_t)(L + 8))); PT_CU(cudaMalloc(&s.wq8, 4 * (size_t)(L + 8)));
    PT_CU(cudaMalloc(&s.counters, 4 * 16)); PT_CU(cudaMemsetAsync(s.counters, 0, 64, st));
    PT_CU(cudaMalloc(&s.range, sizeof(uint2) * L));
    PT_CU(cudaMalloc(&s.prim_prefix, 4 * (max_prims + 1)));
    PT_CU(cub::DeviceRadixSort::SortPairs(nullptr, s.cub_bytes, s.keys_in, s.keys_out, s.vals_in, s.vals_out, (int)L, 0, 63, st));
    {
        size_t scan_bytes = 0;
        PT_CU(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, s.keep, s.pos, (int)L, st));
        s.cub_bytes = std::max(s.cub_bytes, scan_bytes);
    }
    PT_CU(cudaMalloc(&s.cub_temp, std::max<size_t>(s.cub_bytes, 16)));
    s.lbvh = !use_ploc;
    s.sah = (builder_flags & PT_FLAG_BVH_SAH) != 0;
    s.optimal = (builder_flags & PT_FLAG_BVH_OPTIMAL_COLLAPSE) != 0;
    if (s.sah) {
        const size_t S = (size_t)L / 2 + 1;
        for (int k = 0; k < 2; ++k) {
            PT_CU(cudaMalloc(&s.sahs.idx[k], 4 * (size_t)L)); PT_CU(cudaMalloc(&s.sahs.seg[k], 4 * (size_t)L));
            PT_CU(cudaMalloc(&s.sahs.segs[k], sizeof(SahSeg) * S));
        }
        PT_CU(cudaMalloc(&s.sahs.counters, 8)); PT_CU(cudaMemsetAsync(s.sahs.counters, 0, 8, st));
        PT_CU(cudaMalloc(&s.sahs.cb, sizeof(float) * 6 * S));
        PT_CU(cudaMalloc(&s.sahs.bins, sizeof(SahBin) * 3 * kSahBins * S));
        PT_CU(cudaMalloc(&s.sahs.flag, 4 * (size_t)L)); PT_CU(cudaMalloc(&s.sahs.scan, 4 * (size_t)L));
        s.sahs.scan_temp = s.cub_temp; s.sahs.scan_temp_bytes = s.cub_bytes;
    }
    if (s.optimal) {
        PT_CU(cudaMalloc(&s.dp_F, sizeof(float) * 8 * (size_t)L));
        PT_CU(cudaMalloc(&s.dp_D, 8 * (size_t)L));
    }
    if (!s.lbvh && !s.sah) {
        for (int k = 0; k < 2; ++k) {
            PT_CU(cudaMalloc(&s.cid[k], 4 * (size_t)L)); PT_CU(cudaMalloc(&s.clo[k], 12 * (size_t)L)); PT_CU(cudaMalloc(&s.chi[k], 12 * (size_t)L));
        }
        PT_CU(cudaMalloc(&s.nn, 4 * (size_t)L)); PT_CU(cudaMalloc(&s.keep, 4 * (size_t)L)); PT_CU(cudaMalloc(&s.pos, 4 * (size_t)L));
    }
    return PT_OK;
}

// TLAS over the instances that reference a non-empty mesh (Tlas.init, Tlas.zig:15-182).  Its nodes are
// appended at bvh->tlas_node_base, so a rebuild overwrites the previous TLAS in place.
int build_tlas(Scratch& s, SceneBvh* out, const PtInstance* d_instances, cudaStream_t st) {
    const uint32_t NI = (uint32_t)out->valid_inst.size();
    if (NI == 0) return PT_OK;
    uint32_t* node_counter = s.counters + 0;
    uint32_t* tlas_leaf_counter = s.counters + 2;
    const uint32_t zero = 0;
    PT_CU(cudaMemcpyAsync(node_counter, &out->tlas_node_base, 4, cudaMemcpyHostToDevice, st));
    PT_CU(cudaMemcpyAsync(tlas_leaf_counter, &zero, 4, cudaMemcpyHostToDevice, st));
    PT_CU(cudaMemcpyAsync(s.counters + 5, &zero, 4, cudaMemcpyHostToDevice, st));
    PT_CU(cudaMemcpyAsync(s.vals_in, out->valid_inst.data(), 4 * (size_t)NI, cudaMemcpyHostToDevice, st));
    PT_CU(cudaStreamSynchronize(st));
    uint32_t* ids_unsorted = nullptr;
    PT_CU(cudaMalloc(&ids_unsorted, 4 * (size_t)NI));
    int rc = PT_OK;
    if (cudaMemcpyAsync(ids_unsorted, s.vals_in, 4 * (size_t)NI, cudaMemcpyDeviceToDevice, st) != cudaSuccess) rc = PT_ERR_CUDA;
    if (rc == PT_OK) {
        k_init_bounds<<<1, 32, 0, st>>>(s.bounds);
        if (out->exact_instance_boxes) {
            k_instance_boxes_init<<<blocks_for(3 * NI), kThreads, 0, st>>>(NI, s.leaf_lo, s.leaf_hi);
            const uint32_t chunks = std::max(1u, std::min(64u, (out->max_mesh_tris + 2047u) / 2048u));
            k_instance_boxes_exact<<<dim3(chunks, NI), 256, 0, st>>>(d_instances, ids_unsorted, out->tris2, out->mesh_tri_range, s.leaf_lo, s.leaf_hi);
            k_instance_boxes_finish<<<blocks_for(NI), kThreads, 0, st>>>(NI, s.leaf_lo, s.leaf_hi, s.bounds);
        } else {
            k_instance_boxes<<<blocks_for(NI), kThreads, 0, st>>>(d_instances, ids_unsorted, NI, out->mesh_bounds, s.leaf_lo, s.leaf_hi, s.bounds);
        }
        rc = build_tree(s, NI, out->tlas2, st);
    }
    if (rc == PT_OK) {
        // sorted leaf j -> instance id
        k_gather_u32<<<blocks_for(NI), kThreads, 0, st>>>(ids_unsorted, s.vals_out, NI, out->tlas2_leaf_instance);
        CollapseArgs a{};
        a.nodes2 = out->tlas2; a.range = s.range; a.order = s.vals_out; a.leaf_lo = s.leaf_lo; a.leaf_hi = s.leaf_hi;
        a.max_leaf = 1; a.nodes8 = out->nodes8; a.node_counter = node_counter; a.tri_counter = tlas_leaf_counter;
        a.tris_sorted = nullptr; a.tris8 = nullptr; a.leaf_ids = ids_unsorted; a.leaf_out = out->tlas_leaf_instance;
        a.wq2 = s.wq2; a.wq8 = s.wq8; a.tail = s.counters + 4; a.capacity = s.capacity; a.overflow = s.counters + 5;
        rc = collapse_tree(s, a, NI, &out->tlas_root8, st);
    }
    cudaStreamSynchronize(st);
    cudaFree(ids_unsorted);
    return rc;
}
#undef PT_CU

}  // namespace

int build_scene_bvh(const uint8_t* d_blob, const PrimRec* d_prims, const std::vector<PrimRec>& h_prims,
                    const std::vector<PtMesh>& meshes, const PtInstance* d_instances,
                    const std::vector<PtInstance>& h_instances, uint32_t builder_flags, cudaStream_t st, SceneBvh* out) {
    const size_t M = meshes.size();
    out->tri_off.assign(M, 0); out->tri_count.assign(M, 0); out->blas_root8.assign(M, kInvalidId);
    uint32_t total = 0, max_n = 1;
    size_t max_prims = 1;
    for (size_t m = 0; m < M; ++m) {
        uint32_t n = 0;
        for (uint32_t p = meshes[m].start; p < meshes[m].end; ++p) n += h_prims[p].triangle_count;
        out->tri_off[m] = total; out->tri_count[m] = n; total += n;
        max_n = std::max(max_n, n);
        max_prims = std::max<size_t>(max_prims, meshes[m].end - meshes[m].start);
    }
    std::vector<uint32_t> valid_inst;
    for (size_t i = 0; i < h_instances.size(); ++i)
        if (h_instances[i].mesh_index < M && out->tri_count[h_instances[i].mesh_index] > 0) valid_inst.push_back((uint32_t)i);
    const uint32_t NI = (uint32_t)valid_inst.size();
    const uint32_t L = std::max(max_n, std::max(NI, 1u));
    out->num_tris = total;
    out->tlas2_count = NI;

    cudaEvent_t ev0, ev1;
    PT_CUDA_CHECK(cudaEventCreate(&ev0)); PT_CUDA_CHECK(cudaEventCreate(&ev1));
    PT_CUDA_CHECK(cudaEventRecord(ev0, st));

    Scratch s;
    int rc = PT_OK;
    auto fail = [&](int code) { s.release(); cudaEventDestroy(ev0); cudaEventDestroy(ev1); return code; };
#define PT_TRY(expr) do { rc = (expr); if (rc != PT_OK) return fail(rc); } while (0)
#define PT_CU(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { set_error(std::string(#expr) + ": " + cudaGetErrorString(_e)); return fail(PT_ERR_CUDA); } } while (0)
    PT_TRY(alloc_scratch(s, L, max_prims, builder_flags, st));
    s.capacity = L + 8;

    const size_t cap_nodes = (size_t)total + M + NI + 8;
    PT_CU(cudaMalloc(&out->nodes8, sizeof(Bvh8Node) * cap_nodes));
    PT_CU(cudaMalloc(&out->tris8, sizeof(Bvh8Tri) * std::max(total, 1u)));
    PT_CU(cudaMalloc(&out->tris2, sizeof(Bvh8Tri) * std::max(total, 1u)));
    PT_CU(cudaMalloc(&out->nodes2, sizeof(Bvh2Node) * std::max(total, 1u)));
    PT_CU(cudaMalloc(&out->tlas2, sizeof(Bvh2Node) * std::max(NI, 1u)));
    PT_CU(cudaMalloc(&out->tlas_leaf_instance, 4 * (size_t)std::max(NI, 1u)));
    PT_CU(cudaMalloc(&out->tlas2_leaf_instance, 4 * (size_t)std::max(NI, 1u)));
    PT_CU(cudaMalloc(&out->mesh_bounds, sizeof(float) * 6 * std::max<size_t>(M, 1)));

    uint32_t* node_counter = s.counters + 0;
    uint32_t* tri_counter = s.counters + 1;
    uint32_t* tail = s.counters + 4;      // [4] tail, [5] overflow (read together)
    uint32_t* overflow = s.counters + 5;

    // ---- one BLAS per mesh
    for (size_t m = 0; m < M; ++m) {
        const uint32_t n = out->tri_count[m];
        if (n == 0) continue;
        const uint32_t np = meshes[m].end - meshes[m].start;
        std::vector<uint32_t> prefix(np + 1, 0);
        for (uint32_t k = 0; k < np; ++k) prefix[k + 1] = prefix[k] + h_prims[meshes[m].start + k].triangle_count;
        PT_CU(cudaMemcpyAsync(s.prim_prefix, prefix.data(), 4 * (np + 1), cudaMemcpyHostToDevice, st));
        PT_CU(cudaStreamSynchronize(st));  // prefix is a stack temporary
        k_init_bounds<<<1, 32, 0, st>>>(s.bounds);
        k_gather_tris<<<blocks_for(n), kThreads, 0, st>>>(d_blob, d_prims, meshes[m].start, np, s.prim_prefix, n, s.tmp_tris,
                                                           s.leaf_lo, s.leaf_hi, s.bounds);
        PT_CU(cudaMemcpyAsync(out->mesh_bounds + 6 * m, s.bounds, sizeof(float) * 6, cudaMemcpyDeviceToDevice, st));
        Bvh2Node* n2 = out->nodes2 + out->tri_off[m];
        PT_TRY(build_tree(s, n, n2, st));
        k_reorder_tris<<<blocks_for(n), kThreads, 0, st>>>(s.tmp_tris, s.vals_out, n, out->tris2 + out->tri_off[m]);
        CollapseArgs a{};
        a.nodes2 = n2; a.range = s.range; a.order = s.vals_out; a.leaf_lo = s.leaf_lo; a.leaf_hi = s.leaf_hi;
        a.max_leaf = s.max_leaf; a.nodes8 = out->nodes8; a.node_counter = node_counter; a.tri_counter = tri_counter;
        a.tris_sorted = out->tris2 + out->tri_off[m]; a.tris8 = out->tris8; a.leaf_ids = nullptr; a.leaf_out = nullptr;
        a.wq2 = s.wq2; a.wq8 = s.wq8; a.tail = tail; a.capacity = L + 8; a.overflow = overflow;
        PT_TRY(collapse_tree(s, a, n, &out->blas_root8[m], st));
    }
    // ---- TLAS
    {
        std::vector<uint2> rg(std::max<size_t>(M, 1), make_uint2(0u, 0u));
        uint64_t instanced = 0;
        for (size_t m = 0; m < M; ++m) { rg[m] = make_uint2(out->tri_off[m], out->tri_count[m]); out->max_mesh_tris = std::max(out->max_mesh_tris, out->tri_count[m]); }
        for (uint32_t i : valid_inst) instanced += out->tri_count[h_instances[i].mesh_index];
        PT_CU(cudaMalloc(&out->mesh_tri_range, sizeof(uint2) * rg.size()));
        PT_CU(cudaMemcpyAsync(out->mesh_tri_range, rg.data(), sizeof(uint2) * rg.size(), cudaMemcpyHostToDevice, st));
        PT_CU(cudaStreamSynchronize(st));
        // exact boxes touch every instanced triangle once per (re)build: bounded so that it stays well under a millisecond
        out->exact_instance_boxes = instanced <= (64ull << 20);
    }
    PT_CU(cudaMemcpyAsync(&out->tlas_node_base, node_counter, 4, cudaMemcpyDeviceToHost, st));
    PT_CU(cudaStreamSynchronize(st));
    out->valid_inst = valid_inst;
    out->builder_flags = builder_flags;
    PT_TRY(build_tlas(s, out, d_instances, st));
    PT_CU(cudaMemcpyAsync(&out->num_nodes8, node_counter, 4, cudaMemcpyDeviceToHost, st));
    PT_CU(cudaEventRecord(ev1, st));
    PT_CU(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&out->ms_build, ev0, ev1);
    out->refit_repairs = s.refit_repairs;
    if (s.refit_repairs) fprintf(stderr, "b200pt: BVH refit verification corrected %u nodes\n", s.refit_repairs);
    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    s.release();
#undef PT_TRY
#undef PT_CU
    return PT_OK;
}

// Animated instances: the instance transforms changed (d_instances holds the new ones); mesh indices and the
// BLASes are unchanged.  The reference has no TLAS update path (Tlas.init is its only builder), so this is
// a full rebuild of the top level, which costs well under a millisecond for 1 000 instances.
int rebuild_tlas(SceneBvh* bvh, const PtInstance* d_instances, cudaStream_t st, float* ms_out) {
    const uint32_t NI = (uint32_t)bvh->valid_inst.size();
    if (ms_out) *ms_out = 0.0f;
    if (NI == 0) return PT_OK;
    cudaEvent_t ev0, ev1;
    PT_CUDA_CHECK(cudaEventCreate(&ev0)); PT_CUDA_CHECK(cudaEventCreate(&ev1));
    Scratch s;
    int rc = alloc_scratch(s, NI, 1, bvh->builder_flags, st);
    s.capacity = NI + 8;
    if (rc == PT_OK) {
        cudaEventRecord(ev0, st);
        rc = build_tlas(s, bvh, d_instances, st);
        cudaEventRecord(ev1, st);
        cudaStreamSynchronize(st);
        if (rc == PT_OK && ms_out) cudaEventElapsedTime(ms_out, ev0, ev1);
    }
    bvh->refit_repairs += s.refit_repairs;
    s.release();
    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    return rc;
}

}  // namespace pt
