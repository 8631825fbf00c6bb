// bvh_sah.cuh — two opt-in builder stages (included by bvh_build.cu; replaces nothing by default):
//
//  (1) build_tree_sah(): top-down binned-SAH binary hierarchy on the GPU, level-synchronous.  Every level
//      bins the centroids of all live segments at once (32 bins x 3 axes per segment, float min/max + count
//      atomics), one thread per segment sweeps its bins for the cheapest split, and a stable partition
//      (flags -> exclusive scan -> scatter) moves the leaves.  Leaves end up in hierarchy order, every node
//      covers a contiguous run of them, node ids are derived from the split position (the gap between two
//      adjacent leaves; root swapped to 0), so the result — topology, numbering and leaf order — does not
//      depend on atomic ordering.  Output has the layout of the Karras stage (Bvh2Node with child references,
//      range, parent links), so refit, collapse and the BVH2 walker run unchanged.
//  (2) k_collapse_dp: the cost table of the cost-optimal BVH2 -> BVH8 collapse (Ylitie, Karras, Laine 2017,
//      section 4.1): F[n][i] = least SAH cost of representing the subtree of binary node n by at most i slots
//      of one wide node; k_collapse then follows the recorded decisions instead of expanding greedily.
//
// Why: tools/bvh_lab.cpp (a CPU model of the shipped builder, validated against the GPU's node count and
// per-ray counts on C2) predicts 14.0 + 3.8 -> 11.5 + 4.2 (nodes + triangles per ray) for SAH + optimal collapse,
// i.e. -13 % of K-trace's cost units; profiles/r1_experiments.md.
//
// The kernel bodies use nothing but a global thread index, 32-bit atomics and fences, so the same source is
// also compiled as plain C++ (PT_EMULATE: threads run one after the other) by tools/bvh_lab.cpp and
// tests/test_bvh_sah_model.py, which check the trees it builds on the CPU.  That model is a design/test tool:
// the product runs the CUDA build only.
#pragma once
#include <stdint.h>

#ifdef PT_EMULATE
#include <cmath>
#include <cstring>
namespace pt {
struct alignas(16) Bvh2Node { float lo[3]; uint32_t left; float hi[3]; uint32_t right; };
struct uint2 { uint32_t x, y; };
static inline uint2 make_uint2(uint32_t x, uint32_t y) { uint2 r; r.x = x; r.y = y; return r; }
static thread_local uint32_t pt_emu_tid = 0;
#define PT_KERNEL static void
#define PT_DEV static inline
#define PT_GLOBAL_TID pt_emu_tid
#define PT_LAUNCH(kernel, total, st, ...) do { for (uint32_t _t = 0; _t < (uint32_t)(total); ++_t) { pt::pt_emu_tid = _t; kernel(__VA_ARGS__); } } while (0)
typedef int cudaStream_t;
static inline uint32_t atomicAdd(uint32_t* p, uint32_t v) { uint32_t o = *p; *p += v; return o; }
static inline void atomic_min_f(float* a, float v) { if (v < *a) *a = v; }
static inline void atomic_max_f(float* a, float v) { if (v > *a) *a = v; }
static inline void __threadfence() {}
static inline int sah_memset(void* p, int byte, size_t bytes, cudaStream_t) { std::memset(p, byte, bytes); return 0; }
static inline int sah_d2h(void* dst, const void* src, size_t bytes, cudaStream_t) { std::memcpy(dst, src, bytes); return 0; }
static inline int sah_d2d(void* dst, const void* src, size_t bytes, cudaStream_t) { std::memcpy(dst, src, bytes); return 0; }
static inline int sah_scan(const uint32_t* in, uint32_t* out, uint32_t n, cudaStream_t, void*, size_t) {
    uint32_t acc = 0;
    for (uint32_t i = 0; i < n; ++i) { out[i] = acc; acc += in[i]; }
    return 0;
}
#else
namespace pt {
#define PT_KERNEL __global__ void
#define PT_DEV __device__ __forceinline__
#define PT_GLOBAL_TID (blockIdx.x * blockDim.x + threadIdx.x)
#define PT_LAUNCH(kernel, total, st, ...) kernel<<<blocks_for(total), kThreads, 0, st>>>(__VA_ARGS__)
static inline int sah_memset(void* p, int byte, size_t bytes, cudaStream_t st) { return cudaMemsetAsync(p, byte, bytes, st) == cudaSuccess ? 0 : 1; }
static inline int sah_d2h(void* dst, const void* src, size_t bytes, cudaStream_t st) {
    if (cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st) != cudaSuccess) return 1;
    return cudaStreamSynchronize(st) == cudaSuccess ? 0 : 1;
}
static inline int sah_d2d(void* dst, const void* src, size_t bytes, cudaStream_t st) { return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, st) == cudaSuccess ? 0 : 1; }
static inline int sah_scan(const uint32_t* in, uint32_t* out, uint32_t n, cudaStream_t st, void* temp, size_t temp_bytes) {
    return cub::DeviceScan::ExclusiveSum(temp, temp_bytes, in, out, (int)n, st) == cudaSuccess ? 0 : 1;
}
#endif

constexpr int kSahBins = 32;     // 16 / 32 / 64 / 128 bins give the same SAH cost and per-ray counts within +-5 % in the CPU model (tools/bvh_lab)
constexpr uint32_t kSahNone = 0xFFFFFFFFu;
constexpr uint32_t kSahNoSplit = 3u;     // SahSeg::axis: leaves stay where they are, cut at `mid`

struct SahBin { float lo[3], hi[3]; uint32_t count; };                 // 28 B
struct SahSeg {                      // one live segment of the current level
    uint32_t first, last;            // leaf positions, inclusive
    uint32_t parent;                 // node id of the parent, kSahNone for the root
    uint32_t side;                   // 0 = left child of parent, 1 = right
    // decided by k_sah_split:
    uint32_t axis, bin, mid;         // positions [first, mid) go left
    uint32_t child_slot[2];          // next level's segment of each child, kSahNone for a single leaf
};

struct SahArgs {
    uint32_t n;
    const float* leaf_lo;            // per source leaf
    const float* leaf_hi;
    // ping-pong per level
    const uint32_t* idx_in;          // position -> source leaf
    const uint32_t* seg_in;          // position -> live segment of this level, kSahNone when settled
    uint32_t* idx_out;
    uint32_t* seg_out;
    SahSeg* segs;                    // this level
    SahSeg* next_segs;               // next level
    uint32_t* counters;              // [0] segments of the next level, [1] root gap
    float* cb;                       // 6 per segment: centroid bounds
    SahBin* bins;                    // 3 * kSahBins per segment
    uint32_t* flag;                  // per position: goes left
    const uint32_t* scan;            // exclusive sum of flag
    uint32_t force_median;           // level budget exhausted: split every segment in the middle
    // outputs
    Bvh2Node* nodes;                 // n - 1 internal nodes, boxes filled in by the refit
    uint2* range;                    // per node (first, last)
    uint32_t* parent;                // Karras layout: internal j at [j], leaf position j at [(n - 1) + j]
};

PT_DEV float sah_centroid(const SahArgs& a, uint32_t leaf, int axis) {
    return 0.5f * (a.leaf_lo[3 * (size_t)leaf + axis] + a.leaf_hi[3 * (size_t)leaf + axis]);
}
PT_DEV int sah_bin_of(float c, float lo, float ext) {
    int b = (int)((c - lo) * ((float)kSahBins / ext));
    return b < 0 ? 0 : (b > kSahBins - 1 ? kSahBins - 1 : b);
}
PT_DEV float sah_half_area(const float* lo, const float* hi) {
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

PT_KERNEL k_sah_init(SahArgs a, uint32_t num_segs) {
    const uint32_t s = PT_GLOBAL_TID;
    if (s >= num_segs) return;
    for (int k = 0; k < 3; ++k) { a.cb[6 * (size_t)s + k] = 3.0e38f; a.cb[6 * (size_t)s + 3 + k] = -3.0e38f; }
    SahBin* b = a.bins + (size_t)s * 3 * kSahBins;
    for (int i = 0; i < 3 * kSahBins; ++i) {
        for (int k = 0; k < 3; ++k) { b[i].lo[k] = 3.0e38f; b[i].hi[k] = -3.0e38f; }
        b[i].count = 0u;
    }
}

// centroid bounds of every live segment with more than two leaves
PT_KERNEL k_sah_cbounds(SahArgs a) {
    const uint32_t p = PT_GLOBAL_TID;
    if (p >= a.n) return;
    const uint32_t s = a.seg_in[p];
    if (s == kSahNone) return;
    if (a.segs[s].last - a.segs[s].first < 2u) return;
    const uint32_t leaf = a.idx_in[p];
    for (int k = 0; k < 3; ++k) {
        const float c = sah_centroid(a, leaf, k);
        atomic_min_f(a.cb + 6 * (size_t)s + k, c);
        atomic_max_f(a.cb + 6 * (size_t)s + 3 + k, c);
    }
}

PT_KERNEL k_sah_bin(SahArgs a) {
    const uint32_t p = PT_GLOBAL_TID;
    if (p >= a.n) return;
    const uint32_t s = a.seg_in[p];
    if (s == kSahNone) return;
    if (a.segs[s].last - a.segs[s].first < 2u) return;
    const uint32_t leaf = a.idx_in[p];
    float lo[3], hi[3];
    for (int k = 0; k < 3; ++k) { lo[k] = a.leaf_lo[3 * (size_t)leaf + k]; hi[k] = a.leaf_hi[3 * (size_t)leaf + k]; }
    for (int ax = 0; ax < 3; ++ax) {
        const float clo = a.cb[6 * (size_t)s + ax], ext = a.cb[6 * (size_t)s + 3 + ax] - clo;
        if (!(ext > 0.0f)) continue;
        SahBin* b = a.bins + ((size_t)s * 3 + ax) * kSahBins + sah_bin_of(0.5f * (lo[ax] + hi[ax]), clo, ext);
        for (int k = 0; k < 3; ++k) { atomic_min_f(&b->lo[k], lo[k]); atomic_max_f(&b->hi[k], hi[k]); }
        atomicAdd(&b->count, 1u);
    }
}

// id of the internal node whose split lies between leaf positions g and g + 1.  Every internal node of a binary
// tree over an ordered run of leaves owns exactly one such gap; the root takes id 0 and gap 0's node the root's gap.
PT_DEV uint32_t sah_node_id(uint32_t g, uint32_t root_gap) { return g == root_gap ? 0u : (g == 0u ? root_gap : g); }

// one thread per live segment: choose the split, create the node, register the children
PT_KERNEL k_sah_split(SahArgs a, uint32_t num_segs) {
    const uint32_t s = PT_GLOBAL_TID;
    if (s >= num_segs) return;
    SahSeg sg = a.segs[s];
    const uint32_t cnt = sg.last - sg.first + 1u;
    uint32_t axis = kSahNoSplit, bin = 0u, mid = sg.first + cnt / 2u;
    if (cnt > 2u && !a.force_median) {
        float best = 3.0e38f;
        for (int ax = 0; ax < 3; ++ax) {
            const float ext = a.cb[6 * (size_t)s + 3 + ax] - a.cb[6 * (size_t)s + ax];
            if (!(ext > 0.0f)) continue;
            const SahBin* b = a.bins + ((size_t)s * 3 + ax) * kSahBins;
            float right_cost[kSahBins];
            float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
            uint32_t c = 0u;
            for (int i = kSahBins - 1; i > 0; --i) {
                for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], b[i].lo[k]); hi[k] = fmaxf(hi[k], b[i].hi[k]); }
                c += b[i].count;
                right_cost[i] = c ? sah_half_area(lo, hi) * (float)c : 3.0e38f;
            }
            for (int k = 0; k < 3; ++k) { lo[k] = 3.0e38f; hi[k] = -3.0e38f; }
            c = 0u;
            for (int i = 0; i < kSahBins - 1; ++i) {
                for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], b[i].lo[k]); hi[k] = fmaxf(hi[k], b[i].hi[k]); }
                c += b[i].count;
                if (c == 0u || c == cnt) continue;
                const float cost = sah_half_area(lo, hi) * (float)c + right_cost[i + 1];
                if (cost < best) { best = cost; axis = (uint32_t)ax; bin = (uint32_t)i; mid = sg.first + c; }
            }
        }
    }
    // level 0 has exactly one segment, the root: its gap is published before any other node needs it
    const uint32_t gap = mid - 1u;
    if (sg.parent == kSahNone) a.counters[1] = gap;
    const uint32_t id = sg.parent == kSahNone ? 0u : sah_node_id(gap, a.counters[1]);
    a.range[id] = make_uint2(sg.first, sg.last);
    a.parent[id] = sg.parent;
    if (sg.parent != kSahNone) { if (sg.side) a.nodes[sg.parent].right = id; else a.nodes[sg.parent].left = id; }
    const uint32_t cf[2] = {sg.first, mid}, cl[2] = {mid - 1u, sg.last};
    for (uint32_t side = 0; side < 2u; ++side) {
        if (cf[side] == cl[side]) {
            // single leaf: settled.  (With kSahNoSplit the leaf is already at its position; otherwise the scatter puts it there.)
            const uint32_t ref = cf[side] | 0x80000000u;
            if (side) a.nodes[id].right = ref; else a.nodes[id].left = ref;
            a.parent[(size_t)(a.n - 1u) + cf[side]] = id;
            sg.child_slot[side] = kSahNone;
        } else {
            const uint32_t ns = atomicAdd(a.counters, 1u);
            SahSeg c;
            c.first = cf[side]; c.last = cl[side]; c.parent = id; c.side = side;
            c.axis = kSahNoSplit; c.bin = 0u; c.mid = 0u; c.child_slot[0] = c.child_slot[1] = kSahNone;
            a.next_segs[ns] = c;
            sg.child_slot[side] = ns;
        }
    }
    sg.axis = axis; sg.bin = bin; sg.mid = mid;
    a.segs[s] = sg;
}

PT_KERNEL k_sah_flags(SahArgs a) {
    const uint32_t p = PT_GLOBAL_TID;
    if (p >= a.n) return;
    const uint32_t s = a.seg_in[p];
    uint32_t f = 0u;
    if (s != kSahNone) {
        const SahSeg& sg = a.segs[s];
        if (sg.axis == kSahNoSplit) f = p < sg.mid ? 1u : 0u;
        else {
            const int ax = (int)sg.axis;
            const float clo = a.cb[6 * (size_t)s + ax], ext = a.cb[6 * (size_t)s + 3 + ax] - clo;
            f = sah_bin_of(sah_centroid(a, a.idx_in[p], ax), clo, ext) <= (int)sg.bin ? 1u : 0u;
        }
    }
    a.flag[p] = f;
}

// stable partition of every live segment; settled positions are copied through
PT_KERNEL k_sah_scatter(SahArgs a) {
    const uint32_t p = PT_GLOBAL_TID;
    if (p >= a.n) return;
    const uint32_t s = a.seg_in[p];
    if (s == kSahNone) { a.idx_out[p] = a.idx_in[p]; a.seg_out[p] = kSahNone; return; }
    const SahSeg& sg = a.segs[s];
    const uint32_t left_before = a.scan[p] - a.scan[sg.first];
    const uint32_t dst = a.flag[p] ? sg.first + left_before : sg.mid + ((p - sg.first) - left_before);
    a.idx_out[dst] = a.idx_in[p];
    a.seg_out[dst] = sg.child_slot[dst < sg.mid ? 0 : 1];
}

PT_KERNEL k_sah_start(SahArgs a, uint32_t* idx, uint32_t* seg) {
    const uint32_t p = PT_GLOBAL_TID;
    if (p >= a.n) return;
    idx[p] = p; seg[p] = 0u;
    if (p == 0u) {
        SahSeg r;
        r.first = 0u; r.last = a.n - 1u; r.parent = kSahNone; r.side = 0u;
        r.axis = kSahNoSplit; r.bin = 0u; r.mid = 0u; r.child_slot[0] = r.child_slot[1] = kSahNone;
        a.segs[0] = r;
    }
}

struct SahScratch {
    uint32_t* idx[2];        // n each
    uint32_t* seg[2];        // n each
    SahSeg* segs[2];         // n / 2 + 1 each
    uint32_t* counters;      // 2
    float* cb;               // 6 * (n / 2 + 1)
    SahBin* bins;            // 3 * kSahBins * (n / 2 + 1)
    uint32_t* flag;          // n
    uint32_t* scan;          // n
    void* scan_temp; size_t scan_temp_bytes;
};

// n >= 2 leaves.  On return: nodes / range / parent describe the hierarchy (boxes not yet set), *order_out points at
// the buffer of sc.idx that holds position -> source leaf.  Returns 0, or 1 on a CUDA error, 2 when the level budget ran out.
static int build_tree_sah(const SahScratch& sc, uint32_t n, const float* leaf_lo, const float* leaf_hi, Bvh2Node* nodes, uint2* range,
                          uint32_t* parent, const uint32_t** order_out, uint32_t* levels_out, cudaStream_t st) {
    SahArgs a;
    a.n = n; a.leaf_lo = leaf_lo; a.leaf_hi = leaf_hi;
    a.counters = sc.counters; a.cb = sc.cb; a.bins = sc.bins; a.flag = sc.flag; a.scan = sc.scan;
    a.nodes = nodes; a.range = range; a.parent = parent;
    a.force_median = 0u;
    a.segs = sc.segs[0]; a.next_segs = sc.segs[1];
    a.idx_in = sc.idx[0]; a.seg_in = sc.seg[0]; a.idx_out = sc.idx[1]; a.seg_out = sc.seg[1];
    PT_LAUNCH(k_sah_start, n, st, a, sc.idx[0], sc.seg[0]);
    uint32_t num_segs = 1u, level = 0u;
    int cur = 0;
    // a SAH hierarchy over n leaves is at most a few times log2(n) deep on real geometry; past the budget every
    // segment is halved, which ends after log2(n) more levels
    const uint32_t budget = 96u;
    while (num_segs) {
        if (level > budget + 40u) return 2;
        a.force_median = level >= budget ? 1u : 0u;
        a.segs = sc.segs[cur]; a.next_segs = sc.segs[cur ^ 1];
        a.idx_in = sc.idx[cur]; a.seg_in = sc.seg[cur]; a.idx_out = sc.idx[cur ^ 1]; a.seg_out = sc.seg[cur ^ 1];
        if (sah_memset(sc.counters, 0, 4, st)) return 1;
        PT_LAUNCH(k_sah_init, num_segs, st, a, num_segs);
        PT_LAUNCH(k_sah_cbounds, n, st, a);
        PT_LAUNCH(k_sah_bin, n, st, a);
        PT_LAUNCH(k_sah_split, num_segs, st, a, num_segs);
        PT_LAUNCH(k_sah_flags, n, st, a);
        if (sah_scan(sc.flag, sc.scan, n, st, sc.scan_temp, sc.scan_temp_bytes)) return 1;
        PT_LAUNCH(k_sah_scatter, n, st, a);
        uint32_t next = 0u;
        if (sah_d2h(&next, sc.counters, 4, st)) return 1;
        if (next > n / 2u + 1u) return 2;
        num_segs = next;
        cur ^= 1;
        ++level;
    }
    *order_out = sc.idx[cur];
    if (levels_out) *levels_out = level;
    return 0;
}

// ------------------------------------------------------------------ cost table of the optimal collapse
// F / D: 8 entries per internal node.  F[i], i = 1..7: least cost of the subtree as at most i slots.
// D[0] = slots given to the left child when the node becomes a wide node (8 in total); D[1] = 0: one slot holding a
// leaf, 8: one slot holding an inner wide node; D[i >= 2] = 0: same as i - 1, else slots given to the left child.
struct DpArgs {
    uint32_t n;                  // leaves
    const Bvh2Node* nodes;       // boxes valid (after the refit)
    const uint2* range;          // count = y - x + 1
    const uint32_t* parent;      // Karras layout
    const uint32_t* order;       // leaf position -> source leaf
    const float* leaf_lo;
    const float* leaf_hi;
    uint32_t* visit;             // n - 1, zeroed
    float* F;
    uint8_t* D;
    uint32_t max_leaf;
    float c_node, c_tri;
};

PT_DEV float dp_child_cost(const DpArgs& a, uint32_t ref, int i) {
    if (ref & 0x80000000u) {
        const size_t src = a.order[ref & 0x7fffffffu];
        return a.c_tri * sah_half_area(a.leaf_lo + 3 * src, a.leaf_hi + 3 * src);
    }
    return reinterpret_cast<const volatile float*>(a.F)[(size_t)ref * 8 + i];   // written by another thread of this launch
}

PT_KERNEL k_collapse_dp(DpArgs a) {
    const uint32_t leaf = PT_GLOBAL_TID;
    if (leaf >= a.n) return;
    uint32_t cur = a.parent[(size_t)(a.n - 1u) + leaf];
    while (cur != 0xFFFFFFFFu) {
        __threadfence();
        if (atomicAdd(&a.visit[cur], 1u) == 0u) return;      // first arrival: the sibling subtree is not done
        __threadfence();
        const uint32_t l = a.nodes[cur].left, r = a.nodes[cur].right;
        const float A = sah_half_area(a.nodes[cur].lo, a.nodes[cur].hi);
        const uint32_t count = a.range[cur].y - a.range[cur].x + 1u;
        float fl[8], fr[8];
        for (int i = 1; i <= 7; ++i) { fl[i] = dp_child_cost(a, l, i); fr[i] = dp_child_cost(a, r, i); }
        float best8 = 3.0e38f; int k8 = 1;
        for (int k = 1; k <= 7; ++k) {
            const float c = fl[k] + fr[8 - k];
            if (c < best8) { best8 = c; k8 = k; }
        }
        const float c_inner = a.c_node * A + best8;
        const float c_leaf = count <= a.max_leaf ? a.c_tri * A * (float)count : 3.0e38f;
        float f[8]; uint8_t d[8];
        d[0] = (uint8_t)k8;
        if (c_leaf <= c_inner) { f[1] = c_leaf; d[1] = 0; } else { f[1] = c_inner; d[1] = 8; }
        for (int i = 2; i <= 7; ++i) {
            f[i] = f[i - 1]; d[i] = 0;
            for (int k = 1; k < i; ++k) {
                const float c = fl[k] + fr[i - k];
                if (c < f[i]) { f[i] = c; d[i] = (uint8_t)k; }
            }
        }
        f[0] = c_inner;
        for (int i = 0; i < 8; ++i) { a.F[(size_t)cur * 8 + i] = f[i]; a.D[(size_t)cur * 8 + i] = d[i]; }
        cur = a.parent[cur];
    }
}

// Children of the wide node rooted at binary node b2 under the decisions D of k_collapse_dp: the node's 8 slots are split
// D[0] : 8 - D[0] between its two subtrees, and a subtree that was given i slots either stays one slot (a leaf, or an
// inner wide node of its own) or hands k : i - k to its children.  refs[] receives binary child references left to
// right, is_leaf[] whether the slot is a leaf; returns the number of slots used (<= 8).  Used by k_collapse.
PT_DEV int dp_cover(const uint8_t* D, const Bvh2Node* nodes, uint32_t b2, uint32_t* refs, bool* is_leaf) {
    uint32_t st_ref[16]; int st_i[16];
    int sp = 0, cnt = 0;
    const int k8 = D[8 * (size_t)b2];
    st_ref[sp] = nodes[b2].right; st_i[sp] = 8 - k8; ++sp;
    st_ref[sp] = nodes[b2].left; st_i[sp] = k8; ++sp;
    while (sp && cnt < 8) {
        --sp;
        const uint32_t ref = st_ref[sp];
        int i = st_i[sp];
        if (ref & 0x80000000u) { refs[cnt] = ref; is_leaf[cnt] = true; ++cnt; continue; }
        const uint8_t* d = D + 8 * (size_t)ref;
        while (i >= 2 && d[i] == 0) --i;
        if (i <= 1) { refs[cnt] = ref; is_leaf[cnt] = d[1] == 0; ++cnt; continue; }
        st_ref[sp] = nodes[ref].right; st_i[sp] = i - (int)d[i]; ++sp;
        st_ref[sp] = nodes[ref].left; st_i[sp] = (int)d[i]; ++sp;
    }
    return cnt;
}

}  // namespace pt
