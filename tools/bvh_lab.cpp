// tools/bvh_lab.cpp — CPU model of the shipped acceleration-structure builder and of candidate replacements,
// scored with the ordered walker's per-ray counts (nodes decoded, triangles tested: the quantities
// pt_count_traversal reports on the GPU and that K-trace's time follows, profiles/r1_experiments.md).
//
// NOT product code and not the oracle: a design tool for picking the next builder change without spending GPU
// minutes.  It restates, with the same formulas as csrc/bvh_build.cu:
//   * the 63-bit Morton order + Karras radix tree (k_morton / k_karras), boxes refitted bottom-up;
//   * the greedy surface-area collapse to 8-wide nodes (k_collapse), octant slot assignment, 8-bit quantised
//     child boxes (the walker tests the DEQUANTISED boxes, so the counts include the quantisation slack);
//   * the walker's order (trace.cuh:trace_bvh8): a node's hit triangles first, then its hit inner children by
//     octant priority, each tested against the closest hit so far.
// and adds: a top-down binned-SAH binary builder, and the cost-optimal collapse of Ylitie et al. 2017 (dynamic
// programme over "cover this subtree with at most i slots").
//
// usage: bvh_lab <tris.bin> <width> <height> <pixel step>     (tris.bin: u32 n, then n x 9 floats)
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <numeric>
#include <vector>

#define PT_EMULATE 1
#include "../vulkan-pathtracer_b200/csrc/bvh_sah.cuh"

struct Box {
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    void grow(const Box& b) { for (int k = 0; k < 3; ++k) { lo[k] = std::fmin(lo[k], b.lo[k]); hi[k] = std::fmax(hi[k], b.hi[k]); } }
    float half_area() const { float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2]; return dx * dy + dy * dz + dz * dx; }
};

// binary hierarchy: node 0 is the root; a child reference is an internal node index, or (sorted leaf | LEAF)
constexpr uint32_t LEAF = 0x80000000u;
struct B2 { Box b; uint32_t left, right; uint32_t count; };
struct Bvh2 { std::vector<B2> nodes; std::vector<uint32_t> order; /* sorted leaf -> triangle */ };

static std::vector<float> g_tris;     // 9 floats per triangle
static std::vector<Box> g_leaf;       // per triangle
static uint32_t g_n = 0;

// ------------------------------------------------------------------ LBVH (k_morton + k_karras + k_refit)
static uint64_t expand21(uint32_t v) {
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

static int delta(const std::vector<uint64_t>& keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __builtin_clz((uint32_t)i ^ (uint32_t)j);
    return __builtin_clzll(a ^ b);
}

static Box ref_box(const Bvh2& t, uint32_t ref) { return (ref & LEAF) ? g_leaf[t.order[ref & ~LEAF]] : t.nodes[ref].b; }
static uint32_t ref_count(const Bvh2& t, uint32_t ref) { return (ref & LEAF) ? 1u : t.nodes[ref].count; }

static void refit(Bvh2& t, uint32_t i) {      // post-order, explicit stack
    std::vector<std::pair<uint32_t, int>> st;
    st.push_back({i, 0});
    while (!st.empty()) {
        auto [n, phase] = st.back();
        st.pop_back();
        B2& nd = t.nodes[n];
        if (phase == 0) {
            st.push_back({n, 1});
            if (!(nd.left & LEAF)) st.push_back({nd.left, 0});
            if (!(nd.right & LEAF)) st.push_back({nd.right, 0});
        } else {
            nd.b = Box();
            nd.b.grow(ref_box(t, nd.left)); nd.b.grow(ref_box(t, nd.right));
            nd.count = ref_count(t, nd.left) + ref_count(t, nd.right);
        }
    }
}

static Bvh2 build_lbvh() {
    Bvh2 t;
    const int n = (int)g_n;
    Box all;
    for (auto& b : g_leaf) all.grow(b);
    std::vector<uint64_t> key(n);
    for (int i = 0; i < n; ++i) {
        uint32_t q[3];
        for (int a = 0; a < 3; ++a) {
            float lo = all.lo[a], ext = all.hi[a] - all.lo[a];
            float c = 0.5f * (g_leaf[i].lo[a] + g_leaf[i].hi[a]);
            float u = ext > 0.0f ? (c - lo) / ext : 0.0f;
            u = std::fmin(std::fmax(u, 0.0f), 1.0f);
            q[a] = std::min(2097151u, (uint32_t)(u * 2097152.0f));
        }
        key[i] = (expand21(q[0]) << 2) | (expand21(q[1]) << 1) | expand21(q[2]);
    }
    t.order.resize(n);
    std::iota(t.order.begin(), t.order.end(), 0u);
    std::stable_sort(t.order.begin(), t.order.end(), [&](uint32_t a, uint32_t b) { return key[a] < key[b]; });
    std::vector<uint64_t> keys(n);
    for (int i = 0; i < n; ++i) keys[i] = key[t.order[i]];
    t.nodes.resize(n - 1);
    for (int i = 0; i < n - 1; ++i) {
        int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
        int dmin = delta(keys, n, i, i - d);
        int lmax = 2;
        while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
        int l = 0;
        for (int s = lmax >> 1; s >= 1; s >>= 1)
            if (delta(keys, n, i, i + (l + s) * d) > dmin) l += s;
        int j = i + l * d;
        int dnode = delta(keys, n, i, j);
        int s = 0, tt = l;
        do {
            tt = (tt + 1) >> 1;
            if (delta(keys, n, i, i + (s + tt) * d) > dnode) s += tt;
        } while (tt > 1);
        int gamma = i + s * d + std::min(d, 0);
        int first = std::min(i, j), last = std::max(i, j);
        t.nodes[i].left = (first == gamma) ? ((uint32_t)gamma | LEAF) : (uint32_t)gamma;
        t.nodes[i].right = (last == gamma + 1) ? ((uint32_t)(gamma + 1) | LEAF) : (uint32_t)(gamma + 1);
    }
    refit(t, 0);
    return t;
}

// ------------------------------------------------------------------ top-down binned SAH (to single-triangle leaves)
static Bvh2 build_sah(int bins) {
    Bvh2 t;
    const uint32_t n = g_n;
    t.order.resize(n);
    std::iota(t.order.begin(), t.order.end(), 0u);
    t.nodes.reserve(n);
    std::vector<float> cen(3 * (size_t)n);
    for (uint32_t i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a) cen[3 * (size_t)i + a] = 0.5f * (g_leaf[i].lo[a] + g_leaf[i].hi[a]);
    struct Job { uint32_t node, first, last; };   // [first, last)
    t.nodes.push_back(B2());
    std::vector<Job> jobs{{0u, 0u, n}};
    std::vector<Box> bb(bins);
    std::vector<uint32_t> bc(bins);
    std::vector<float> right_area(bins);
    while (!jobs.empty()) {
        Job j = jobs.back();
        jobs.pop_back();
        Box nb, cb;
        for (uint32_t i = j.first; i < j.last; ++i) {
            const uint32_t tr = t.order[i];
            nb.grow(g_leaf[tr]);
            for (int a = 0; a < 3; ++a) { cb.lo[a] = std::fmin(cb.lo[a], cen[3 * (size_t)tr + a]); cb.hi[a] = std::fmax(cb.hi[a], cen[3 * (size_t)tr + a]); }
        }
        const uint32_t cnt = j.last - j.first;
        uint32_t mid = j.first + cnt / 2;
        if (cnt > 2) {
            float best = 3.0e38f; int best_axis = -1, best_bin = -1;
            for (int a = 0; a < 3; ++a) {
                const float ext = cb.hi[a] - cb.lo[a];
                if (!(ext > 0.0f)) continue;
                for (int b = 0; b < bins; ++b) { bb[b] = Box(); bc[b] = 0; }
                const float k = (float)bins / ext;
                for (uint32_t i = j.first; i < j.last; ++i) {
                    const uint32_t tr = t.order[i];
                    int b = std::min(bins - 1, (int)((cen[3 * (size_t)tr + a] - cb.lo[a]) * k));
                    bb[b].grow(g_leaf[tr]); bc[b]++;
                }
                Box r; uint32_t rc = 0;
                for (int b = bins - 1; b > 0; --b) { r.grow(bb[b]); rc += bc[b]; right_area[b] = rc ? r.half_area() * (float)rc : 3.0e38f; }
                Box l; uint32_t lc = 0;
                for (int b = 0; b < bins - 1; ++b) {
                    l.grow(bb[b]); lc += bc[b];
                    if (lc == 0 || lc == cnt) continue;
                    const float c = l.half_area() * (float)lc + right_area[b + 1];
                    if (c < best) { best = c; best_axis = a; best_bin = b; }
                }
            }
            if (best_axis >= 0) {
                const int a = best_axis;
                const float k = (float)bins / (cb.hi[a] - cb.lo[a]);
                auto it = std::partition(t.order.begin() + j.first, t.order.begin() + j.last, [&](uint32_t tr) {
                    return std::min(bins - 1, (int)((cen[3 * (size_t)tr + a] - cb.lo[a]) * k)) <= best_bin;
                });
                mid = (uint32_t)(it - t.order.begin());
            } else {
                // all centroids coincide: median split
            }
        }
        if (mid == j.first || mid == j.last) mid = j.first + cnt / 2;
        B2& nd = t.nodes[j.node];
        nd.b = nb; nd.count = cnt;
        auto child = [&](uint32_t f, uint32_t l) -> uint32_t {
            if (l - f == 1) return f | LEAF;
            const uint32_t idx = (uint32_t)t.nodes.size();
            t.nodes.push_back(B2());
            jobs.push_back({idx, f, l});
            return idx;
        };
        const uint32_t lref = child(j.first, mid), rref = child(mid, j.last);
        t.nodes[j.node].left = lref; t.nodes[j.node].right = rref;
    }
    return t;
}

// ------------------------------------------------------------------ the product's SAH stage (csrc/bvh_sah.cuh), threads run serially
static Bvh2 build_sah_kernel(uint32_t* levels) {
    const uint32_t n = g_n, S = n / 2 + 1;
    std::vector<float> lo(3 * (size_t)n), hi(3 * (size_t)n);
    for (uint32_t i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) { lo[3 * (size_t)i + k] = g_leaf[i].lo[k]; hi[3 * (size_t)i + k] = g_leaf[i].hi[k]; }
    std::vector<uint32_t> idx0(n), idx1(n), seg0(n), seg1(n), flag(n), scan(n), parent(2 * (size_t)n + 2), counters(2);
    std::vector<pt::SahSeg> s0(S), s1(S);
    std::vector<float> cb(6 * (size_t)S);
    std::vector<pt::SahBin> bins((size_t)S * 3 * pt::kSahBins);
    std::vector<pt::Bvh2Node> nodes(n - 1);
    std::vector<pt::uint2> range(n - 1);
    pt::SahScratch sc;
    sc.idx[0] = idx0.data(); sc.idx[1] = idx1.data(); sc.seg[0] = seg0.data(); sc.seg[1] = seg1.data();
    sc.segs[0] = s0.data(); sc.segs[1] = s1.data(); sc.counters = counters.data(); sc.cb = cb.data(); sc.bins = bins.data();
    sc.flag = flag.data(); sc.scan = scan.data(); sc.scan_temp = nullptr; sc.scan_temp_bytes = 0;
    const uint32_t* order = nullptr;
    const int rc = pt::build_tree_sah(sc, n, lo.data(), hi.data(), nodes.data(), range.data(), parent.data(), &order, levels, 0);
    if (rc) { std::fprintf(stderr, "build_tree_sah failed: %d\n", rc); std::exit(1); }
    Bvh2 t;
    t.order.assign(order, order + n);
    t.nodes.resize(n - 1);
    for (uint32_t i = 0; i + 1 < n; ++i) { t.nodes[i].left = nodes[i].left; t.nodes[i].right = nodes[i].right; }
    refit(t, 0);
    // invariants the collapse relies on: contiguous ranges, parent links, every leaf exactly once
    std::vector<uint8_t> seen(n, 0);
    for (uint32_t j = 0; j < n; ++j) { if (t.order[j] >= n || seen[t.order[j]]) { std::fprintf(stderr, "order is not a permutation\n"); std::exit(1); } seen[t.order[j]] = 1; }
    for (uint32_t i = 0; i + 1 < n; ++i) {
        if (range[i].y - range[i].x + 1 != t.nodes[i].count) { std::fprintf(stderr, "range/count mismatch at node %u\n", i); std::exit(1); }
        for (uint32_t ref : {t.nodes[i].left, t.nodes[i].right}) {
            const uint32_t pp = (ref & LEAF) ? parent[(size_t)(n - 1) + (ref & ~LEAF)] : parent[ref];
            if (pp != i) { std::fprintf(stderr, "parent link mismatch at node %u\n", i); std::exit(1); }
        }
    }
    if (parent[0] != 0xFFFFFFFFu) { std::fprintf(stderr, "root has a parent\n"); std::exit(1); }
    return t;
}

// decision table of the product's k_collapse_dp (serial), layout D[8 * node + i]
static std::vector<uint8_t> dp_kernel(const Bvh2& t, uint32_t max_leaf, float c_node, float c_tri) {
    const uint32_t n = g_n;
    std::vector<float> lo(3 * (size_t)n), hi(3 * (size_t)n), F(8 * (size_t)(n - 1));
    for (uint32_t i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) { lo[3 * (size_t)i + k] = g_leaf[i].lo[k]; hi[3 * (size_t)i + k] = g_leaf[i].hi[k]; }
    std::vector<pt::Bvh2Node> nodes(n - 1);
    std::vector<pt::uint2> range(n - 1);
    std::vector<uint32_t> parent(2 * (size_t)n + 2, 0xFFFFFFFFu), visit(n - 1, 0);
    for (uint32_t i = 0; i + 1 < n; ++i) {
        for (int k = 0; k < 3; ++k) { nodes[i].lo[k] = t.nodes[i].b.lo[k]; nodes[i].hi[k] = t.nodes[i].b.hi[k]; }
        nodes[i].left = t.nodes[i].left; nodes[i].right = t.nodes[i].right;
        range[i] = pt::make_uint2(0u, t.nodes[i].count - 1u);
        for (uint32_t ref : {t.nodes[i].left, t.nodes[i].right}) {
            if (ref & LEAF) parent[(size_t)(n - 1) + (ref & ~LEAF)] = i; else parent[ref] = i;
        }
    }
    std::vector<uint8_t> D(8 * (size_t)(n - 1));
    pt::DpArgs a;
    a.n = n; a.nodes = nodes.data(); a.range = range.data(); a.parent = parent.data(); a.order = t.order.data();
    a.leaf_lo = lo.data(); a.leaf_hi = hi.data(); a.visit = visit.data(); a.F = F.data(); a.D = D.data();
    a.max_leaf = max_leaf; a.c_node = c_node; a.c_tri = c_tri;
    PT_LAUNCH(pt::k_collapse_dp, n, 0, a);
    for (uint32_t i = 0; i + 1 < n; ++i) if (visit[i] != 2) { std::fprintf(stderr, "k_collapse_dp did not reach node %u\n", i); std::exit(1); }
    return D;
}

// ------------------------------------------------------------------ wide tree
struct WChild { Box qb, xb; int inner; uint32_t tri_first, tri_cnt; int slot; };   // quantised and exact box
struct WNode { int nc; WChild c[8]; };
struct Wide { std::vector<WNode> nodes; std::vector<uint32_t> tris; double sah = 0.0; };

static bool g_optimal_slots = false;
struct CC { uint32_t ref; uint32_t count; Box b; bool leaf; };

static void leaves_under(const Bvh2& t, uint32_t ref, std::vector<uint32_t>& out) {
    std::vector<uint32_t> st{ref};
    while (!st.empty()) {
        uint32_t r = st.back(); st.pop_back();
        if (r & LEAF) out.push_back(t.order[r & ~LEAF]);
        else { st.push_back(t.nodes[r].right); st.push_back(t.nodes[r].left); }
    }
}

// children of the wide node rooted at binary node b2 -> quantised node (same formulas as k_collapse)
static void emit_node(const Bvh2& t, Wide& w, int widx, std::vector<CC>& ch, std::vector<std::pair<uint32_t, int>>& todo) {
    const int cnt = (int)ch.size();
    Box nb;
    for (auto& c : ch) nb.grow(c.b);
    float p[3], scale[3];
    for (int k = 0; k < 3; ++k) {
        float ext = nb.hi[k] - nb.lo[k];
        float mag = std::fmax(std::fabs(nb.lo[k]), std::fabs(nb.hi[k]));
        ext = std::fmax(std::fmax(ext, mag * 1.2e-4f), 1e-30f) * 1.01f;
        int e;
        std::frexp(ext / 250.0f, &e);
        e = std::max(-125, std::min(126, e));
        scale[k] = std::ldexp(1.0f, e);
        p[k] = nb.lo[k] - 0.5f * scale[k];
    }
    int slot_of[8]; bool slot_used[8] = {false}, assigned[8] = {false};
    if (g_optimal_slots) {
        // exact assignment: maximise the sum of dot(child centroid - node centroid, octant direction) over all injective
        // child -> slot maps (dynamic programme over slot subsets, children in order)
        const float ncx[3] = {0.5f * (nb.lo[0] + nb.hi[0]), 0.5f * (nb.lo[1] + nb.hi[1]), 0.5f * (nb.lo[2] + nb.hi[2])};
        float score[8][8];
        for (int c = 0; c < cnt; ++c) {
            float d[3];
            for (int k = 0; k < 3; ++k) d[k] = 0.5f * (ch[c].b.lo[k] + ch[c].b.hi[k]) - ncx[k];
            for (int sl = 0; sl < 8; ++sl) score[c][sl] = ((sl & 1) ? d[0] : -d[0]) + ((sl & 2) ? d[1] : -d[1]) + ((sl & 4) ? d[2] : -d[2]);
        }
        static float best[9][256]; static int8_t pick[9][256];
        for (int m = 0; m < 256; ++m) best[cnt][m] = 0.0f;
        for (int c = cnt - 1; c >= 0; --c)
            for (int m = 0; m < 256; ++m) {
                best[c][m] = -3.0e38f; pick[c][m] = -1;
                if (__builtin_popcount(m) != c) continue;
                for (int sl = 0; sl < 8; ++sl) {
                    if (m >> sl & 1) continue;
                    const float v = score[c][sl] + best[c + 1][m | (1 << sl)];
                    if (v > best[c][m]) { best[c][m] = v; pick[c][m] = (int8_t)sl; }
                }
            }
        int m = 0;
        for (int c = 0; c < cnt; ++c) { slot_of[c] = pick[c][m]; m |= 1 << slot_of[c]; assigned[c] = true; }
    }
    const float nc[3] = {0.5f * (nb.lo[0] + nb.hi[0]), 0.5f * (nb.lo[1] + nb.hi[1]), 0.5f * (nb.lo[2] + nb.hi[2])};
    for (int it = 0; it < cnt && !g_optimal_slots; ++it) {
        float best = -3.0e38f; int bc = -1, bs = -1;
        for (int c = 0; c < cnt; ++c) {
            if (assigned[c]) continue;
            float dx = 0.5f * (ch[c].b.lo[0] + ch[c].b.hi[0]) - nc[0];
            float dy = 0.5f * (ch[c].b.lo[1] + ch[c].b.hi[1]) - nc[1];
            float dz = 0.5f * (ch[c].b.lo[2] + ch[c].b.hi[2]) - nc[2];
            for (int s = 0; s < 8; ++s) {
                if (slot_used[s]) continue;
                float cost = ((s & 1) ? dx : -dx) + ((s & 2) ? dy : -dy) + ((s & 4) ? dz : -dz);
                if (cost > best) { best = cost; bc = c; bs = s; }
            }
        }
        assigned[bc] = true; slot_used[bs] = true; slot_of[bc] = bs;
    }
    WNode out;
    out.nc = cnt;
    const float dq = 1.0f / 1024.0f;
    for (int c = 0; c < cnt; ++c) {
        WChild& o = out.c[c];
        o.slot = slot_of[c];
        o.xb = ch[c].b;
        for (int k = 0; k < 3; ++k) {
            float fl = std::floor((ch[c].b.lo[k] - p[k]) / scale[k] - dq);
            float fh = std::ceil((ch[c].b.hi[k] - p[k]) / scale[k] + dq);
            fl = std::fmin(std::fmax(fl, 0.0f), 255.0f); fh = std::fmin(std::fmax(fh, 0.0f), 255.0f);
            o.qb.lo[k] = p[k] + fl * scale[k]; o.qb.hi[k] = p[k] + fh * scale[k];
        }
        if (ch[c].leaf) {
            o.inner = -1;
            o.tri_first = (uint32_t)w.tris.size();
            leaves_under(t, ch[c].ref, w.tris);
            o.tri_cnt = (uint32_t)w.tris.size() - o.tri_first;
            w.sah += 0.7 * (double)ch[c].b.half_area() * o.tri_cnt;
        } else {
            o.inner = (int)w.nodes.size() + (int)todo.size() - 0;   // fixed below
            o.tri_first = o.tri_cnt = 0;
        }
    }
    // allocate inner children after the node itself is placed
    w.nodes[widx] = out;
    for (int c = 0; c < cnt; ++c) {
        if (ch[c].leaf) continue;
        const int idx = (int)w.nodes.size();
        w.nodes.push_back(WNode());
        w.nodes[widx].c[c].inner = idx;
        todo.push_back({ch[c].ref, idx});
        w.sah += (double)ch[c].b.half_area();
    }
}

// greedy surface-area expansion (the shipped collapse)
static Wide collapse_greedy(const Bvh2& t, uint32_t max_leaf) {
    Wide w;
    w.nodes.push_back(WNode());
    w.sah = t.nodes[0].b.half_area();
    std::vector<std::pair<uint32_t, int>> todo{{0u, 0}};
    size_t head = 0;
    auto load = [&](uint32_t ref) { CC c; c.ref = ref; c.count = ref_count(t, ref); c.b = ref_box(t, ref); c.leaf = (ref & LEAF) || c.count <= max_leaf; return c; };
    while (head < todo.size()) {
        auto [b2, widx] = todo[head++];
        std::vector<CC> ch;
        ch.push_back(load(t.nodes[b2].left)); ch.push_back(load(t.nodes[b2].right));
        while (ch.size() < 8) {
            int best = -1; float ba = -1.0f;
            for (int c = 0; c < (int)ch.size(); ++c) {
                if (ch[c].leaf) continue;
                float a = ch[c].b.half_area();
                if (a > ba) { ba = a; best = c; }
            }
            if (best < 0) break;
            const B2& nd = t.nodes[ch[best].ref];
            const uint32_t l = nd.left, r = nd.right;
            ch[best] = load(l);
            ch.push_back(load(r));
        }
        emit_node(t, w, widx, ch, todo);
    }
    return w;
}

// cost-optimal collapse (Ylitie, Karras, Laine 2017, section 4.1).  F[n][i] = least cost of representing the subtree
// of binary node n by at most i slots of one wide node (each slot a leaf of <= max_leaf triangles or an inner wide node).
static Wide collapse_optimal(const Bvh2& t, uint32_t max_leaf, float c_node, float c_tri, bool kernel_table = false) {
    const size_t N = t.nodes.size();
    std::vector<float> F(N * 8, 3.0e38f);
    std::vector<uint8_t> D(N * 8, 0);        // i >= 2: 0 = same as i-1, k = left gets k slots; i == 1: 0 = leaf, 8 = inner
    std::vector<uint8_t> D8(N, 0);           // inner: left gets k of the 8 slots
    const float INF = 3.0e38f;
    auto Fref = [&](uint32_t ref, int i) -> float {
        if (ref & LEAF) return c_tri * g_leaf[t.order[ref & ~LEAF]].half_area();
        return F[(size_t)ref * 8 + i];
    };
    // post-order over the binary tree
    std::vector<uint32_t> post; post.reserve(N);
    {
        std::vector<uint32_t> st{0u};
        while (!st.empty()) {
            uint32_t n = st.back(); st.pop_back(); post.push_back(n);
            if (!(t.nodes[n].left & LEAF)) st.push_back(t.nodes[n].left);
            if (!(t.nodes[n].right & LEAF)) st.push_back(t.nodes[n].right);
        }
        std::reverse(post.begin(), post.end());
    }
    for (uint32_t n : post) {
        const B2& nd = t.nodes[n];
        const float A = nd.b.half_area();
        const float c_leaf = nd.count <= max_leaf ? c_tri * A * (float)nd.count : INF;
        float best8 = INF; int k8 = 1;
        for (int k = 1; k <= 7; ++k) {
            float c = Fref(nd.left, k) + Fref(nd.right, 8 - k);
            if (c < best8) { best8 = c; k8 = k; }
        }
        const float c_inner = c_node * A + best8;
        D8[n] = (uint8_t)k8;
        float* f = &F[(size_t)n * 8];
        uint8_t* d = &D[(size_t)n * 8];
        if (c_leaf <= c_inner) { f[1] = c_leaf; d[1] = 0; } else { f[1] = c_inner; d[1] = 8; }
        for (int i = 2; i <= 7; ++i) {
            f[i] = f[i - 1]; d[i] = 0;
            for (int k = 1; k < i; ++k) {
                float c = Fref(nd.left, k) + Fref(nd.right, i - k);
                if (c < f[i]) { f[i] = c; d[i] = (uint8_t)k; }
            }
        }
    }
    if (kernel_table) {
        // the product's table must make the same decisions as the recursion above
        const std::vector<uint8_t> K = dp_kernel(t, max_leaf, c_node, c_tri);
        size_t diff = 0;
        for (size_t n = 0; n < N; ++n) {
            if (K[8 * n] != D8[n]) diff++;
            for (int i = 1; i <= 7; ++i) if (K[8 * n + i] != D[8 * n + i]) diff++;
            D8[n] = K[8 * n];
            for (int i = 1; i <= 7; ++i) D[8 * n + i] = K[8 * n + i];
        }
        std::printf("k_collapse_dp vs the lab's recursion: %zu differing decisions of %zu\n", diff, 8 * N);
    }
    Wide w;
    w.nodes.push_back(WNode());
    w.sah = t.nodes[0].b.half_area();
    std::vector<std::pair<uint32_t, int>> todo{{0u, 0}};
    size_t head = 0;
    std::function<void(uint32_t, int, std::vector<CC>&)> cover = [&](uint32_t ref, int i, std::vector<CC>& out) {
        if (ref & LEAF) { CC c; c.ref = ref; c.count = 1; c.b = ref_box(t, ref); c.leaf = true; out.push_back(c); return; }
        while (i >= 2 && D[(size_t)ref * 8 + i] == 0) --i;
        if (i == 1) {
            CC c; c.ref = ref; c.count = t.nodes[ref].count; c.b = t.nodes[ref].b; c.leaf = D[(size_t)ref * 8 + 1] == 0;
            out.push_back(c);
            return;
        }
        const int k = D[(size_t)ref * 8 + i];
        cover(t.nodes[ref].left, k, out);
        cover(t.nodes[ref].right, i - k, out);
    };
    std::vector<pt::Bvh2Node> kn;
    std::vector<uint8_t> KD;
    if (kernel_table) {   // k_collapse's own child selection (pt::dp_cover) on the kernel's table
        kn.resize(N); KD.resize(8 * N);
        for (size_t n = 0; n < N; ++n) {
            kn[n].left = t.nodes[n].left; kn[n].right = t.nodes[n].right;
            KD[8 * n] = D8[n];
            for (int i = 1; i <= 7; ++i) KD[8 * n + i] = D[8 * n + i];
        }
    }
    size_t cover_diff = 0;
    while (head < todo.size()) {
        auto [b2, widx] = todo[head++];
        std::vector<CC> ch;
        cover(t.nodes[b2].left, D8[b2], ch);
        cover(t.nodes[b2].right, 8 - D8[b2], ch);
        if (kernel_table) {
            uint32_t refs[8]; bool leaf[8];
            const int cnt = pt::dp_cover(KD.data(), kn.data(), b2, refs, leaf);
            bool same = cnt == (int)ch.size();
            for (int c = 0; same && c < cnt; ++c) same = refs[c] == ch[c].ref && leaf[c] == ch[c].leaf;
            if (!same) cover_diff++;
            ch.clear();
            for (int c = 0; c < cnt; ++c) { CC x; x.ref = refs[c]; x.count = ref_count(t, refs[c]); x.b = ref_box(t, refs[c]); x.leaf = leaf[c]; ch.push_back(x); }
        }
        emit_node(t, w, widx, ch, todo);
    }
    if (kernel_table) std::printf("dp_cover vs the lab's recursion: %zu differing wide nodes of %zu\n", cover_diff, w.nodes.size());
    return w;
}

// ------------------------------------------------------------------ rays + walker
struct Ray { float o[3], d[3], tmin, tmax; };
struct HitR { float t; uint32_t tri; };

static const float* g_tri_ptr = nullptr;     // triangles the walker tests (g_tris unless a BLAS of the two-level model is walked)

static bool tri_hit(uint32_t tr, const Ray& r, float tmax, float& t_out) {
    const float* v = (g_tri_ptr ? g_tri_ptr : g_tris.data()) + 9 * (size_t)tr;
    double e1[3], e2[3], n[3];
    for (int k = 0; k < 3; ++k) { e1[k] = (double)v[3 + k] - v[k]; e2[k] = (double)v[6 + k] - v[k]; }
    n[0] = e1[1] * e2[2] - e1[2] * e2[1]; n[1] = e1[2] * e2[0] - e1[0] * e2[2]; n[2] = e1[0] * e2[1] - e1[1] * e2[0];
    const double nd = n[0] * r.d[0] + n[1] * r.d[1] + n[2] * r.d[2];
    if (!(nd < 0.0)) return false;     // front faces only (CullBackFacingTriangles)
    double pv[3], q[3], s[3];
    pv[0] = r.d[1] * e2[2] - r.d[2] * e2[1]; pv[1] = r.d[2] * e2[0] - r.d[0] * e2[2]; pv[2] = r.d[0] * e2[1] - r.d[1] * e2[0];
    const double det = e1[0] * pv[0] + e1[1] * pv[1] + e1[2] * pv[2];
    if (det == 0.0) return false;
    const double inv = 1.0 / det;
    for (int k = 0; k < 3; ++k) s[k] = (double)r.o[k] - v[k];
    const double u = (s[0] * pv[0] + s[1] * pv[1] + s[2] * pv[2]) * inv;
    if (u < 0.0 || u > 1.0) return false;
    q[0] = s[1] * e1[2] - s[2] * e1[1]; q[1] = s[2] * e1[0] - s[0] * e1[2]; q[2] = s[0] * e1[1] - s[1] * e1[0];
    const double vv = (r.d[0] * q[0] + r.d[1] * q[1] + r.d[2] * q[2]) * inv;
    if (vv < 0.0 || u + vv > 1.0) return false;
    const double tt = (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]) * inv;
    if (!(tt > r.tmin && tt < tmax)) return false;
    t_out = (float)tt;
    return true;
}

struct Counts { uint64_t nodes = 0, tris = 0, rays = 0, big_tris = 0; };
static float g_big_area = 3.0e38f;   // triangle tests on leaves with a larger box half-area are counted separately

// diagnostics (what-if bounds, not buildable variants): exact child boxes instead of the 8-bit grid; inner children
// visited by true entry distance instead of octant priority; a popped node skipped when the closest hit found since
// it was pushed already lies in front of its entry point (needs the entry distance on the stack)
static bool g_exact_boxes = false, g_sort_by_distance = false, g_cull_on_pop = false;
static bool g_any_hit = false;   // stop at the first intersection found (what an occlusion query would cost)

static HitR walk(const Wide& w, const Ray& r, Counts& cnt) {
    HitR best{1e32f, 0xFFFFFFFFu};
    float id[3]; uint32_t rs = 0;
    for (int k = 0; k < 3; ++k) { id[k] = 1.0f / (r.d[k] == 0.0f ? 1e-30f : r.d[k]); if (id[k] >= 0.0f) rs |= 1u << k; }
    int stack[512]; float stack_t[512]; int sp = 0;
    stack_t[sp] = 0.0f; stack[sp++] = 0;
    cnt.rays++;
    while (sp) {
        --sp;
        if (g_cull_on_pop && stack_t[sp] > best.t) continue;
        const WNode& n = w.nodes[stack[sp]];
        cnt.nodes++;
        const float tlim = std::fmin(r.tmax, best.t);
        struct { int prio, node; float tn; } inner[8]; int ni = 0;
        uint32_t leaf_hit = 0;
        for (int c = 0; c < n.nc; ++c) {
            const Box& b = g_exact_boxes ? n.c[c].xb : n.c[c].qb;
            float tn = r.tmin, tf = tlim;
            for (int k = 0; k < 3; ++k) {
                float a = (b.lo[k] - r.o[k]) * id[k], bb = (b.hi[k] - r.o[k]) * id[k];
                tn = std::fmax(tn, std::fmin(a, bb)); tf = std::fmin(tf, std::fmax(a, bb));
            }
            if (!(tn <= tf * 1.0000019f)) continue;
            if (n.c[c].inner >= 0) {
                inner[ni].prio = g_sort_by_distance ? -(int)(std::fmin(tn, 1.0e6f) * 1024.0f) : (int)((uint32_t)n.c[c].slot ^ rs);
                inner[ni].node = n.c[c].inner; inner[ni].tn = tn; ni++;
            }
            else leaf_hit |= 1u << c;
        }
        // the node's hit triangles first (all of them, against the ray's own tmax like test_triangle)
        for (int c = 0; c < n.nc; ++c) {
            if (!(leaf_hit >> c & 1u)) continue;
            for (uint32_t k = 0; k < n.c[c].tri_cnt; ++k) {
                cnt.tris++;
                if (g_big_area < 3.0e38f && g_leaf[w.tris[n.c[c].tri_first + k]].half_area() > g_big_area) cnt.big_tris++;
                float tt;
                const uint32_t tr = w.tris[n.c[c].tri_first + k];
                if (tri_hit(tr, r, r.tmax, tt) && (tt < best.t || (tt == best.t && tr < best.tri))) { best.t = tt; best.tri = tr; }
            }
        }
        if (g_any_hit && best.tri != 0xFFFFFFFFu) break;
        // inner children: highest priority is visited first, so it is pushed last
        for (int i = 1; i < ni; ++i) { auto e = inner[i]; int j = i - 1; while (j >= 0 && inner[j].prio > e.prio) { inner[j + 1] = inner[j]; --j; } inner[j + 1] = e; }
        for (int k = 0; k < ni; ++k) { stack_t[sp] = inner[k].tn; stack[sp++] = inner[k].node; }
    }
    return best;
}

static uint32_t hash32(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

// ------------------------------------------------------------------ two-level model (instanced scenes, C4)
// scene file: u32 M; per mesh u32 n + n x 9 floats; u32 I; per instance u32 mesh + 12 floats (rows of the 3x4 transform)
struct Inst { uint32_t mesh; float m[12]; double inv[9]; };
struct Mesh { std::vector<float> tris; Wide shipped, better; };

static void set_leaves(const std::vector<float>& tris) {
    g_tris = tris;
    g_n = (uint32_t)(tris.size() / 9);
    g_leaf.assign(g_n, Box());
    for (uint32_t i = 0; i < g_n; ++i)
        for (int v = 0; v < 3; ++v)
            for (int k = 0; k < 3; ++k) {
                g_leaf[i].lo[k] = std::fmin(g_leaf[i].lo[k], g_tris[9 * (size_t)i + 3 * v + k]);
                g_leaf[i].hi[k] = std::fmax(g_leaf[i].hi[k], g_tris[9 * (size_t)i + 3 * v + k]);
            }
}

static Wide build_variant(bool better, uint32_t max_leaf) {
    if (g_n < 2) {   // a single leaf: one wide node with one slot
        Wide w; WNode n; n.nc = 1; n.c[0].qb = n.c[0].xb = g_leaf[0]; n.c[0].inner = -1; n.c[0].tri_first = 0; n.c[0].tri_cnt = 1; n.c[0].slot = 0;
        w.nodes.push_back(n); w.tris.push_back(0); return w;
    }
    if (!better) { Bvh2 t = build_lbvh(); return collapse_greedy(t, max_leaf); }
    uint32_t levels = 0;
    Bvh2 t = build_sah_kernel(&levels);
    return collapse_optimal(t, max_leaf, 1.0f, 0.7f, true);
}

struct Counts2 { uint64_t nodes = 0, tris = 0, entries = 0, rays = 0; };

static HitR walk2(const Wide& tlas, const std::vector<Mesh>& meshes, const std::vector<Inst>& inst, bool better, const Ray& r, Counts2& c2) {
    HitR best{1e32f, 0xFFFFFFFFu};
    float id[3]; uint32_t rs = 0;
    for (int k = 0; k < 3; ++k) { id[k] = 1.0f / (r.d[k] == 0.0f ? 1e-30f : r.d[k]); if (id[k] >= 0.0f) rs |= 1u << k; }
    int stack[512]; int sp = 0;
    stack[sp++] = 0;
    c2.rays++;
    while (sp) {
        const WNode& n = tlas.nodes[stack[--sp]];
        c2.nodes++;
        const float tlim = std::fmin(r.tmax, best.t);
        struct { int prio, node; } inner[8]; int ni = 0;
        uint32_t leaf_hit = 0;
        for (int c = 0; c < n.nc; ++c) {
            const Box& b = n.c[c].qb;
            float tn = r.tmin, tf = tlim;
            for (int k = 0; k < 3; ++k) {
                float a = (b.lo[k] - r.o[k]) * id[k], bb = (b.hi[k] - r.o[k]) * id[k];
                tn = std::fmax(tn, std::fmin(a, bb)); tf = std::fmin(tf, std::fmax(a, bb));
            }
            if (!(tn <= tf * 1.0000019f)) continue;
            if (n.c[c].inner >= 0) { inner[ni].prio = (int)((uint32_t)n.c[c].slot ^ rs); inner[ni].node = n.c[c].inner; ni++; }
            else leaf_hit |= 1u << c;
        }
        for (int c = 0; c < n.nc; ++c) {
            if (!(leaf_hit >> c & 1u)) continue;
            const Inst& in = inst[tlas.tris[n.c[c].tri_first]];
            c2.entries++;
            Ray o;
            o.tmin = r.tmin; o.tmax = std::fmin(r.tmax, best.t);
            const double q[3] = {(double)r.o[0] - in.m[3], (double)r.o[1] - in.m[7], (double)r.o[2] - in.m[11]};
            for (int k = 0; k < 3; ++k) {
                o.o[k] = (float)(in.inv[3 * k] * q[0] + in.inv[3 * k + 1] * q[1] + in.inv[3 * k + 2] * q[2]);
                o.d[k] = (float)(in.inv[3 * k] * r.d[0] + in.inv[3 * k + 1] * r.d[1] + in.inv[3 * k + 2] * r.d[2]);
            }
            const Mesh& m = meshes[in.mesh];
            g_tri_ptr = m.tris.data();
            Counts cb;
            const HitR h = walk(better ? m.better : m.shipped, o, cb);
            g_tri_ptr = nullptr;
            c2.nodes += cb.nodes; c2.tris += cb.tris;
            if (h.tri != 0xFFFFFFFFu && h.t < best.t) { best.t = h.t; best.tri = h.tri; }
        }
        for (int i = 1; i < ni; ++i) { auto e = inner[i]; int j = i - 1; while (j >= 0 && inner[j].prio > e.prio) { inner[j + 1] = inner[j]; --j; } inner[j + 1] = e; }
        for (int k = 0; k < ni; ++k) stack[sp++] = inner[k].node;
    }
    return best;
}

static int scene_mode(const char* path, int W, int H, int step) {
    FILE* fp = std::fopen(path, "rb");
    if (!fp) { std::perror(path); return 1; }
    uint32_t M = 0, I = 0;
    if (std::fread(&M, 4, 1, fp) != 1) return 1;
    std::vector<Mesh> meshes(M);
    for (auto& m : meshes) {
        uint32_t n = 0;
        if (std::fread(&n, 4, 1, fp) != 1) return 1;
        m.tris.resize(9 * (size_t)n);
        if (std::fread(m.tris.data(), 36, n, fp) != n) return 1;
    }
    if (std::fread(&I, 4, 1, fp) != 1) return 1;
    std::vector<Inst> inst(I);
    for (auto& in : inst) {
        if (std::fread(&in.mesh, 4, 1, fp) != 1 || std::fread(in.m, 4, 12, fp) != 12) return 1;
        const double a[9] = {in.m[0], in.m[1], in.m[2], in.m[4], in.m[5], in.m[6], in.m[8], in.m[9], in.m[10]};
        const double det = a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
        const double c[9] = {a[4] * a[8] - a[5] * a[7], a[2] * a[7] - a[1] * a[8], a[1] * a[5] - a[2] * a[4],
                             a[5] * a[6] - a[3] * a[8], a[0] * a[8] - a[2] * a[6], a[2] * a[3] - a[0] * a[5],
                             a[3] * a[7] - a[4] * a[6], a[1] * a[6] - a[0] * a[7], a[0] * a[4] - a[1] * a[3]};
        for (int k = 0; k < 9; ++k) in.inv[k] = c[k] / det;
    }
    std::fclose(fp);
    std::printf("%u meshes, %u instances, %dx%d every %d pixels\n", M, I, W, H, step);
    for (auto& m : meshes) { set_leaves(m.tris); m.shipped = build_variant(false, 2); m.better = build_variant(true, 2); }
    // world-space triangles of every instance: the flattened scene, and the exact instance boxes (as the shipped TLAS uses)
    std::vector<float> flat;
    std::vector<float> boxes_as_tris;     // one degenerate "triangle" (lo, hi, lo) per instance: its box is the instance box
    for (auto& in : inst) {
        const std::vector<float>& t = meshes[in.mesh].tris;
        Box b;
        for (size_t v = 0; v < t.size() / 3; ++v) {
            float w[3];
            for (int r = 0; r < 3; ++r) w[r] = in.m[4 * r] * t[3 * v] + in.m[4 * r + 1] * t[3 * v + 1] + in.m[4 * r + 2] * t[3 * v + 2] + in.m[4 * r + 3];
            for (int k = 0; k < 3; ++k) { flat.push_back(w[k]); b.lo[k] = std::fmin(b.lo[k], w[k]); b.hi[k] = std::fmax(b.hi[k], w[k]); }
        }
        for (int k = 0; k < 3; ++k) boxes_as_tris.push_back(b.lo[k]);
        for (int k = 0; k < 3; ++k) boxes_as_tris.push_back(b.hi[k]);
        for (int k = 0; k < 3; ++k) boxes_as_tris.push_back(b.lo[k]);
    }
    set_leaves(boxes_as_tris);
    const Wide tlas_shipped = build_variant(false, 1), tlas_better = build_variant(true, 1);
    std::printf("TLAS: %zu / %zu wide nodes (shipped / SAH + optimal)\n", tlas_shipped.nodes.size(), tlas_better.nodes.size());

    std::vector<Ray> primary, bounce;
    const float vh = 2.0f * std::tan(0.5f * 0.78539816339f), vw = vh * (float)W / (float)H;
    for (int y = step / 2; y < H; y += step)
        for (int x = step / 2; x < W; x += step) {
            const float dx = 2.0f * ((float)x + 0.5f) / (float)W - 1.0f, dy = 2.0f * ((float)y + 0.5f) / (float)H - 1.0f;
            float d[3] = {-vw * dx, -vh * dy, 1.0f};
            const float inv = 1.0f / std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
            Ray r; r.tmin = 0.001f; r.tmax = 10000.0f;
            for (int k = 0; k < 3; ++k) { r.o[k] = 0.0f; r.d[k] = d[k] * inv; }
            primary.push_back(r);
        }
    {
        // bounce rays: uniform directions in the hemisphere facing back along the incoming ray (no normals at hand here)
        Counts2 dummy; uint32_t seed = 1;
        for (auto& r : primary) {
            const HitR h = walk2(tlas_shipped, meshes, inst, false, r, dummy);
            if (h.tri == 0xFFFFFFFFu) continue;
            Ray b; b.tmin = 0.001f; b.tmax = 10000.0f;
            float d[3], len2;
            do {
                for (int k = 0; k < 3; ++k) d[k] = 2.0f * (float)(hash32(seed++) >> 8) / 16777216.0f - 1.0f;
                len2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
            } while (len2 > 1.0f || len2 < 1e-4f);
            const float il = 1.0f / std::sqrt(len2);
            float dot = 0.0f;
            for (int k = 0; k < 3; ++k) { d[k] *= il; dot += d[k] * r.d[k]; }
            for (int k = 0; k < 3; ++k) { if (dot > 0.0f) d[k] = -d[k]; b.d[k] = d[k]; b.o[k] = r.o[k] + (h.t - 1e-3f) * r.d[k]; }
            bounce.push_back(b);
        }
    }
    std::printf("%zu primary rays, %zu bounce rays\n\n", primary.size(), bounce.size());
    std::printf("%-52s %9s %7s %8s %8s\n", "variant", "nodes/ray", "tris", "entries", "n+.7t");
    std::vector<float> ref_t;
    for (int v = 0; v < 2; ++v) {
        Counts2 c;
        size_t idx = 0, mism = 0;
        for (auto* rays : {&primary, &bounce})
            for (auto& r : *rays) {
                const HitR h = walk2(v ? tlas_better : tlas_shipped, meshes, inst, v != 0, r, c);
                if (v == 0) ref_t.push_back(h.t); else if (std::fabs(h.t - ref_t[idx]) > 1e-3f * std::fmax(1.0f, ref_t[idx]) && !(h.t > 1e31f && ref_t[idx] > 1e31f)) mism++;
                idx++;
            }
        const double n = (double)c.nodes / c.rays, t = (double)c.tris / c.rays;
        std::printf("%-52s %9.2f %7.2f %8.2f %8.2f%s\n", v ? "two-level, SAH + optimal collapse on both levels" : "two-level as shipped (LBVH + greedy, exact boxes)",
                    n, t, (double)c.entries / c.rays, n + 0.7 * t, mism ? "  HIT MISMATCH" : "");
    }
    // flattened: one hierarchy over the world-space triangles of all instances
    set_leaves(flat);
    std::printf("flattened scene: %u world-space triangles\n", g_n);
    for (int v = 0; v < 2; ++v) {
        const Wide w = build_variant(v != 0, 2);
        Counts c;
        size_t idx = 0, mism = 0;
        for (auto* rays : {&primary, &bounce})
            for (auto& r : *rays) {
                const HitR h = walk(w, r, c);
                if (std::fabs(h.t - ref_t[idx]) > 1e-3f * std::fmax(1.0f, ref_t[idx]) && !(h.t > 1e31f && ref_t[idx] > 1e31f)) mism++;
                idx++;
            }
        const double n = (double)c.nodes / c.rays, t = (double)c.tris / c.rays;
        std::printf("%-52s %9.2f %7.2f %8s %8.2f   (%zu wide nodes)%s\n", v ? "flattened, SAH + optimal collapse" : "flattened, LBVH + greedy collapse", n, t, "-", n + 0.7 * t,
                    w.nodes.size(), mism ? "  HIT MISMATCH" : "");
        if (mism) std::printf("   (%zu rays)\n", mism);
    }
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 5) { std::fprintf(stderr, "usage: bvh_lab tris.bin|scene.scene width height step\n"); return 2; }
    if (std::strstr(argv[1], ".scene")) return scene_mode(argv[1], std::atoi(argv[2]), std::atoi(argv[3]), std::atoi(argv[4]));
    FILE* fp = std::fopen(argv[1], "rb");
    if (!fp) { std::perror(argv[1]); return 1; }
    if (std::fread(&g_n, 4, 1, fp) != 1) return 1;
    g_tris.resize(9 * (size_t)g_n);
    if (std::fread(g_tris.data(), 36, g_n, fp) != g_n) return 1;
    std::fclose(fp);
    const int W = std::atoi(argv[2]), H = std::atoi(argv[3]), step = std::atoi(argv[4]);
    g_leaf.resize(g_n);
    for (uint32_t i = 0; i < g_n; ++i)
        for (int v = 0; v < 3; ++v)
            for (int k = 0; k < 3; ++k) {
                g_leaf[i].lo[k] = std::fmin(g_leaf[i].lo[k], g_tris[9 * (size_t)i + 3 * v + k]);
                g_leaf[i].hi[k] = std::fmax(g_leaf[i].hi[k], g_tris[9 * (size_t)i + 3 * v + k]);
            }
    std::printf("triangles %u, %dx%d every %d pixels\n", g_n, W, H, step);

    // rays: the reference camera at its defaults (Camera.zig:86-102, ray_gen.rgen:34-48); bounce rays are
    // cosine-weighted about the geometric normal from the primary hits (a stand-in for the BSDF's directions)
    std::vector<Ray> primary;
    const float vh = 2.0f * std::tan(0.5f * 0.78539816339f), vw = vh * (float)W / (float)H;
    for (int y = step / 2; y < H; y += step)
        for (int x = step / 2; x < W; x += step) {
            const float dx = 2.0f * ((float)x + 0.5f) / (float)W - 1.0f, dy = 2.0f * ((float)y + 0.5f) / (float)H - 1.0f;
            float d[3] = {-vw * dx, -vh * dy, 1.0f};
            const float inv = 1.0f / std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
            Ray r; r.tmin = 0.001f; r.tmax = 10000.0f;
            for (int k = 0; k < 3; ++k) { r.o[k] = 0.0f; r.d[k] = d[k] * inv; }
            primary.push_back(r);
        }

    struct Variant { const char* name; Wide w; };
    std::vector<Variant> vs;
    Bvh2 lb = build_lbvh();
    std::printf("LBVH built (%zu internal nodes)\n", lb.nodes.size());
    Bvh2 sah = build_sah(32);
    std::printf("binned SAH built (%zu internal nodes)\n", sah.nodes.size());
    vs.push_back({"LBVH + greedy, 2 tris/slot (shipped)", collapse_greedy(lb, 2)});
    vs.push_back({"LBVH + greedy, 3 tris/slot", collapse_greedy(lb, 3)});
    vs.push_back({"LBVH + optimal, 2 tris/slot", collapse_optimal(lb, 2, 1.0f, 0.7f)});
    vs.push_back({"LBVH + optimal, 3 tris/slot", collapse_optimal(lb, 3, 1.0f, 0.7f)});
    vs.push_back({"SAH32 + greedy, 2 tris/slot", collapse_greedy(sah, 2)});
    vs.push_back({"SAH32 + optimal, 1 tri/slot", collapse_optimal(sah, 1, 1.0f, 0.7f)});
    vs.push_back({"SAH32 + optimal, 2 tris/slot", collapse_optimal(sah, 2, 1.0f, 0.7f)});
    vs.push_back({"SAH32 + optimal, 3 tris/slot", collapse_optimal(sah, 3, 1.0f, 0.7f)});
    {
        Bvh2 sah64 = build_sah(64), sah128 = build_sah(128);
        vs.push_back({"SAH64 + optimal, 2 tris/slot", collapse_optimal(sah64, 2, 1.0f, 0.7f)});
        vs.push_back({"SAH128 + optimal, 2 tris/slot", collapse_optimal(sah128, 2, 1.0f, 0.7f)});
    }
    Bvh2 sah16 = build_sah(16);
    vs.push_back({"SAH16 + optimal, 2 tris/slot", collapse_optimal(sah16, 2, 1.0f, 0.7f)});
    uint32_t levels = 0;
    Bvh2 ks = build_sah_kernel(&levels);
    std::printf("csrc/bvh_sah.cuh (serial model): %u levels\n", levels);
    vs.push_back({"bvh_sah.cuh + greedy, 2 tris/slot", collapse_greedy(ks, 2)});
    vs.push_back({"bvh_sah.cuh + k_collapse_dp, 2 tris/slot", collapse_optimal(ks, 2, 1.0f, 0.7f, true)});
    vs.push_back({"bvh_sah.cuh + k_collapse_dp, 3 tris/slot", collapse_optimal(ks, 3, 1.0f, 0.7f, true)});
    vs.push_back({"LBVH + k_collapse_dp, 2 tris/slot", collapse_optimal(lb, 2, 1.0f, 0.7f, true)});
    vs.push_back({"bvh_sah.cuh + dp, c_tri 0.4", collapse_optimal(ks, 2, 1.0f, 0.4f)});
    vs.push_back({"bvh_sah.cuh + dp, c_tri 1.0", collapse_optimal(ks, 2, 1.0f, 1.0f)});
    vs.push_back({"bvh_sah.cuh + dp, c_tri 1.5", collapse_optimal(ks, 2, 1.0f, 1.5f)});
    g_optimal_slots = true;
    vs.push_back({"bvh_sah.cuh + dp, exact slot assignment", collapse_optimal(ks, 2, 1.0f, 0.7f)});
    vs.push_back({"LBVH + greedy, exact slot assignment", collapse_greedy(lb, 2)});
    g_optimal_slots = false;

    // bounce rays from the first variant's primary hits (the closest hit does not depend on the hierarchy)
    std::vector<Ray> bounce;
    {
        Counts dummy;
        uint32_t seed = 1;
        for (auto& r : primary) {
            HitR h = walk(vs[0].w, r, dummy);
            if (h.tri == 0xFFFFFFFFu) continue;
            const float* v = &g_tris[9 * (size_t)h.tri];
            float e1[3], e2[3], n[3];
            for (int k = 0; k < 3; ++k) { e1[k] = v[3 + k] - v[k]; e2[k] = v[6 + k] - v[k]; }
            n[0] = e1[1] * e2[2] - e1[2] * e2[1]; n[1] = e1[2] * e2[0] - e1[0] * e2[2]; n[2] = e1[0] * e2[1] - e1[1] * e2[0];
            const float il = 1.0f / std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
            for (int k = 0; k < 3; ++k) n[k] *= il;
            const float u1 = (float)(hash32(seed++) >> 8) / 16777216.0f, u2 = (float)(hash32(seed++) >> 8) / 16777216.0f;
            const float rr = std::sqrt(u1), ph = 6.2831853f * u2;
            const float lx = rr * std::cos(ph), ly = rr * std::sin(ph), lz = std::sqrt(std::fmax(0.0f, 1.0f - u1));
            float a[3] = {std::fabs(n[0]) < 0.9f ? 1.0f : 0.0f, std::fabs(n[0]) < 0.9f ? 0.0f : 1.0f, 0.0f}, tx[3], ty[3];
            tx[0] = a[1] * n[2] - a[2] * n[1]; tx[1] = a[2] * n[0] - a[0] * n[2]; tx[2] = a[0] * n[1] - a[1] * n[0];
            const float it = 1.0f / std::sqrt(tx[0] * tx[0] + tx[1] * tx[1] + tx[2] * tx[2]);
            for (int k = 0; k < 3; ++k) tx[k] *= it;
            ty[0] = n[1] * tx[2] - n[2] * tx[1]; ty[1] = n[2] * tx[0] - n[0] * tx[2]; ty[2] = n[0] * tx[1] - n[1] * tx[0];
            Ray b; b.tmin = 0.001f; b.tmax = 10000.0f;
            for (int k = 0; k < 3; ++k) { b.o[k] = r.o[k] + h.t * r.d[k] + n[k] * 1e-5f; b.d[k] = lx * tx[k] + ly * ty[k] + lz * n[k]; }
            bounce.push_back(b);
        }
    }
    std::printf("%zu primary rays, %zu bounce rays\n\n", primary.size(), bounce.size());
    std::printf("%-42s %8s %6s %10s | %7s %6s | %7s %6s | %7s %6s %7s\n", "variant", "nodes", "fill", "SAH", "prim n", "tri", "bnc n", "tri", "all n", "tri", "n+.7t");
    std::vector<HitR> ref_hits;
    for (size_t vi = 0; vi < vs.size(); ++vi) {
        const Wide& w = vs[vi].w;
        uint64_t slots = 0;
        for (auto& n : w.nodes) slots += n.nc;
        Counts cp, cb;
        size_t mism = 0, idx = 0;
        for (auto& r : primary) { HitR h = walk(w, r, cp); if (vi == 0) ref_hits.push_back(h); else if (h.tri != ref_hits[idx].tri) mism++; idx++; }
        for (auto& r : bounce) { HitR h = walk(w, r, cb); if (vi == 0) ref_hits.push_back(h); else if (h.tri != ref_hits[idx].tri) mism++; idx++; }
        const double an = (double)(cp.nodes + cb.nodes) / (double)(cp.rays + cb.rays), at = (double)(cp.tris + cb.tris) / (double)(cp.rays + cb.rays);
        std::printf("%-42s %8zu %6.2f %10.1f | %7.2f %6.2f | %7.2f %6.2f | %7.2f %6.2f %7.2f%s\n", vs[vi].name, w.nodes.size(),
                    (double)slots / (double)w.nodes.size(), w.sah / (double)vs[vi].w.nodes.size() * 0.0 + w.sah / (double)lb.nodes[0].b.half_area(),
                    (double)cp.nodes / cp.rays, (double)cp.tris / cp.rays, (double)cb.nodes / std::max<uint64_t>(1, cb.rays),
                    (double)cb.tris / std::max<uint64_t>(1, cb.rays), an, at, an + 0.7 * at, mism ? "  HIT MISMATCH" : "");
        if (mism) std::printf("   (%zu rays with a different closest hit)\n", mism);
    }
    // what-if bounds on the last SAH + optimal variant
    {
        const Wide* w = nullptr; const char* nm = nullptr;
        for (auto& v : vs) if (std::strstr(v.name, "bvh_sah.cuh + k_collapse_dp, 2")) { w = &v.w; nm = v.name; }
        if (w) {
            std::printf("\nwhat-if bounds on '%s' (not buildable variants):\n", nm);
            {
                // how many triangle tests go to the few large triangles (candidates for pre-splitting)
                std::vector<float> areas(g_n);
                for (uint32_t i = 0; i < g_n; ++i) areas[i] = g_leaf[i].half_area();
                std::vector<float> sorted = areas;
                std::sort(sorted.begin(), sorted.end());
                for (double q : {0.90, 0.99, 0.999}) {
                    g_big_area = sorted[(size_t)(q * (g_n - 1))];
                    Counts c;
                    for (auto& r : primary) walk(*w, r, c);
                    for (auto& r : bounce) walk(*w, r, c);
                    std::printf("  triangle tests on the largest %.1f %% of triangles (box half-area > %.3g): %.2f of %.2f per ray\n", 100.0 * (1.0 - q),
                                g_big_area, (double)c.big_tris / c.rays, (double)c.tris / c.rays);
                }
                g_big_area = 3.0e38f;
            }
            for (int mode = 0; mode < 4; ++mode) {
                g_exact_boxes = mode == 1; g_sort_by_distance = mode == 2; g_cull_on_pop = mode == 3;
                Counts c;
                for (auto& r : primary) walk(*w, r, c);
                for (auto& r : bounce) walk(*w, r, c);
                const double an = (double)c.nodes / c.rays, at = (double)c.tris / c.rays;
                std::printf("  %-52s %7.2f %6.2f %7.2f\n", mode == 0 ? "as shipped (8-bit boxes, octant order)" : mode == 1 ? "exact child boxes" :
                            mode == 2 ? "inner children by entry distance" : "popped node skipped when behind the closest hit", an, at, an + 0.7 * at);
            }
            g_exact_boxes = g_sort_by_distance = g_cull_on_pop = false;
            {
                // the last segment of a sample only adds the emission of what it hits (ray_gen.rgen:75-77) or the sky: with
                // few emissive triangles it is an occlusion query plus a test of those triangles
                Counts c0, c1;
                for (auto& r : bounce) walk(*w, r, c0);
                g_any_hit = true;
                for (auto& r : bounce) walk(*w, r, c1);
                g_any_hit = false;
                std::printf("  bounce rays as closest-hit queries: %.2f nodes %.2f tris (%.2f units); as first-hit queries: %.2f nodes %.2f tris (%.2f units)\n",
                            (double)c0.nodes / c0.rays, (double)c0.tris / c0.rays, ((double)c0.nodes + 0.7 * c0.tris) / c0.rays,
                            (double)c1.nodes / c1.rays, (double)c1.tris / c1.rays, ((double)c1.nodes + 0.7 * c1.tris) / c1.rays);
            }
        }
    }
    return 0;
}
