// oracle.cpp — CPU ORACLE for the trace loop of antlilja/vulkan-pathtracer.
//
// TEST INFRASTRUCTURE ONLY.  Nothing in the product (libb200pt.so, the CLI, the python host
// package) may include, link or call this file; it is used by tests/, by
// __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs as the checker
// and the timed CPU baseline.
//
// PARITY STATUS: the reference ships no tests, golden vectors or fixtures for this path
// (SURVEY.md §4, §8c), and its BVH build / traversal / ray-triangle test live inside the Vulkan
// driver.  Pinned against the reference's own code: prng_seed, pcg2d and every function of
// brdf.glsl — oracle/ref_glsl/ compiles src/shaders/rand.glsl and brdf.glsl from /root/reference as
// C++ (oracle/_ref/libref_glsl.so) and tests/test_oracle_vs_ref_glsl.py compares them with this
// restatement.  Everything else (closest-hit rule, attribute fetch, material fetch, texture
// filtering, the loop in ray_gen.rgen) is restated by hand from the cited lines: PARITY UNPINNED.
//
// Arithmetic contract shared with the CUDA kernels (DESIGN.md §"Arithmetic contract"): IEEE
// binary32, no FMA contraction (-ffp-contract=off), expressions evaluated left to right exactly
// as written here, sqrt and / correctly rounded, sin/cos = correctly rounded (computed in double,
// rounded to float), min/max with GLSL semantics (min(x,y) = y<x ? y : x).
//
// Each function cites the reference file:line it follows (paths relative to /root/reference).

#include "../include/b200pt.h"
#include "../include/b200pt_contract.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace orc {

// ------------------------------------------------------------------------------------------
// small vector algebra (GLSL semantics, explicit evaluation order)
// ------------------------------------------------------------------------------------------
struct V2 { float x, y; };
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };
struct U2 { uint32_t x, y; };

static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 v3(float s) { return V3{s, s, s}; }
static inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
static inline V3 operator*(float s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }
static inline V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
static inline V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
static inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) {
    return V3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
// normalize(v) = v * (1 / sqrt(dot(v, v))): one correctly rounded division (arithmetic contract)
static inline V3 normalize(V3 a) { return a * (1.0f / std::sqrt(dot(a, a))); }
static inline float gmin(float x, float y) { return (y < x) ? y : x; }
static inline float gmax(float x, float y) { return (x < y) ? y : x; }
static inline float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
static inline V3 mix(V3 a, V3 b, float t) { return V3{mixf(a.x, b.x, t), mixf(a.y, b.y, t), mixf(a.z, b.z, t)}; }
static inline float idx(V3 a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }

// fp32 sin / cos of the arithmetic contract (include/b200pt_contract.h: explicit double arithmetic, no library call)
static inline float sin_cr(float x) { float s, c; pt_contract_sincos(x, &s, &c); return s; }
static inline float cos_cr(float x) { float s, c; pt_contract_sincos(x, &s, &c); return c; }

static const float kPi = 3.141592653589793238462643f;  // brdf.glsl:4 (GLSL literal = fp32)

// 3x3 matrix, row-major: m[r][c]
struct M3 { float m[3][3]; };

static inline V3 mul(const M3& a, V3 v) {
    return V3{a.m[0][0] * v.x + a.m[0][1] * v.y + a.m[0][2] * v.z,
              a.m[1][0] * v.x + a.m[1][1] * v.y + a.m[1][2] * v.z,
              a.m[2][0] * v.x + a.m[2][1] * v.y + a.m[2][2] * v.z};
}

// inverse(mat3) by cofactors / determinant (the definition both sides use).
static inline M3 inverse(const M3& a) {
    const float a00 = a.m[0][0], a01 = a.m[0][1], a02 = a.m[0][2];
    const float a10 = a.m[1][0], a11 = a.m[1][1], a12 = a.m[1][2];
    const float a20 = a.m[2][0], a21 = a.m[2][1], a22 = a.m[2][2];
    const float c00 = a11 * a22 - a12 * a21;
    const float c01 = a12 * a20 - a10 * a22;
    const float c02 = a10 * a21 - a11 * a20;
    const float det = a00 * c00 + a01 * c01 + a02 * c02;
    const float id = 1.0f / det;
    M3 r;
    r.m[0][0] = c00 * id;
    r.m[0][1] = (a02 * a21 - a01 * a22) * id;
    r.m[0][2] = (a01 * a12 - a02 * a11) * id;
    r.m[1][0] = c01 * id;
    r.m[1][1] = (a00 * a22 - a02 * a20) * id;
    r.m[1][2] = (a02 * a10 - a00 * a12) * id;
    r.m[2][0] = c02 * id;
    r.m[2][1] = (a01 * a20 - a00 * a21) * id;
    r.m[2][2] = (a00 * a11 - a01 * a10) * id;
    return r;
}

// ------------------------------------------------------------------------------------------
// RNG — src/shaders/rand.glsl
// ------------------------------------------------------------------------------------------
// rand.glsl:89-94
static inline U2 prng_seed(uint32_t px, uint32_t py, uint32_t frame) {
    U2 s{px, py};
    s.x = s.x * 0x9e3779b9u + frame;
    s.y = s.y * 0x9e3779b9u + frame;
    s.x ^= s.x >> 16;
    s.y ^= s.y >> 16;
    s.x *= 0x9e3779b9u;
    s.y *= 0x9e3779b9u;
    return s;
}

// rand.glsl:76-87.  The literal 2.32830643654e-10 rounds to exactly 2^-32 in fp32.
static inline V2 pcg2d(U2& seed) {
    seed.x = 1664525u * seed.x + 1013904223u;
    seed.y = 1664525u * seed.y + 1013904223u;
    seed.x += 1664525u * seed.y;
    seed.y += 1664525u * seed.x;
    seed.x ^= seed.x >> 16;
    seed.y ^= seed.y >> 16;
    seed.x += 1664525u * seed.y;
    seed.y += 1664525u * seed.x;
    seed.x ^= seed.x >> 16;
    seed.y ^= seed.y >> 16;
    return V2{(float)seed.x * 2.32830643654e-10f, (float)seed.y * 2.32830643654e-10f};
}

// ------------------------------------------------------------------------------------------
// BSDF — src/shaders/brdf.glsl
// ------------------------------------------------------------------------------------------
struct ShadingInfo {  // brdf.glsl:6-13
    V3 normal;
    V3 out_dir;
    float lambert_out;
    V3 diffuse_albedo;
    V3 fresnel_0;
    float roughness;
};

// brdf.glsl:15-20
static inline V3 fresnel_schlick(V3 f0, V3 f90, float cosine_theta) {
    float factor = 1.0f - cosine_theta;
    float factor_squared = factor * factor;
    float factor_fifth = factor_squared * factor_squared * factor;
    return mix(f0, f90, factor_fifth);
}

// brdf.glsl:22-50
static inline V3 brdf(const ShadingInfo& s, V3 light_dir) {
    float n_dot_light = dot(s.normal, light_dir);
    float n_dot_view = s.lambert_out;
    if (gmin(n_dot_light, n_dot_view) < 0.0f) return v3(0.0f);

    V3 half_vector = normalize(light_dir + s.out_dir);
    float half_dot_view = dot(half_vector, s.out_dir);

    float f90 = (half_dot_view * half_dot_view) * (2.0f * s.roughness) + 0.5f;
    float diffuse_fresnel = fresnel_schlick(v3(1.0f), v3(f90), n_dot_view).x *
                            fresnel_schlick(v3(1.0f), v3(f90), n_dot_light).x;
    V3 result = diffuse_fresnel * s.diffuse_albedo;

    float half_dot_normal = dot(half_vector, s.normal);
    float roughness_sq = s.roughness * s.roughness;
    float denominator = half_dot_normal * (roughness_sq - 1.0f) + 1.0f;
    float distribution = roughness_sq / (denominator * denominator);

    float masking = n_dot_light * std::sqrt((n_dot_view - roughness_sq * n_dot_view) * n_dot_view + roughness_sq);
    float shadowing = n_dot_view * std::sqrt((n_dot_light - roughness_sq * n_dot_light) * n_dot_light + roughness_sq);
    float geometry = 0.5f / (masking + shadowing);

    V3 specular_fresnel = fresnel_schlick(s.fresnel_0, v3(1.0f), gmax(0.0f, half_dot_view));
    result = result + (distribution * geometry) * specular_fresnel;
    return result / kPi;
}

// brdf.glsl:52-66
static inline V3 sample_ggx_vndf(V3 view_dir, float rx, float ry, V2 rs) {
    V3 transformed_view = normalize(v3(view_dir.x * rx, view_dir.y * ry, view_dir.z));
    float phi = (2.0f * kPi) * rs.x;
    float z = 1.0f - rs.y * (1.0f + transformed_view.z);
    float sin_theta = std::sqrt(gmax(0.0f, 1.0f - z * z));
    V3 hemisphere_sample = v3(sin_theta * cos_cr(phi), sin_theta * sin_cr(phi), z);
    V3 sum = hemisphere_sample + transformed_view;
    return normalize(v3(sum.x * rx, sum.y * ry, sum.z));
}

// brdf.glsl:68-79
static inline float get_ggx_vndf_density(float n_dot_view, float half_dot_normal, float half_dot_view, float roughness) {
    if (half_dot_normal < 0.0f) return 0.0f;
    float roughness_sq = roughness * roughness;
    float inv_roughness_sq = 1.0f - roughness_sq;
    float denominator = n_dot_view + std::sqrt(roughness_sq + inv_roughness_sq * n_dot_view * n_dot_view);
    float d_vis = gmax(0.0f, half_dot_view) * (2.0f / kPi) / denominator;
    float m_sq_term = 1.0f - inv_roughness_sq * half_dot_normal * half_dot_normal;
    return d_vis * roughness_sq / (m_sq_term * m_sq_term);
}

// brdf.glsl:81-84;  reflect(I,N) = I - 2*dot(N,I)*N
static inline V3 sample_ggx_in_dir(V3 view_dir, float roughness, V2 rs) {
    V3 h = sample_ggx_vndf(view_dir, roughness, roughness, rs);
    V3 refl = view_dir - (2.0f * dot(h, view_dir)) * h;
    return -refl;
}

// brdf.glsl:86-93
static inline float get_ggx_in_dir_density(float n_dot_view, V3 view_dir, V3 light_dir, V3 normal, float roughness) {
    V3 half_vector = normalize(light_dir + view_dir);
    float half_dot_view = dot(half_vector, view_dir);
    float half_dot_normal = dot(half_vector, normal);
    float density = get_ggx_vndf_density(n_dot_view, half_dot_normal, half_dot_view, roughness);
    return density / (4.0f * half_dot_view);
}

// brdf.glsl:95-105 — columns of the returned mat3
struct Onb { V3 tx, ty, tz; };
static inline Onb get_shading_space(V3 n) {
    float sign = n.z > 0.0f ? 1.0f : -1.0f;
    float a = -1.0f / (sign + n.z);
    float b = n.x * n.y * a;
    Onb o;
    o.tx = v3(1.0f + sign * n.x * n.x * a, sign * b, -sign * n.x);
    o.ty = v3(b, sign + n.y * n.y * a, -n.y);
    o.tz = n;
    return o;
}
static inline V3 onb_to_world(const Onb& o, V3 v) { return o.tx * v.x + o.ty * v.y + o.tz * v.z; }
static inline V3 onb_to_local(const Onb& o, V3 w) { return v3(dot(o.tx, w), dot(o.ty, w), dot(o.tz, w)); }

// brdf.glsl:107-113
static inline V3 sample_hemisphere_psa(V2 rs) {
    float phi = (2.0f * kPi) * rs.x;
    float radius = std::sqrt(rs.y);
    float z = std::sqrt(1.0f - radius * radius);
    return v3(radius * cos_cr(phi), radius * sin_cr(phi), z);
}
// brdf.glsl:115-117
static inline float get_hemisphere_psa_density(float sampled_z) { return gmax(0.0f, sampled_z) / kPi; }
// brdf.glsl:119-122
static inline float get_diffuse_sampling_probability(const ShadingInfo& s) {
    float luminance = dot(s.diffuse_albedo, v3(0.2126f, 0.7152f, 0.0722f));
    return gmin(0.5f, luminance);
}

// brdf.glsl:124-140
static inline V3 sample_brdf(const ShadingInfo& s, V2 rs) {
    Onb t2w = get_shading_space(s.normal);
    float diffuse_prob = get_diffuse_sampling_probability(s);
    V3 sampled_dir;
    if (rs.x < diffuse_prob) {
        rs.x /= diffuse_prob;
        sampled_dir = onb_to_world(t2w, sample_hemisphere_psa(rs));
    } else {
        rs.x = (rs.x - diffuse_prob) / (1.0f - diffuse_prob);
        V3 local_view = onb_to_local(t2w, s.out_dir);
        V3 local_light = sample_ggx_in_dir(local_view, s.roughness, rs);
        sampled_dir = onb_to_world(t2w, local_light);
    }
    return sampled_dir;
}

// brdf.glsl:142-150
static inline float get_brdf_density(const ShadingInfo& s, V3 sampled_dir) {
    float diffuse_prob = get_diffuse_sampling_probability(s);
    float specular_density = get_ggx_in_dir_density(s.lambert_out, s.out_dir, sampled_dir, s.normal, s.roughness);
    float diffuse_density = get_hemisphere_psa_density(dot(s.normal, sampled_dir));
    return mixf(specular_density, diffuse_density, diffuse_prob);
}

// ------------------------------------------------------------------------------------------
// Scene view + closest hit (our canonical definition of traceRayEXT)
// ------------------------------------------------------------------------------------------
struct Box {
    float lo[3], hi[3];
    void reset() { for (int k = 0; k < 3; ++k) { lo[k] = std::numeric_limits<float>::infinity(); hi[k] = -lo[k]; } }
    void grow(const float* p) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } }
    void grow(const Box& b) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); } }
    float area() const { float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2]; return 2.0f * (dx * dy + dy * dz + dz * dx); }
};

struct Bvh2Node { Box box; int left, right; int first, count; };  // leaf iff count > 0

// generic binned-SAH BVH2 over boxes
struct Bvh2 {
    std::vector<Bvh2Node> nodes;
    std::vector<int> items;
    void build(const std::vector<Box>& boxes, int leaf_size) {
        const int n = (int)boxes.size();
        items.resize(n);
        for (int i = 0; i < n; ++i) items[i] = i;
        nodes.clear();
        if (n == 0) return;
        nodes.reserve(2 * n);
        std::vector<float> cent(3 * (size_t)n);
        for (int i = 0; i < n; ++i)
            for (int k = 0; k < 3; ++k) cent[3 * (size_t)i + k] = 0.5f * (boxes[i].lo[k] + boxes[i].hi[k]);
        nodes.push_back(Bvh2Node{});
        struct Task { int node, first, count; };
        std::vector<Task> stack{{0, 0, n}};
        while (!stack.empty()) {
            Task t = stack.back(); stack.pop_back();
            Box b; b.reset(); Box cb; cb.reset();
            for (int i = t.first; i < t.first + t.count; ++i) { b.grow(boxes[items[i]]); cb.grow(&cent[3 * (size_t)items[i]]); }
            Bvh2Node& nd = nodes[t.node];
            nd.box = b; nd.left = nd.right = -1; nd.first = t.first; nd.count = t.count;
            if (t.count <= leaf_size) continue;
            // binned SAH
            const int NB = 16;
            int best_axis = -1, best_bin = -1; float best_cost = std::numeric_limits<float>::infinity();
            for (int ax = 0; ax < 3; ++ax) {
                float lo = cb.lo[ax], hi = cb.hi[ax];
                if (!(hi > lo)) continue;
                float scale = NB / (hi - lo);
                Box bb[NB]; int bc[NB];
                for (int i = 0; i < NB; ++i) { bb[i].reset(); bc[i] = 0; }
                for (int i = t.first; i < t.first + t.count; ++i) {
                    int id = items[i];
                    int bi = std::min(NB - 1, std::max(0, (int)((cent[3 * (size_t)id + ax] - lo) * scale)));
                    bb[bi].grow(boxes[id]); bc[bi]++;
                }
                float la[NB - 1], ra[NB - 1]; int lc[NB - 1], rc[NB - 1];
                Box acc; acc.reset(); int c = 0;
                for (int i = 0; i < NB - 1; ++i) { acc.grow(bb[i]); c += bc[i]; la[i] = c ? acc.area() : 0.f; lc[i] = c; }
                acc.reset(); c = 0;
                for (int i = NB - 1; i > 0; --i) { acc.grow(bb[i]); c += bc[i]; ra[i - 1] = c ? acc.area() : 0.f; rc[i - 1] = c; }
                for (int i = 0; i < NB - 1; ++i) {
                    if (lc[i] == 0 || rc[i] == 0) continue;
                    float cost = la[i] * lc[i] + ra[i] * rc[i];
                    if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = i; }
                }
            }
            int mid;
            if (best_axis < 0) {
                mid = t.first + t.count / 2;  // all centroids identical: split by index
            } else {
                float lo = cb.lo[best_axis], hi = cb.hi[best_axis];
                float scale = NB / (hi - lo);
                auto it = std::partition(items.begin() + t.first, items.begin() + t.first + t.count, [&](int id) {
                    int bi = std::min(NB - 1, std::max(0, (int)((cent[3 * (size_t)id + best_axis] - lo) * scale)));
                    return bi <= best_bin;
                });
                mid = (int)(it - items.begin());
                if (mid == t.first || mid == t.first + t.count) mid = t.first + t.count / 2;
            }
            int l = (int)nodes.size();
            nodes.push_back(Bvh2Node{}); nodes.push_back(Bvh2Node{});
            nodes[t.node].left = l; nodes[t.node].right = l + 1; nodes[t.node].count = 0;
            stack.push_back({l, t.first, mid - t.first});
            stack.push_back({l + 1, mid, t.first + t.count - mid});
        }
    }
};

struct TriRef { uint32_t prim; uint32_t tri; };  // prim = global primitive index, tri = triangle in primitive

struct MeshAccel {
    std::vector<TriRef> refs;       // all triangles of the mesh, (geometry, triangle) order
    std::vector<float> verts;       // 9 floats per ref
    Bvh2 bvh;
    Box box;
};

struct InstanceXf {
    M3 A;        // object -> world linear part  (rows of transform, Tlas.zig:33-41)
    V3 T;        // translation
    M3 Ainv;     // world -> object
    uint32_t mesh;
};

struct SceneView {
    PtSceneDesc d;
    std::vector<MeshAccel> meshes;
    std::vector<InstanceXf> inst;
    Bvh2 tlas;
    std::vector<Box> inst_boxes;
    bool have_bvh = false;
};

static inline void load3(const uint8_t* base, uint64_t off, uint32_t i, float* out) {
    std::memcpy(out, base + off + 12ull * i, 12);
}
static inline void tri_indices(const PtSceneDesc& d, const PtPrimitive& p, uint32_t tri, uint32_t idx[3]) {
    // triangle.glsl:56-69
    if ((p.info & PT_UINT32_INDICES_MASK) == 0) {
        uint16_t v[3];
        std::memcpy(v, d.triangle_data + p.indices_offset + 6ull * tri, 6);
        idx[0] = v[0]; idx[1] = v[1]; idx[2] = v[2];
    } else {
        std::memcpy(idx, d.triangle_data + p.indices_offset + 12ull * tri, 12);
    }
}

static void build_scene_view(SceneView& sv, const PtSceneDesc& d, bool with_bvh) {
    sv.d = d;
    sv.meshes.resize(d.num_meshes);
    for (size_t m = 0; m < d.num_meshes; ++m) {
        MeshAccel& ma = sv.meshes[m];
        ma.box.reset();
        for (uint32_t p = d.meshes[m].start; p < d.meshes[m].end; ++p) {
            const PtPrimitive& pr = d.primitives[p];
            for (uint32_t t = 0; t < pr.triangle_count; ++t) {
                uint32_t id[3]; tri_indices(d, pr, t, id);
                float v[9];
                for (int k = 0; k < 3; ++k) load3(d.triangle_data, pr.positions_offset, id[k], v + 3 * k);
                ma.refs.push_back(TriRef{p, t});
                ma.verts.insert(ma.verts.end(), v, v + 9);
                for (int k = 0; k < 3; ++k) ma.box.grow(v + 3 * k);
            }
        }
        if (with_bvh) {
            std::vector<Box> boxes(ma.refs.size());
            for (size_t i = 0; i < ma.refs.size(); ++i) { boxes[i].reset(); for (int k = 0; k < 3; ++k) boxes[i].grow(&ma.verts[9 * i + 3 * k]); }
            ma.bvh.build(boxes, 4);
        }
    }
    sv.inst.resize(d.num_instances);
    sv.inst_boxes.resize(d.num_instances);
    for (size_t i = 0; i < d.num_instances; ++i) {
        const float* t = d.instances[i].transform;
        InstanceXf& x = sv.inst[i];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) x.A.m[r][c] = t[4 * r + c];
        x.T = v3(t[3], t[7], t[11]);
        x.Ainv = inverse(x.A);
        x.mesh = (uint32_t)d.instances[i].mesh_index;
        // conservative world box: transform the 8 corners in double, pad
        Box wb; wb.reset();
        const Box& ob = sv.meshes[x.mesh].box;
        if (sv.meshes[x.mesh].refs.empty()) { sv.inst_boxes[i] = wb; continue; }
        for (int c = 0; c < 8; ++c) {
            double p[3] = {(c & 1) ? ob.hi[0] : ob.lo[0], (c & 2) ? ob.hi[1] : ob.lo[1], (c & 4) ? ob.hi[2] : ob.lo[2]};
            float w[3];
            for (int r = 0; r < 3; ++r) w[r] = (float)(x.A.m[r][0] * p[0] + x.A.m[r][1] * p[1] + x.A.m[r][2] * p[2] + (double)t[4 * r + 3]);
            wb.grow(w);
        }
        for (int k = 0; k < 3; ++k) {
            float pad = 1e-5f * std::max(std::max(std::fabs(wb.lo[k]), std::fabs(wb.hi[k])), wb.hi[k] - wb.lo[k]) + 1e-30f;
            wb.lo[k] -= pad; wb.hi[k] += pad;
        }
        sv.inst_boxes[i] = wb;
    }
    if (with_bvh) sv.tlas.build(sv.inst_boxes, 1);
    sv.have_bvh = with_bvh;
}

// Watertight ray/triangle test (Woop, Benthin, Wald 2013) — the shared definition of the
// ray-triangle intersection (DESIGN.md §"Closest-hit rule").
struct RayPre { int kx, ky, kz; float Sx, Sy, Sz; };
static inline RayPre ray_precompute(V3 d) {
    RayPre r;
    int kz = 0; float m = std::fabs(d.x);
    if (std::fabs(d.y) > m) { kz = 1; m = std::fabs(d.y); }
    if (std::fabs(d.z) > m) { kz = 2; }
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    if (idx(d, kz) < 0.0f) std::swap(kx, ky);
    r.kx = kx; r.ky = ky; r.kz = kz;
    // one correctly rounded division; the shear constants only have to be the same for every
    // triangle the ray meets (that is what makes the test watertight), not exact quotients
    r.Sz = 1.0f / idx(d, kz);
    r.Sx = idx(d, kx) * r.Sz;
    r.Sy = idx(d, ky) * r.Sz;
    return r;
}

// returns true and (t,b1,b2) if the front face of (v0,v1,v2) is hit in (tmin,tmax)
static inline bool intersect_tri(V3 o, const RayPre& rp, const float* v, float tmin, float tmax, float& t_out, float& b1, float& b2) {
    const V3 A = v3(v[0], v[1], v[2]) - o, B = v3(v[3], v[4], v[5]) - o, C = v3(v[6], v[7], v[8]) - o;
    const float Ax = idx(A, rp.kx) - rp.Sx * idx(A, rp.kz), Ay = idx(A, rp.ky) - rp.Sy * idx(A, rp.kz);
    const float Bx = idx(B, rp.kx) - rp.Sx * idx(B, rp.kz), By = idx(B, rp.ky) - rp.Sy * idx(B, rp.kz);
    const float Cx = idx(C, rp.kx) - rp.Sx * idx(C, rp.kz), Cy = idx(C, rp.ky) - rp.Sy * idx(C, rp.kz);
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    // front face (geometric normal cross(v1-v0, v2-v0) towards the ray origin) <=> U,V,W >= 0
    if (U < 0.0f || V < 0.0f || W < 0.0f) return false;
    const float det = U + V + W;
    if (det == 0.0f) return false;
    const float Az = rp.Sz * idx(A, rp.kz), Bz = rp.Sz * idx(B, rp.kz), Cz = rp.Sz * idx(C, rp.kz);
    const float T = U * Az + V * Bz + W * Cz;
    const float t = T / det;
    if (!(t > tmin && t < tmax)) return false;
    t_out = t; b1 = V / det; b2 = W / det;
    return true;
}

static inline bool hit_better(float t, uint32_t inst, uint32_t prim, uint32_t tri, const PtHit& h) {
    if (t < h.t) return true;
    if (t > h.t) return false;
    if (inst != h.instance) return inst < h.instance;
    if (prim != h.primitive_index) return prim < h.primitive_index;
    return tri < h.triangle_index;
}

// conservative slab test (padded): may report false positives, never false negatives
static inline bool hit_box(const Box& b, V3 o, V3 inv_d, float tmin, float tmax) {
    float t0 = tmin, t1 = tmax;
    const float oo[3] = {o.x, o.y, o.z}, id[3] = {inv_d.x, inv_d.y, inv_d.z};
    for (int k = 0; k < 3; ++k) {
        float a = (b.lo[k] - oo[k]) * id[k];
        float c = (b.hi[k] - oo[k]) * id[k];
        if (a != a || c != c) continue;  // 0 * inf: origin on a slab plane of a parallel ray: cannot cull on this axis
        float n = std::min(a, c), f = std::max(a, c);
        n = n - std::fabs(n) * 4e-7f; f = f + std::fabs(f) * 4e-7f;
        t0 = std::max(t0, n); t1 = std::min(t1, f);
    }
    return t0 <= t1;
}
static inline V3 safe_inv(V3 d) {
    auto inv = [](float x) { return 1.0f / x; };  // +-0 -> +-inf, handled in hit_box
    return v3(inv(d.x), inv(d.y), inv(d.z));
}

static inline void intersect_instance(const SceneView& sv, uint32_t ii, V3 o, V3 d, float tmin, float tmax, bool use_bvh, PtHit& best, uint64_t* tri_tests) {
    const InstanceXf& x = sv.inst[ii];
    const MeshAccel& ma = sv.meshes[x.mesh];
    if (ma.refs.empty()) return;
    // world -> object: o' = Ainv*(o - T), d' = Ainv*d (not renormalised: t is shared)
    const V3 oo = mul(x.Ainv, o - x.T);
    const V3 dd = mul(x.Ainv, d);
    const RayPre rp = ray_precompute(dd);
    auto test = [&](int ri) {
        float t, b1, b2;
        if (tri_tests) ++*tri_tests;
        if (intersect_tri(oo, rp, &ma.verts[9 * (size_t)ri], tmin, tmax, t, b1, b2)) {
            const TriRef& tr = ma.refs[ri];
            if (hit_better(t, ii, tr.prim, tr.tri, best)) {
                best.t = t; best.b1 = b1; best.b2 = b2; best.instance = ii; best.primitive_index = tr.prim; best.triangle_index = tr.tri;
            }
        }
    };
    if (!use_bvh) {
        for (int i = 0; i < (int)ma.refs.size(); ++i) test(i);
        return;
    }
    const V3 inv_d = safe_inv(dd);
    int stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        const Bvh2Node& n = ma.bvh.nodes[stack[--sp]];
        // prune with best.t (inclusive: equal-t candidates must still be visited for the tie-break)
        if (!hit_box(n.box, oo, inv_d, tmin, std::min(tmax, best.t))) continue;
        if (n.count > 0) {
            for (int i = n.first; i < n.first + n.count; ++i) test(ma.bvh.items[i]);
        } else {
            stack[sp++] = n.left; stack[sp++] = n.right;
        }
    }
}

static inline PtHit closest_hit(const SceneView& sv, V3 o, V3 d, float tmin, float tmax, bool use_bvh, uint64_t* tri_tests = nullptr) {
    PtHit best; best.t = 1e32f; best.b1 = best.b2 = 0.0f;
    best.instance = best.primitive_index = best.triangle_index = 0xFFFFFFFFu;
    if (!use_bvh || !sv.have_bvh) {
        for (uint32_t i = 0; i < sv.inst.size(); ++i) intersect_instance(sv, i, o, d, tmin, tmax, false, best, tri_tests);
        return best;
    }
    if (sv.tlas.nodes.empty()) return best;
    const V3 inv_d = safe_inv(d);
    int stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        const Bvh2Node& n = sv.tlas.nodes[stack[--sp]];
        if (!hit_box(n.box, o, inv_d, tmin, std::min(tmax, best.t))) continue;
        if (n.count > 0) {
            for (int i = n.first; i < n.first + n.count; ++i) intersect_instance(sv, (uint32_t)sv.tlas.items[i], o, d, tmin, tmax, true, best, tri_tests);
        } else {
            stack[sp++] = n.left; stack[sp++] = n.right;
        }
    }
    return best;
}

static bool g_count_dead = false;
static std::atomic<uint64_t> g_dead_rays{0};

// ------------------------------------------------------------------------------------------
// Hit attribute fetch — src/shaders/triangle.glsl:46-112
// ------------------------------------------------------------------------------------------
struct TriangleData { V2 uv; V3 normal; V4 tangent; V3 geometry_normal; uint32_t material_index; };

static inline TriangleData get_triangle_data(const SceneView& sv, const PtHit& h) {
    const PtSceneDesc& d = sv.d;
    const PtPrimitive& p = d.primitives[h.primitive_index];
    TriangleData td;
    const V3 c = v3(1.0f - h.b1 - h.b2, h.b1, h.b2);              // :50-54
    uint32_t id[3]; tri_indices(d, p, h.triangle_index, id);       // :56-69
    {                                                              // :71-77
        float u[3][2];
        for (int k = 0; k < 3; ++k) std::memcpy(u[k], d.triangle_data + p.uvs_offset + 8ull * id[k], 8);
        td.uv.x = u[0][0] * c.x + u[1][0] * c.y + u[2][0] * c.z;
        td.uv.y = u[0][1] * c.x + u[1][1] * c.y + u[2][1] * c.z;
    }
    // :79  `v * transpose(inverse(A))` == inverse(A) * v  (quirk Q3/Q4, copied literally)
    const M3& Ainv = sv.inst[h.instance].Ainv;
    {                                                              // :80-87
        float n[3][3];
        for (int k = 0; k < 3; ++k) load3(d.triangle_data, p.normals_offset, id[k], n[k]);
        V3 nn = v3(n[0][0], n[0][1], n[0][2]) * c.x + v3(n[1][0], n[1][1], n[1][2]) * c.y + v3(n[2][0], n[2][1], n[2][2]) * c.z;
        td.normal = normalize(mul(Ainv, nn));
    }
    {                                                              // :89-97
        float t[3][4];
        for (int k = 0; k < 3; ++k) std::memcpy(t[k], d.triangle_data + p.tangents_offset + 16ull * id[k], 16);
        V3 tt = v3(t[0][0], t[0][1], t[0][2]) * c.x + v3(t[1][0], t[1][1], t[1][2]) * c.y + v3(t[2][0], t[2][1], t[2][2]) * c.z;
        float tw = t[0][3] * c.x + t[1][3] * c.y + t[2][3] * c.z;
        V3 txyz = normalize(mul(Ainv, tt));
        td.tangent = V4{txyz.x, txyz.y, txyz.z, tw};
    }
    {                                                              // :99-107
        float q[3][3];
        for (int k = 0; k < 3; ++k) load3(d.triangle_data, p.positions_offset, id[k], q[k]);
        V3 p0 = v3(q[0][0], q[0][1], q[0][2]), p1 = v3(q[1][0], q[1][1], q[1][2]), p2 = v3(q[2][0], q[2][1], q[2][2]);
        td.geometry_normal = normalize(mul(Ainv, cross(p1 - p0, p2 - p0)));
    }
    td.material_index = p.info & PT_MATERIAL_INDEX_MASK;           // :109
    return td;
}

// ------------------------------------------------------------------------------------------
// Material fetch — src/shaders/material.glsl:31-69; sampler src/RayTracingPipeline.zig:54-70
// ------------------------------------------------------------------------------------------
static inline V4 unpack_unorm4x8(uint32_t p) {
    return V4{(float)(p & 0xff) / 255.0f, (float)((p >> 8) & 0xff) / 255.0f, (float)((p >> 16) & 0xff) / 255.0f, (float)(p >> 24) / 255.0f};
}

// texture(): bilinear, REPEAT, LOD 0, RGBA8 UNORM (Vulkan spec texel filtering: u*W - 0.5, floor,
// fract, wrap mod W).
static inline V4 sample_texture(const PtTexture& tex, V2 uv) {
    const float W = (float)tex.width, H = (float)tex.height;
    const float u = uv.x * W - 0.5f, v = uv.y * H - 0.5f;
    const float fu = std::floor(u), fv = std::floor(v);
    const float a = u - fu, b = v - fv;
    auto wrap = [](float f, uint32_t n) -> uint32_t {
        // f is integral; exact modulo for |f| < 2^31, otherwise texel 0 (never reached by sane uvs)
        if (!(f >= -2147483648.0f && f < 2147483648.0f)) return 0u;
        int64_t i = (int64_t)f; int64_t m = i % (int64_t)n; if (m < 0) m += n; return (uint32_t)m;
    };
    const uint32_t x0 = wrap(fu, tex.width), x1 = (x0 + 1 == tex.width) ? 0 : x0 + 1;
    const uint32_t y0 = wrap(fv, tex.height), y1 = (y0 + 1 == tex.height) ? 0 : y0 + 1;
    auto texel = [&](uint32_t x, uint32_t y) -> V4 {
        uint32_t p; std::memcpy(&p, tex.rgba + 4ull * ((uint64_t)y * tex.width + x), 4);
        return unpack_unorm4x8(p);
    };
    const V4 t00 = texel(x0, y0), t10 = texel(x1, y0), t01 = texel(x0, y1), t11 = texel(x1, y1);
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    V4 r;
    r.x = t00.x * w00 + t10.x * w10 + t01.x * w01 + t11.x * w11;
    r.y = t00.y * w00 + t10.y * w10 + t01.y * w01 + t11.y * w11;
    r.z = t00.z * w00 + t10.z * w10 + t01.z * w01 + t11.z * w11;
    r.w = t00.w * w00 + t10.w * w10 + t01.w * w01 + t11.w * w11;
    return r;
}

struct MaterialData { V4 albedo; V3 normal; V3 emissive; float roughness, metallic; };

static inline MaterialData get_material_data(const SceneView& sv, const TriangleData& td) {
    const PtMaterial& m = sv.d.materials[td.material_index];
    MaterialData md;
    md.albedo = unpack_unorm4x8(m.albedo_factor);                                   // :36
    if (m.albedo_texture_index != PT_INVALID_TEXTURE) {                              // :37-39
        V4 t = sample_texture(sv.d.textures[m.albedo_texture_index], td.uv);
        md.albedo = V4{md.albedo.x * t.x, md.albedo.y * t.y, md.albedo.z * t.z, md.albedo.w * t.w};
    }
    const V4 mr = unpack_unorm4x8(m.metal_roughness_factor);                         // :41-43
    md.roughness = mr.y; md.metallic = mr.z;
    if (m.metal_roughness_texture_index != PT_INVALID_TEXTURE) {                     // :44-48
        V4 t = sample_texture(sv.d.textures[m.metal_roughness_texture_index], td.uv);
        md.roughness *= t.y; md.metallic *= t.z;
    }
    const V4 e = unpack_unorm4x8(m.emissive_factor);                                 // :50
    md.emissive = v3(e.x, e.y, e.z);
    if (m.emissive_texture_index != PT_INVALID_TEXTURE) {                            // :51-53
        V4 t = sample_texture(sv.d.textures[m.emissive_texture_index], td.uv);
        md.emissive = md.emissive * v3(t.x, t.y, t.z);
    }
    md.normal = td.normal;                                                           // :55
    if (m.normal_texture_index != PT_INVALID_TEXTURE) {                              // :56-66
        V4 t = sample_texture(sv.d.textures[m.normal_texture_index], td.uv);
        const V3 nts = v3(t.x, t.y, t.z);
        const V3 T = v3(td.tangent.x, td.tangent.y, td.tangent.z);
        const V3 bitangent = cross(td.normal, T) * td.tangent.w;
        const V3 n = normalize(2.0f * nts - v3(1.0f));
        md.normal = T * n.x + bitangent * n.y + td.normal * n.z;
    }
    return md;
}

// ------------------------------------------------------------------------------------------
// The trace loop — src/shaders/ray_gen.rgen:33-105
// ------------------------------------------------------------------------------------------
struct PixelOut { V3 color; PtHit primary; uint64_t rays; };

static inline PixelOut shade_pixel(const SceneView& sv, const PtCamera& cam, uint32_t x, uint32_t y, uint32_t W, uint32_t H,
                                   uint32_t S, uint32_t B, bool use_bvh, uint32_t flags = 0) {
    PixelOut out; out.rays = 0;
    const float px = (float)x + 0.5f, py = (float)y + 0.5f;                     // :34
    const float u = px / (float)W, v = py / (float)H;                           // :35
    const float dx = u * 2.0f - 1.0f, dy = v * 2.0f - 1.0f;                     // :36
    const float tmin = 0.001f, tmax = 10000.0f;                                 // :38-39
    U2 seed = prng_seed(x, y, cam.frame_count);                                 // :41
    const V3 pos = v3(cam.position[0], cam.position[1], cam.position[2]);
    const V3 hor = v3(cam.horizontal[0], cam.horizontal[1], cam.horizontal[2]);
    const V3 ver = v3(cam.vertical[0], cam.vertical[1], cam.vertical[2]);
    const V3 fwd = v3(cam.forward[0], cam.forward[1], cam.forward[2]);
    V3 final_color = v3(0.0f);                                                  // :43
    out.primary.t = 1e32f; out.primary.b1 = out.primary.b2 = 0.f;
    out.primary.instance = out.primary.primitive_index = out.primary.triangle_index = 0xFFFFFFFFu;
    for (uint32_t i = 0; i < S; ++i) {                                          // :44
        V3 color = v3(0.0f), atten = v3(1.0f);                                  // :45-46
        V3 origin = pos;                                                        // :47
        V3 direction = normalize(fwd + hor * dx - ver * dy);                    // :48
        for (uint32_t j = 0; j < B; ++j) {                                      // :49
            // PT_FLAG_SKIP_DEAD_SEGMENTS (include/b200pt.h): the last segment draws no random numbers and only adds
            // atten * (emission or sky), both finite, so with a throughput of exactly zero it adds exactly nothing
            if ((flags & PT_FLAG_SKIP_DEAD_SEGMENTS) && j == B - 1 && j > 0 && atten.x == 0.0f && atten.y == 0.0f && atten.z == 0.0f) break;
            const PtHit h = closest_hit(sv, origin, direction, tmin, tmax, use_bvh);  // :50-63
            ++out.rays;
            if (i == 0 && j == 0) out.primary = h;
            if (h.t == 1e32f) {                                                 // :65-70
                float t = 0.5f * (direction.y + 1.0f);
                color = color + atten * mix(v3(1.0f, 1.0f, 1.0f), v3(0.5f, 0.7f, 1.0f), t);
                break;
            }
            const TriangleData td = get_triangle_data(sv, h);                   // :72
            const MaterialData md = get_material_data(sv, td);                  // :73
            color = color + md.emissive * 10.0f * atten;                        // :75
            if (j < B - 1) {                                                    // :77
                ShadingInfo s;
                s.out_dir = -direction;                                         // :79
                s.normal = md.normal;                                           // :80
                s.lambert_out = dot(s.normal, s.out_dir);                       // :81
                const V3 alb = v3(md.albedo.x, md.albedo.y, md.albedo.z);
                s.fresnel_0 = mix(v3(0.02f), alb, md.metallic);                 // :82
                s.diffuse_albedo = alb - md.metallic * alb;                     // :83
                s.roughness = gmax(0.006f, md.roughness);                       // :84
                origin = origin + h.t * direction + td.geometry_normal * 0.00001f;  // :86
                direction = sample_brdf(s, pcg2d(seed));                        // :87
                float density = get_brdf_density(s, direction);                 // :89
                float lambert_in = dot(s.normal, direction);                    // :90
                if (lambert_in <= 0.0f) break;                                  // :91-93
                atten = atten * (brdf(s, direction) * lambert_in / density);    // :95
                // PT_FLAG_RUSSIAN_ROULETTE (extension, not in the reference; include/b200pt.h): from the second bounce on the
                // path survives with probability p = clamp(max(atten), 0.05, 1), one more pcg2d draw decides
                if ((flags & PT_FLAG_RUSSIAN_ROULETTE) && j >= 1) {
                    const float p = gmin(1.0f, gmax(0.05f, gmax(atten.x, gmax(atten.y, atten.z))));
                    if (p < 1.0f) {
                        const V2 xr = pcg2d(seed);
                        if (!(xr.x < p)) break;
                        atten = atten / p;
                    }
                }
                // statistic for DESIGN.md section 10: last segments about to be traced with a throughput of exactly zero (brdf()
                // returns 0 when the shading normal faces away from the viewer)
                if (g_count_dead && j + 2 == B && atten.x == 0.0f && atten.y == 0.0f && atten.z == 0.0f) g_dead_rays.fetch_add(1, std::memory_order_relaxed);
            }
        }
        final_color = final_color + color;                                      // :99
    }
    out.color = final_color / (float)S;                                         // :102
    return out;
}

}  // namespace orc

// ==========================================================================================
// C entry points (ctypes)
// ==========================================================================================
using namespace orc;

extern "C" {

struct OrcScene { SceneView sv; };

OrcScene* orc_scene_create(const PtSceneDesc* d, int with_bvh) {
    OrcScene* s = new OrcScene();
    build_scene_view(s->sv, *d, with_bvh != 0);
    return s;
}
void orc_scene_destroy(OrcScene* s) { delete s; }

// Render one frame.  rgba: W*H*4 floats (may be null); ids: 3 arrays of W*H u32 (may be null).
// Pixels with (tile owner != rank) are skipped when world > 1 (tile = 32x8, owner = tile % world)
// and left untouched.  Returns the number of rays (closest-hit queries) executed.
uint64_t orc_render(const OrcScene* s, const PtRenderConfig* cfg, const PtCamera* cam, int use_bvh, int nthreads,
                    float* rgba, uint32_t* id_inst, uint32_t* id_prim, uint32_t* id_tri) {
    const uint32_t div = cfg->divider ? cfg->divider : 1;
    const uint32_t W = cfg->width / div, H = cfg->height / div;
    if (nthreads < 1) nthreads = 1;
    std::atomic<uint64_t> rays{0};
    std::atomic<uint32_t> next_row{0};
    const uint32_t tiles_x = (W + 31) / 32;
    auto worker = [&]() {
        uint64_t local = 0;
        for (;;) {
            uint32_t y = next_row.fetch_add(1);
            if (y >= H) break;
            for (uint32_t x = 0; x < W; ++x) {
                if (cfg->partition == PT_PARTITION_TILES && cfg->world > 1) {
                    uint32_t tile = (y / 8) * tiles_x + (x / 32);
                    if ((int)(tile % (uint32_t)cfg->world) != cfg->rank) continue;
                }
                PixelOut po = shade_pixel(s->sv, *cam, x, y, W, H, cfg->num_samples, cfg->num_bounces, use_bvh != 0, cfg->flags);
                local += po.rays;
                size_t i = (size_t)y * W + x;
                if (rgba) { rgba[4 * i + 0] = po.color.x; rgba[4 * i + 1] = po.color.y; rgba[4 * i + 2] = po.color.z; rgba[4 * i + 3] = 1.0f; }
                if (id_inst) id_inst[i] = po.primary.instance;
                if (id_prim) id_prim[i] = po.primary.primitive_index;
                if (id_tri) id_tri[i] = po.primary.triangle_index;
            }
        }
        rays += local;
    };
    std::vector<std::thread> th;
    for (int i = 1; i < nthreads; ++i) th.emplace_back(worker);
    worker();
    for (auto& t : th) t.join();
    return rays.load();
}

// Render only rows [y0, y1) (bounded CPU-baseline samples).  Same outputs/indexing as orc_render.
uint64_t orc_render_rows(const OrcScene* s, const PtRenderConfig* cfg, const PtCamera* cam, int use_bvh, int nthreads,
                         uint32_t y0, uint32_t y1, uint32_t row_step, float* rgba) {
    const uint32_t div = cfg->divider ? cfg->divider : 1;
    const uint32_t W = cfg->width / div, H = cfg->height / div;
    if (y1 > H) y1 = H;
    if (row_step == 0) row_step = 1;
    if (nthreads < 1) nthreads = 1;
    std::atomic<uint64_t> rays{0};
    std::atomic<uint32_t> next_row{0};
    const uint32_t nrows = (y1 > y0) ? (y1 - y0 + row_step - 1) / row_step : 0;
    auto worker = [&]() {
        uint64_t local = 0;
        for (;;) {
            uint32_t r = next_row.fetch_add(1);
            if (r >= nrows) break;
            uint32_t y = y0 + r * row_step;
            for (uint32_t x = 0; x < W; ++x) {
                PixelOut po = shade_pixel(s->sv, *cam, x, y, W, H, cfg->num_samples, cfg->num_bounces, use_bvh != 0, cfg->flags);
                local += po.rays;
                size_t i = (size_t)y * W + x;
                if (rgba) { rgba[4 * i + 0] = po.color.x; rgba[4 * i + 1] = po.color.y; rgba[4 * i + 2] = po.color.z; rgba[4 * i + 3] = 1.0f; }
            }
        }
        rays += local;
    };
    std::vector<std::thread> th;
    for (int i = 1; i < nthreads; ++i) th.emplace_back(worker);
    worker();
    for (auto& t : th) t.join();
    return rays.load();
}

// Trace arbitrary rays (traversal parity tests).
void orc_trace(const OrcScene* s, const PtRay* rays, size_t n, int use_bvh, int nthreads, PtHit* hits) {
    if (nthreads < 1) nthreads = 1;
    std::atomic<size_t> next{0};
    auto worker = [&]() {
        for (;;) {
            size_t b = next.fetch_add(256);
            if (b >= n) break;
            size_t e = std::min(n, b + 256);
            for (size_t i = b; i < e; ++i) {
                const PtRay& r = rays[i];
                hits[i] = closest_hit(s->sv, v3(r.origin[0], r.origin[1], r.origin[2]), v3(r.dir[0], r.dir[1], r.dir[2]), r.tmin, r.tmax, use_bvh != 0);
            }
        }
    };
    std::vector<std::thread> th;
    for (int i = 1; i < nthreads; ++i) th.emplace_back(worker);
    worker();
    for (auto& t : th) t.join();
}

// One closest-hit query: the "driver" side of traceRayEXT for the reference's own shaders compiled as C++
// (oracle/ref_glsl/ref_glsl_api.cpp takes this function's address as its trace callback).
int orc_closest_hit_one(const void* scene, const float o[3], float tmin, const float d[3], float tmax, int use_bvh, PtHit* out) {
    const OrcScene* s = static_cast<const OrcScene*>(scene);
    *out = closest_hit(s->sv, v3(o[0], o[1], o[2]), v3(d[0], d[1], d[2]), tmin, tmax, use_bvh != 0);
    return out->t != 1e32f;
}

// getTriangleData + getMaterialData of n hits, 25 floats each: uv(2) normal(3) tangent(4) geometry_normal(3)
// material_index(1, raw bits) | albedo(4) normal(3) emissive(3) roughness metallic
void orc_shade_hits(const OrcScene* s, const PtHit* hits, size_t n, float* out) {
    for (size_t i = 0; i < n; ++i) {
        const TriangleData td = get_triangle_data(s->sv, hits[i]);
        const MaterialData md = get_material_data(s->sv, td);
        float* o = out + 25 * i;
        o[0] = td.uv.x; o[1] = td.uv.y; o[2] = td.normal.x; o[3] = td.normal.y; o[4] = td.normal.z;
        o[5] = td.tangent.x; o[6] = td.tangent.y; o[7] = td.tangent.z; o[8] = td.tangent.w;
        o[9] = td.geometry_normal.x; o[10] = td.geometry_normal.y; o[11] = td.geometry_normal.z;
        std::memcpy(o + 12, &td.material_index, 4);
        o[13] = md.albedo.x; o[14] = md.albedo.y; o[15] = md.albedo.z; o[16] = md.albedo.w;
        o[17] = md.normal.x; o[18] = md.normal.y; o[19] = md.normal.z;
        o[20] = md.emissive.x; o[21] = md.emissive.y; o[22] = md.emissive.z; o[23] = md.roughness; o[24] = md.metallic;
    }
}

// --- scalar entry points for known-answer tests -------------------------------------------
void orc_prng_seed(uint32_t x, uint32_t y, uint32_t frame, uint32_t out[2]) { U2 s = prng_seed(x, y, frame); out[0] = s.x; out[1] = s.y; }
void orc_pcg2d(uint32_t seed[2], float out[2]) { U2 s{seed[0], seed[1]}; V2 r = pcg2d(s); seed[0] = s.x; seed[1] = s.y; out[0] = r.x; out[1] = r.y; }

static ShadingInfo make_si(const float n[3], const float v[3], const float kd[3], const float f0[3], float roughness) {
    ShadingInfo s;
    s.normal = v3(n[0], n[1], n[2]); s.out_dir = v3(v[0], v[1], v[2]);
    s.lambert_out = dot(s.normal, s.out_dir);
    s.diffuse_albedo = v3(kd[0], kd[1], kd[2]); s.fresnel_0 = v3(f0[0], f0[1], f0[2]); s.roughness = roughness;
    return s;
}
// Batched: arrays of n records.  sample_brdf -> L; get_brdf_density -> pdf; brdf -> f.
void orc_bsdf_batch(size_t n, const float* N, const float* V, const float* kd, const float* f0, const float* rough, const float* xi,
                    float* L_out, float* pdf_out, float* f_out) {
    for (size_t i = 0; i < n; ++i) {
        ShadingInfo s = make_si(N + 3 * i, V + 3 * i, kd + 3 * i, f0 + 3 * i, rough[i]);
        V3 L = sample_brdf(s, V2{xi[2 * i], xi[2 * i + 1]});
        float pdf = get_brdf_density(s, L);
        V3 f = brdf(s, L);
        L_out[3 * i] = L.x; L_out[3 * i + 1] = L.y; L_out[3 * i + 2] = L.z;
        pdf_out[i] = pdf;
        f_out[3 * i] = f.x; f_out[3 * i + 1] = f.y; f_out[3 * i + 2] = f.z;
    }
}
void orc_fresnel_schlick(const float f0[3], const float f90[3], float c, float out[3]) {
    V3 r = fresnel_schlick(v3(f0[0], f0[1], f0[2]), v3(f90[0], f90[1], f90[2]), c); out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void orc_shading_space(const float n[3], float out[9]) {
    Onb o = get_shading_space(v3(n[0], n[1], n[2]));
    out[0] = o.tx.x; out[1] = o.tx.y; out[2] = o.tx.z; out[3] = o.ty.x; out[4] = o.ty.y; out[5] = o.ty.z; out[6] = o.tz.x; out[7] = o.tz.y; out[8] = o.tz.z;
}
void orc_sample_texture(const uint8_t* rgba, uint32_t w, uint32_t h, float u, float v, float out[4]) {
    PtTexture t{rgba, w, h}; V4 r = sample_texture(t, V2{u, v}); out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
// single ray / single triangle (object space), for the watertightness / facing tests
int orc_intersect_tri(const float o[3], const float d[3], const float v[9], float tmin, float tmax, float out[3]) {
    RayPre rp = ray_precompute(v3(d[0], d[1], d[2]));
    float t, b1, b2;
    if (!intersect_tri(v3(o[0], o[1], o[2]), rp, v, tmin, tmax, t, b1, b2)) return 0;
    out[0] = t; out[1] = b1; out[2] = b2; return 1;
}
uint64_t orc_dead_ray_statistic(int enable) { g_count_dead = enable != 0; return g_dead_rays.exchange(0); }

int orc_hardware_threads(void) { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
