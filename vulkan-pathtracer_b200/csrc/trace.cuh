// trace.cuh — closest-hit query: replaces traceRayEXT (src/shaders/ray_gen.rgen:50-63) +
// closest_hit.rchit:10-17 + miss.rmiss.  TLAS -> BLAS traversal of the compressed BVH8, instance
// transform on leaf entry, watertight ray/triangle test, back-face culling, (tmin, tmax) open
// interval, order-independent closest hit (lexicographic (t, instance, primitive, triangle)).
#pragma once
#include "pt_internal.h"

namespace pt {

// ------------------------------------------------------------------ watertight ray/triangle test
// Woop, Benthin, Wald 2013.  Same expression tree as the oracle (DESIGN.md "Closest-hit rule").
// The ray-space shear is carried as two matrix rows m0 = e_kx - Sx*e_kz, m1 = e_ky - Sy*e_kz, so
// that the sheared coordinates are branch-free dot products: (m0x*A.x + m0y*A.y) + m0z*A.z has one
// exact term (1*A[kx]), one rounded product (-(Sx*A[kz])), one zero and then one rounded sum — bit
// for bit the oracle's A[kx] - Sx*A[kz] (no FMA contraction: --fmad=false) without the dynamic
// component selects, which ptxas turned into seven divergent branch regions per test (profiles/r1a:
// comp() alone was 8 % of K-trace's issue slots at 5.6 lanes).
struct RayPre { float m0x, m0y, m0z, m1x, m1y, m1z, Sz; int kz; };

PT_D RayPre ray_precompute(V3 d) {
    RayPre r;
    int kz = 0; float m = fabsf(d.x);
    if (fabsf(d.y) > m) { kz = 1; m = fabsf(d.y); }
    if (fabsf(d.z) > m) { kz = 2; }
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    const float dz = kz == 0 ? d.x : (kz == 1 ? d.y : d.z);
    if (dz < 0.0f) { int t = kx; kx = ky; ky = t; }
    const float dx = kx == 0 ? d.x : (kx == 1 ? d.y : d.z), dy = ky == 0 ? d.x : (ky == 1 ? d.y : d.z);
    // one correctly rounded division (same definition as the oracle): watertightness only needs the
    // same shear constants for every triangle the ray meets
    r.Sz = 1.0f / dz;
    const float nSx = -(dx * r.Sz), nSy = -(dy * r.Sz);
    r.kz = kz;
    r.m0x = kx == 0 ? 1.0f : (kz == 0 ? nSx : 0.0f); r.m0y = kx == 1 ? 1.0f : (kz == 1 ? nSx : 0.0f); r.m0z = kx == 2 ? 1.0f : (kz == 2 ? nSx : 0.0f);
    r.m1x = ky == 0 ? 1.0f : (kz == 0 ? nSy : 0.0f); r.m1y = ky == 1 ? 1.0f : (kz == 1 ? nSy : 0.0f); r.m1z = ky == 2 ? 1.0f : (kz == 2 ? nSy : 0.0f);
    return r;
}

// hit distance + the unnormalised barycentrics (b1 = V / det, b2 = W / det)
PT_D bool intersect_tri_raw(V3 o, const RayPre& rp, float4 p0, float4 p1, float4 p2, float tmin, float tmax, float& t_out,
                            float& V_out, float& W_out, float& det_out) {
    const V3 A = v3(p0.x, p0.y, p0.z) - o, B = v3(p1.x, p1.y, p1.z) - o, C = v3(p2.x, p2.y, p2.z) - o;
    const float Ax = (rp.m0x * A.x + rp.m0y * A.y) + rp.m0z * A.z, Ay = (rp.m1x * A.x + rp.m1y * A.y) + rp.m1z * A.z;
    const float Bx = (rp.m0x * B.x + rp.m0y * B.y) + rp.m0z * B.z, By = (rp.m1x * B.x + rp.m1y * B.y) + rp.m1z * B.z;
    const float Cx = (rp.m0x * C.x + rp.m0y * C.y) + rp.m0z * C.z, Cy = (rp.m1x * C.x + rp.m1y * C.y) + rp.m1z * C.z;
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    // front face (cross(v1-v0, v2-v0) towards the ray origin) <=> U,V,W >= 0
    if (U < 0.0f || V < 0.0f || W < 0.0f) return false;
    const float det = U + V + W;
    if (det == 0.0f) return false;
    const float Akz = comp(A, rp.kz), Bkz = comp(B, rp.kz), Ckz = comp(C, rp.kz);
    const float Az = rp.Sz * Akz, Bz = rp.Sz * Bkz, Cz = rp.Sz * Ckz;
    const float T = U * Az + V * Bz + W * Cz;
    const float t = T / det;
    if (!(t > tmin && t < tmax)) return false;
    t_out = t; V_out = V; W_out = W; det_out = det;
    return true;
}

PT_D bool intersect_tri(V3 o, const RayPre& rp, float4 p0, float4 p1, float4 p2, float tmin, float tmax, float& t_out,
                        float& b1, float& b2) {
    float V, W, det;
    if (!intersect_tri_raw(o, rp, p0, p1, p2, tmin, tmax, t_out, V, W, det)) return false;
    b1 = V / det; b2 = W / det;
    return true;
}

struct Hit { float t, b1, b2; uint32_t inst, prim, tri; };

PT_D void hit_init(Hit& h) { h.t = kMissT; h.b1 = 0.0f; h.b2 = 0.0f; h.inst = h.prim = h.tri = kInvalidId; }

PT_D bool hit_better(float t, uint32_t inst, uint32_t prim, uint32_t tri, const Hit& h) {
    if (t < h.t) return true;
    if (t > h.t) return false;
    if (inst != h.inst) return inst < h.inst;
    if (prim != h.prim) return prim < h.prim;
    return tri < h.tri;
}

PT_D void test_triangle(const Bvh8Tri* __restrict__ tris, uint32_t idx, V3 o, const RayPre& rp, float tmin, float tmax,
                        uint32_t inst, Hit& best) {
    const float4* q = reinterpret_cast<const float4*>(tris + idx);
    const float4 p0 = __ldg(q), p1 = __ldg(q + 1), p2 = __ldg(q + 2);
    // tmax stays the ray's tmax: candidates with t == best.t must reach the tie-break
    // the two barycentric divisions only for candidates that replace the closest hit (same operands, same bits; measured
    // C2 1.334 -> 1.312 ms, C4 14.98 -> 14.83 ms, profiles/r2a_variants.txt)
    float t, V, W, det;
    if (intersect_tri_raw(o, rp, p0, p1, p2, tmin, tmax, t, V, W, det)) {
        const uint32_t prim = __float_as_uint(p0.w), tri = __float_as_uint(p1.w);
        if (hit_better(t, inst, prim, tri, best)) { best.t = t; best.b1 = V / det; best.b2 = W / det; best.inst = inst; best.prim = prim; best.tri = tri; }
    }
}

// ------------------------------------------------------------------ ray frame for box tests
struct RayFrame {
    V3 o;            // origin in the current space
    V3 idir;         // 1/d with |d| clamped away from 0
    uint32_t rs;     // octant: bit k set when d[k] >= 0
};

// reciprocal for the (conservative) box tests only: MUFU.RCP, <= 1 ulp, covered by the 2^-20 slab padding
PT_D float safe_rcp(float d) {
    const float a = fabsf(d) < 1e-20f ? copysignf(1e-20f, d) : d;
    float r;
    asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
}
PT_D void frame_set(RayFrame& f, V3 o, V3 d) {
    f.o = o;
    f.idir = v3(safe_rcp(d.x), safe_rcp(d.y), safe_rcp(d.z));
    f.rs = (f.idir.x >= 0.0f ? 1u : 0u) | (f.idir.y >= 0.0f ? 2u : 0u) | (f.idir.z >= 0.0f ? 4u : 0u);
}

// Producer side of K-trace's ray fetch: the world-space ray with everything the traversal derives
// from it (box-test reciprocals + octant, watertight shear rows), 80 bytes per slot.  Computed
// where rays are born (K-raygen / K-shade / k_ray_setup, all lanes busy) instead of in K-trace,
// where refilled lanes would do it two at a time (profiles/r1a: 22 % of K-trace's issue slots).
// skip: bit 0 = do not write ray_o, bit 1 = do not write ray_d (FrameParams::primary_lean; 0 everywhere but K-raygen)
PT_D void store_ray(float4* __restrict__ ray_o, float4* __restrict__ ray_d, float4* __restrict__ ray_e, float4* __restrict__ ray_s,
                    float4* __restrict__ ray_m, uint32_t slot, V3 o, V3 d, float tmin, float tmax, uint32_t skip = 0u) {
    RayFrame f;
    frame_set(f, o, d);
    const RayPre rp = ray_precompute(d);
    const uint32_t bits = f.rs | ((uint32_t)rp.kz << 3);
    if (!(skip & 1u)) ray_o[slot] = make_float4(o.x, o.y, o.z, tmin);
    if (!(skip & 2u)) ray_d[slot] = make_float4(d.x, d.y, d.z, tmax);
    ray_e[slot] = make_float4(f.idir.x, f.idir.y, f.idir.z, __uint_as_float(bits));
    ray_s[slot] = make_float4(rp.m0x, rp.m0y, rp.m0z, rp.Sz);
    ray_m[slot] = make_float4(rp.m1x, rp.m1y, rp.m1z, tmax);    // tmax again: K-trace's fetch then needs no ray_d (64 B per ray)
}

// world -> object: o' = Ainv*(o - T), d' = Ainv*d (not renormalised: t is shared between spaces).
// Same expression as the oracle.
PT_D void to_object(const InstanceRec* __restrict__ inst, uint32_t i, V3 o, V3 d, V3& oo, V3& dd) {
    const float4* q = reinterpret_cast<const float4*>(inst[i].inv);
    const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    // a = (m00 m01 m02 m10) b = (m11 m12 m20 m21) c = (m22 Tx Ty Tz)
    const V3 p = o - v3(c.y, c.z, c.w);
    oo = v3(a.x * p.x + a.y * p.y + a.z * p.z, a.w * p.x + b.x * p.y + b.y * p.z, b.z * p.x + b.w * p.y + c.x * p.z);
    dd = v3(a.x * d.x + a.y * d.y + a.z * d.z, a.w * d.x + b.x * d.y + b.y * d.z, b.z * d.x + b.w * d.y + c.x * d.z);
}

// ------------------------------------------------------------------ BVH8 node test
// Returns the hit mask of one 80-byte node: bits 31..24 = inner children in traversal priority
// (bit 24 + (slot ^ rs): nearest octant first when popped from the top), bits 15..0 = leaf items
// relative to tri_base (dense: bit k = item tri_base + k).  Child boxes are decoded from 8-bit grids; the slab test is conservative:
// near planes are pushed down and far planes up by 2^-20 relative (explicit FMAs).
PT_D uint32_t node_test(const Bvh8Node* __restrict__ nodes, uint32_t idx, const RayFrame& f, float tmin, float tmax,
                        uint32_t& child_base, uint32_t& tri_base, uint32_t& imask) {
    const uint4* q = reinterpret_cast<const uint4*>(nodes + idx);
    const uint4 n0 = __ldg(q), n1 = __ldg(q + 1), n2 = __ldg(q + 2), n3 = __ldg(q + 3), n4 = __ldg(q + 4);
    child_base = n1.x; tri_base = n1.y; imask = n0.w >> 24;
    const float eps = 9.5367431640625e-7f;  // 2^-20
    float an[3], af[3], bn[3], bf[3];
    {
        const float sc[3] = {__uint_as_float((n0.w & 0xffu) << 23), __uint_as_float(((n0.w >> 8) & 0xffu) << 23),
                             __uint_as_float(((n0.w >> 16) & 0xffu) << 23)};
        const float pp[3] = {__uint_as_float(n0.x), __uint_as_float(n0.y), __uint_as_float(n0.z)};
        const float oo[3] = {f.o.x, f.o.y, f.o.z}, id[3] = {f.idir.x, f.idir.y, f.idir.z};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float a = sc[k] * id[k];
            const float b = (pp[k] - oo[k]) * id[k];
            an[k] = a - fabsf(a) * eps; af[k] = a + fabsf(a) * eps;
            bn[k] = b - fabsf(b) * eps; bf[k] = b + fabsf(b) * eps;
        }
    }
    // near/far quantised planes per axis by direction sign
    const bool px = (f.rs & 1u) != 0, py = (f.rs & 2u) != 0, pz = (f.rs & 4u) != 0;
    const uint32_t nx[2] = {px ? n2.x : n3.z, px ? n2.y : n3.w}, fx[2] = {px ? n3.z : n2.x, px ? n3.w : n2.y};
    const uint32_t ny[2] = {py ? n2.z : n4.x, py ? n2.w : n4.y}, fy[2] = {py ? n4.x : n2.z, py ? n4.y : n2.w};
    const uint32_t nz[2] = {pz ? n3.x : n4.z, pz ? n3.y : n4.w}, fz[2] = {pz ? n4.z : n3.x, pz ? n4.w : n3.y};
    const uint32_t valid = n1.z;         // Bvh8Node::valid
    uint32_t hits = 0, leaf_off = 0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int i = 4 * h + j;
            const bool inner = ((valid >> (24 + i)) & 1u) != 0u;
            const uint32_t unary = (valid >> (kLeafStride * (uint32_t)i)) & 3u;     // leaf slot: 01 or 11
            const uint32_t off = leaf_off;
            leaf_off += __popc(unary);
            if (!inner && unary == 0u) continue;
            const float t0x = __fmaf_rn((float)((nx[h] >> (8 * j)) & 0xffu), an[0], bn[0]);
            const float t0y = __fmaf_rn((float)((ny[h] >> (8 * j)) & 0xffu), an[1], bn[1]);
            const float t0z = __fmaf_rn((float)((nz[h] >> (8 * j)) & 0xffu), an[2], bn[2]);
            const float t1x = __fmaf_rn((float)((fx[h] >> (8 * j)) & 0xffu), af[0], bf[0]);
            const float t1y = __fmaf_rn((float)((fy[h] >> (8 * j)) & 0xffu), af[1], bf[1]);
            const float t1z = __fmaf_rn((float)((fz[h] >> (8 * j)) & 0xffu), af[2], bf[2]);
            const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));
            const float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
            if (tn <= tf) {
                if (inner) hits |= 1u << (24u + ((uint32_t)i ^ f.rs));
                else hits |= unary << off;
            }
        }
    }
    return hits;
}

// ------------------------------------------------------------------ full two-level closest hit
// Single-ray traversal; the persistent kernel in render.cu calls this per fetched ray.
// counters (optional): [0] += nodes decoded, [1] += triangles tested.
template <bool kCount>
PT_D void trace_bvh8(const DeviceScene& sc, V3 o, V3 d, float tmin, float tmax, Hit& best, uint32_t* n_nodes, uint32_t* n_tris
) {
    hit_init(best);
    if (sc.tlas_root == kInvalidId) return;
    RayFrame world, f;
    frame_set(world, o, d);
    f = world;
    RayPre rp = ray_precompute(d);   // only used inside a BLAS; recomputed on entry
    uint32_t cur_inst = sc.start_inst;
    constexpr int kStack = 64;
    uint2 stack[kStack];
    int sp = 0;
#define PT_PUSH(v) do { if (sp < kStack) stack[sp++] = (v); else *sc.error_flag = 1u; } while (0)
    const uint2 kSentinel = make_uint2(0xFFFFFFFFu, 0u);
    uint2 G = make_uint2(sc.start_root, 0x80000000u);  // node group: (base, hits<<24 | imask)
    uint2 T = make_uint2(0u, 0u);                      // leaf group: (base, item mask)
    for (;;) {
        if (G.y & 0xFF000000u) {
            const uint32_t k = 31u - (uint32_t)__clz(G.y);
            G.y &= ~(1u << k);
            const uint32_t slot = (k - 24u) ^ f.rs;
            const uint32_t rel = __popc(G.y & 0xffu & ((1u << slot) - 1u));
            const uint32_t node = G.x + rel;
            if (G.y & 0xFF000000u) PT_PUSH(G);
            uint32_t cb, tb, im;
            const uint32_t hits = node_test(sc.nodes, node, f, tmin, fminf(tmax, best.t), cb, tb, im);
            if (kCount) ++*n_nodes;
            G = make_uint2(cb, (hits & 0xFF000000u) | im);
            T = make_uint2(tb, hits & 0x00FFFFFFu);
        } else {
            T = G;
            G = make_uint2(0u, 0u);
        }
        if (cur_inst != kInvalidId) {
            // BLAS leaf items are triangles
            while (T.y) {
                const uint32_t b = (uint32_t)__ffs((int)T.y) - 1u;
                T.y &= T.y - 1u;
                test_triangle(sc.tris, T.x + b, f.o, rp, tmin, tmax, cur_inst, best);
                if (kCount) ++*n_tris;
            }
        } else if (T.y) {
            // TLAS leaf item = instance: enter the first, keep the rest for later
            const uint32_t b = (uint32_t)__ffs((int)T.y) - 1u;
            T.y &= T.y - 1u;
            if (T.y) PT_PUSH(T);
            if (G.y & 0xFF000000u) PT_PUSH(G);
            PT_PUSH(kSentinel);
            cur_inst = __ldg(sc.tlas_leaf_instance + T.x + b);
            V3 oo, dd;
            to_object(sc.instances, cur_inst, o, d, oo, dd);
            frame_set(f, oo, dd);
            rp = ray_precompute(dd);
            G = make_uint2(__ldg(&sc.instances[cur_inst].blas_root), 0x80000000u);
            T = make_uint2(0u, 0u);
            continue;
        }
        if ((G.y & 0xFF000000u) == 0u) {
            // pop
            for (;;) {
                if (sp == 0) return;
                G = stack[--sp];
                if (G.x == 0xFFFFFFFFu && G.y == 0u) { f = world; cur_inst = kInvalidId; continue; }
                break;
            }
        }
    }
}

#undef PT_PUSH

// ------------------------------------------------------------------ BVH2 reference walker
PT_D bool box2_hit(const Bvh2Node& n, const RayFrame& f, float tmin, float tmax) {
    const float eps = 9.5367431640625e-7f;
    float tn = tmin, tf = tmax;
    const float lo[3] = {n.lo[0], n.lo[1], n.lo[2]}, hi[3] = {n.hi[0], n.hi[1], n.hi[2]};
    const float oo[3] = {f.o.x, f.o.y, f.o.z}, id[3] = {f.idir.x, f.idir.y, f.idir.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float a = (lo[k] - oo[k]) * id[k], b = (hi[k] - oo[k]) * id[k];
        float mn = fminf(a, b), mx = fmaxf(a, b);
        mn = mn - fabsf(mn) * 4.0f * eps - 1e-30f; mx = mx + fabsf(mx) * 4.0f * eps + 1e-30f;
        tn = fmaxf(tn, mn); tf = fminf(tf, mx);
    }
    return tn <= tf;
}

PT_D void trace_bvh2(const DeviceScene& sc, uint32_t tlas2_count, const uint32_t* __restrict__ tlas2_leaf_instance, V3 o, V3 d,
                     float tmin, float tmax, Hit& best) {
    hit_init(best);
    if (tlas2_count == 0) return;
    RayFrame world;
    frame_set(world, o, d);
    uint32_t tstack[64]; int tsp = 0;
    tstack[tsp++] = (tlas2_count == 1) ? 0x80000000u : 0u;
    if (tlas2_count == 1) tstack[0] = sc.tlas2[0].left;
    while (tsp) {
        const uint32_t ref = tstack[--tsp];
        if (ref == 0xFFFFFFFFu) continue;
        if (!(ref & 0x80000000u)) {
            const Bvh2Node n = sc.tlas2[ref];
            if (!box2_hit(n, world, tmin, fminf(tmax, best.t))) continue;
            if (tsp + 2 > 64) { *sc.error_flag = 1u; continue; }      // a Karras tree over clustered keys can be deeper than the stack
            tstack[tsp++] = n.left; tstack[tsp++] = n.right;
            continue;
        }
        const uint32_t inst = tlas2_leaf_instance[ref & 0x7fffffffu];
        const InstanceRec& ir = sc.instances[inst];
        V3 oo, dd;
        to_object(sc.instances, inst, o, d, oo, dd);
        RayFrame f; frame_set(f, oo, dd);
        const RayPre rp = ray_precompute(dd);
        const Bvh2Node* nodes = sc.nodes2 + ir.bvh2_root;
        const Bvh8Tri* tris = sc.tris2 + ir.bvh2_tri_base;
        uint32_t stack[64]; int sp = 0;
        stack[sp++] = 0u;
        while (sp) {
            const uint32_t r = stack[--sp];
            if (r == 0xFFFFFFFFu) continue;
            if (r & 0x80000000u) { test_triangle(tris, r & 0x7fffffffu, oo, rp, tmin, tmax, inst, best); continue; }
            const Bvh2Node n = nodes[r];
            if (!box2_hit(n, f, tmin, fminf(tmax, best.t))) continue;
            if (sp + 2 > 64) { *sc.error_flag = 1u; continue; }
            stack[sp++] = n.left; stack[sp++] = n.right;
        }
    }
}

}  // namespace pt
