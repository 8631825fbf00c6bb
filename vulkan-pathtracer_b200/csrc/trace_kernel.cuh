// trace_kernel.cuh — K-trace: the persistent-thread BVH8 traversal kernel of the wavefront loop.
//
// Structure (evidence for each choice is in profiles/ and DESIGN.md §K-trace):
//  * persistent warps; each warp owns a chunk of the ray queue taken with ONE global atomic per 32
//    rays; lanes whose ray has terminated are refilled from the warp's chunk with a ballot (dynamic
//    fetch, Aila & Laine 2009), so a warp never waits for its slowest ray;
//  * every loop iteration a lane does at most one leaf item (triangle test / instance entry) and
//    one node test, in two warp-wide phases, so lanes in the same phase execute together;
//  * node test: 80-byte node as 5 x LDG.128 (nothing else is fetched), 8-bit planes converted with one PRMT each (no I2F
//    on the quarter-rate pipe), branch-free hit-mask assembly from compile-time words + the node's `valid` word,
//    conservative slabs (explicit FMAs);
//  * the traversal stack (node / leaf groups, 8 bytes each) lives in thread-local memory;
//  * the closest hit lives in the slot's hit record, only its distance in a register: 64 registers, 8 CTAs per SM.
// The closest-hit semantics are exactly those of trace_bvh8() in trace.cuh (the simple walker that
// the tests and the algorithmic-byte counter use).
#pragma once
#include "trace.cuh"

namespace pt {

// byte j of `word` as the exact float 2^23 + q (bit pattern 0x4B0000qq): one PRMT, no I2F.  The
// bias is folded into the plane offsets (c = b - 2^23*a, rounded away from the box), so
// t = fma(2^23 + q, a, c) needs no subtraction; the directed rounding of c costs at most one grid
// step of slack per plane.
// `bias` holds 0x4B000000 in a register the compiler cannot fold (it comes from the kernel
// parameters): PRMT then takes the selector as its immediate — with a literal bias ptxas puts the
// literal in the immediate slot and re-materialises the selector with a MOV before every PRMT
// (profiles/r1a: 48 extra MOV/IMAD.U32 per node test).
PT_D float q2f(uint32_t word, uint32_t j, uint32_t bias) {
    return __uint_as_float(__byte_perm(word, bias, 0x7650u | j));
}

// bit s of an 8-bit mask moves to bit s ^ rs (the octant permutation of the inner-child hit bits)
PT_HD uint32_t perm8(uint32_t x, uint32_t rs) {
    if (rs & 1u) x = ((x & 0x55u) << 1) | ((x >> 1) & 0x55u);
    if (rs & 2u) x = ((x & 0x33u) << 2) | ((x >> 2) & 0x33u);
    if (rs & 4u) x = ((x & 0x0Fu) << 4) | ((x >> 4) & 0x0Fu);
    return x;
}

// Node test of K-trace.  Hit-mask assembly: a child that passes the slab test contributes a compile-time word — its inner bit
// 24 + slot and both bits of its leaf slot — under one FSETP + SEL, and one AND with the node's `valid` word keeps what the slot
// really holds; the inner byte is brought into traversal priority order (bit 24 + (slot ^ rs)) by one shared-memory lookup.
// (Rounds 1-2 loaded the eight words per node from a second array, `node_bits`: two more LDG.128 that ptxas sank below the plane
// arithmetic — a second memory round trip per node test — and 32 of the 112 bytes a node test fetched.)
// Returns bits 31..24 = inner hits in priority order, bits 15..0 = leaf hits in the slot-strided positions of `valid`; `valid`
// itself is returned for the triangle phase, which turns a bit position into a dense item index with one POPC.
PT_D uint32_t node_test_fast(const Bvh8Node* __restrict__ nodes, uint32_t idx, const RayFrame& f,
                             float tmin, float tmax, uint32_t bias, const uint8_t* __restrict__ perm_lut, uint32_t& child_base,
                             uint32_t& tri_base, uint32_t& imask, uint32_t& valid) {
    const uint4* q = reinterpret_cast<const uint4*>(nodes + idx);
    const uint4 n0 = __ldg(q), n1 = __ldg(q + 1), n2 = __ldg(q + 2), n3 = __ldg(q + 3), n4 = __ldg(q + 4);
    child_base = n1.x; tri_base = n1.y; imask = n0.w >> 24; valid = n1.z;
    const float eps = 9.5367431640625e-7f;  // 2^-20
    float anx, any, anz, afx, afy, afz, bnx, bny, bnz, bfx, bfy, bfz;
    {
        const float sx = __uint_as_float((n0.w & 0xffu) << 23), sy = __uint_as_float(((n0.w >> 8) & 0xffu) << 23),
                    sz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23);
        const float ax = sx * f.idir.x, ay = sy * f.idir.y, az = sz * f.idir.z;
        const float bx = (__uint_as_float(n0.x) - f.o.x) * f.idir.x, by = (__uint_as_float(n0.y) - f.o.y) * f.idir.y,
                    bz = (__uint_as_float(n0.z) - f.o.z) * f.idir.z;
        const float B23 = 8388608.0f;
        // slack of 2^-20 relative on slope and offset, then the 2^23 bias folded into the offset with directed rounding.
        // Explicit FMAs: |v| * 2^-20 and 2^23 * a are exact products (powers of two), so each FMA rounds once, exactly where
        // the FMUL + FADD pair of the walker (trace.cuh:node_test) rounds — same planes, 12 instructions fewer per node
        // (profiles/r2o_variants.txt: trace C2 1.126 -> 1.107 ms, C4 14.22 -> 14.13).
        anx = __fmaf_rn(-fabsf(ax), eps, ax); afx = __fmaf_rn(fabsf(ax), eps, ax);
        any = __fmaf_rn(-fabsf(ay), eps, ay); afy = __fmaf_rn(fabsf(ay), eps, ay);
        anz = __fmaf_rn(-fabsf(az), eps, az); afz = __fmaf_rn(fabsf(az), eps, az);
        bnx = __fmaf_rd(-B23, anx, __fmaf_rn(-fabsf(bx), eps, bx)); bfx = __fmaf_ru(-B23, afx, __fmaf_rn(fabsf(bx), eps, bx));
        bny = __fmaf_rd(-B23, any, __fmaf_rn(-fabsf(by), eps, by)); bfy = __fmaf_ru(-B23, afy, __fmaf_rn(fabsf(by), eps, by));
        bnz = __fmaf_rd(-B23, anz, __fmaf_rn(-fabsf(bz), eps, bz)); bfz = __fmaf_ru(-B23, afz, __fmaf_rn(fabsf(bz), eps, bz));
    }
    const bool px = (f.rs & 1u) != 0, py = (f.rs & 2u) != 0, pz = (f.rs & 4u) != 0;
    uint32_t hits = 0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        // words holding children 4h..4h+3
        const uint32_t lox = h ? n2.y : n2.x, loy = h ? n2.w : n2.z, loz = h ? n3.y : n3.x;
        const uint32_t hix = h ? n3.w : n3.z, hiy = h ? n4.y : n4.x, hiz = h ? n4.w : n4.z;
        const uint32_t nxw = px ? lox : hix, fxw = px ? hix : lox;
        const uint32_t nyw = py ? loy : hiy, fyw = py ? hiy : loy;
        const uint32_t nzw = pz ? loz : hiz, fzw = pz ? hiz : loz;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float t0x = __fmaf_rn(q2f(nxw, j, bias), anx, bnx), t0y = __fmaf_rn(q2f(nyw, j, bias), any, bny), t0z = __fmaf_rn(q2f(nzw, j, bias), anz, bnz);
            const float t1x = __fmaf_rn(q2f(fxw, j, bias), afx, bfx), t1y = __fmaf_rn(q2f(fyw, j, bias), afy, bfy), t1z = __fmaf_rn(q2f(fzw, j, bias), afz, bfz);
            const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));
            const float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
            const uint32_t slot_bits = (1u << (24 + 4 * h + j)) | (3u << (kLeafStride * (4 * h + j)));     // immediate
            hits |= (tn <= tf) ? slot_bits : 0u;
        }
    }
    hits &= valid;
    const uint32_t top = perm_lut[(f.rs << 8) | (hits >> 24)];
    return (top << 24) | (hits & 0x0000FFFFu);
}

struct TraceIO {
    const uint32_t* queue;      // slot per queue entry, or nullptr for identity
    const uint32_t* count_ptr;  // number of entries (device), or nullptr -> count
    uint32_t count;
    uint32_t* cursor;           // dynamic-fetch cursor (zeroed before the launch)
    // per slot, written by the producer of the ray (store_ray(): K-raygen, K-shade, k_ray_setup) at
    // full SIMD width, so that a refilled lane only loads:
    const float4* ray_o;        // origin.xyz, tmin
    const float4* ray_d;        // direction.xyz, tmax
    const float4* ray_e;        // 1/direction (box tests), bits: octant | kz<<3
    const float4* ray_s;        // shear row m0.xyz, Sz
    const float4* ray_m;        // shear row m1.xyz, tmax
    float4* hit_a;              // per slot: t, b1, b2, instance
    uint2* hit_b;               // per slot: primitive_index, triangle_index
    uint32_t hit_a_stride;      // in float4: 1 = SoA, 2 = HitRec array
    uint32_t hit_b_stride;      // in uint2:  1 = SoA, 4 = HitRec array (hit_b = base + 2)
    uint32_t bias_word;
    uint32_t chunk;          // rays a warp takes per global atomic         // 0x4B000000 (see q2f)
    // primary wavefront of a frame: every ray starts at the camera, so K-raygen does not write origins and the fetch takes
    // (origin.xyz, tmin) from here (C2: 33 MB less written by K-raygen and 33 MB less read here per frame)
    uint32_t uniform_origin; // 1: use `origin` instead of ray_o[slot] (flat instantiation only)
    float4 origin;
};

constexpr int kTraceBlock = 128;
#ifndef PT_TRACE_MIN_BLOCKS
#define PT_TRACE_MIN_BLOCKS 8
#endif
constexpr int kTraceMinBlocks = PT_TRACE_MIN_BLOCKS;       // 8: 64 registers, 32 resident warps per SM (see test_triangle_rec)

// Triangle test of K-trace.  The closest hit is kept in the slot's hit record (hit_a / hit_b, global memory) and only its distance
// in a register; bit 31 of `slot` says that a record has been written.  Five registers less across the loop than a Hit in
// registers, which is what lets the kernel fit 64 registers without spills (8 CTAs per SM instead of 7): accepted candidates are
// rare (2-3 per ray, 2-4 lanes at a time), so the extra stores cost less than the occupancy gains.  Measured on B200
// (profiles/r2t_variants.txt, r2u_variants.txt): K-trace C2 1.105 -> 1.073 ms, C3 606 -> 573 ms, C4 14.13 -> 13.60 ms,
// C5 43.5 -> 41.6 ms; the record in memory alone, still at 7 CTAs: C2 1.117 ms.
// Same acceptance rule as test_triangle() / hit_better(): lexicographic (t, instance, primitive, triangle), and the ray's own
// tmax (not the closest distance) bounds the test so that candidates with t == best_t reach the tie-break.
PT_D void test_triangle_rec(const Bvh8Tri* __restrict__ tris, uint32_t idx, V3 o, const RayPre& rp, float tmin, float tmax, uint32_t inst,
                            float& best_t, uint32_t& slot, float4* hit_a, uint2* hit_b) {
    const float4* q = reinterpret_cast<const float4*>(tris + idx);
    const float4 p0 = __ldg(q), p1 = __ldg(q + 1), p2 = __ldg(q + 2);
    float t, V, W, det;
    if (intersect_tri_raw(o, rp, p0, p1, p2, tmin, tmax, t, V, W, det)) {
        const uint32_t prim = __float_as_uint(p0.w), tri = __float_as_uint(p1.w);
        bool better = t < best_t;
        if (t == best_t) {
            if (!(slot & 0x80000000u)) better = true;       // nothing recorded yet: the invalid ids of hit_init() lose every tie
            else {
                // this thread's own earlier stores (program order); volatile keeps the compiler from using the read-only path
                const volatile uint32_t* pa = reinterpret_cast<const volatile uint32_t*>(hit_a);
                const volatile uint32_t* pb = reinterpret_cast<const volatile uint32_t*>(hit_b);
                const uint32_t binst = pa[3], bprim = pb[0], btri = pb[1];
                better = inst != binst ? inst < binst : (prim != bprim ? prim < bprim : tri < btri);
            }
        }
        if (better) {
            // the two barycentric divisions only for candidates that replace the closest hit (trace.cuh:test_triangle)
            best_t = t; slot |= 0x80000000u;
            *hit_a = make_float4(t, V / det, W / det, __uint_as_float(inst));
            *hit_b = make_uint2(prim, tri);
        }
    }
}
constexpr int kStackSize = 64;
#ifdef PT_TRACE_REFILL_MIN
constexpr int kRefillMin = PT_TRACE_REFILL_MIN;
#else
constexpr int kRefillMin = 4;
#endif

// world-space ray frame + shear constants of one slot (what store_ray() wrote)
// The flat kernel fetches four of the five records (64 B): the direction itself is only needed on entry into a transformed
// instance, and tmax rides in ray_m.w too (K-trace C2 1.068 -> 1.065 ms, C3 554.7 -> 548.4, C5 39.9 -> 39.6).  The two-level
// kernel keeps the five-record fetch: without it C4 measured 13.31 -> 13.45 ms (profiles/r2z_variants.txt).
template <bool kFlat>
PT_D void load_ray(const TraceIO& io, uint32_t slot, RayFrame& f, RayPre& rp, float& tmin, float& tmax) {
    const float4 re = io.ray_e[slot], rs = io.ray_s[slot], rm = io.ray_m[slot];
    // (flat kernel only: in the two-level kernel the extra select measured C4 13.28 -> 13.68 ms, profiles/r3a_variants.txt)
    const float4 ro = (kFlat && io.uniform_origin) ? io.origin : io.ray_o[slot];
    tmax = kFlat ? rm.w : io.ray_d[slot].w;
    f.o = v3(ro.x, ro.y, ro.z); f.idir = v3(re.x, re.y, re.z);
    const uint32_t bits = __float_as_uint(re.w);
    f.rs = bits & 7u;
    rp.kz = (int)((bits >> 3) & 3u);
    rp.m0x = rs.x; rp.m0y = rs.y; rp.m0z = rs.z; rp.Sz = rs.w;
    rp.m1x = rm.x; rp.m1y = rm.y; rp.m1z = rm.z;
    tmin = ro.w;
}

#define PT_TRACE_BOUNDS __launch_bounds__(kTraceBlock, kTraceMinBlocks)
// kFlat: the scene is one identity instance (DeviceScene::start_inst is set: C1, C2, C3, C5 and most single-mesh glTF files), so
// traversal never sees the top level: leaf items are always triangles, the stack holds node groups only and there is no
// instance entry / exit.  Its own instantiation keeps that code out of the hot loop.
template <bool kFlat>
__global__ void PT_TRACE_BOUNDS k_trace_persistent(DeviceScene sc, TraceIO io) {
    const uint32_t n = io.count_ptr ? *io.count_ptr : io.count;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    if (sc.tlas_root == kInvalidId) {
        // no geometry: every ray misses
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
            const uint32_t slot = io.queue ? io.queue[i] : i;
            io.hit_a[(size_t)slot * io.hit_a_stride] = make_float4(kMissT, 0.0f, 0.0f, __uint_as_float(kInvalidId));
            io.hit_b[(size_t)slot * io.hit_b_stride] = make_uint2(kInvalidId, kInvalidId);
        }
        return;
    }
    const uint32_t bias = io.bias_word;
    // octant permutation of the inner-child hit byte: lut[rs * 256 + mask]
    __shared__ uint8_t perm_lut[8 * 256];
    for (uint32_t i = threadIdx.x; i < 8u * 256u; i += blockDim.x) perm_lut[i] = (uint8_t)perm8(i & 255u, i >> 8);
    __syncthreads();
    // warp-uniform chunk of the queue: 32 rays per global atomic.  Larger chunks save atomics but leave the last
    // round of a wavefront unbalanced (a 1 M-ray wavefront is only 241 rays per warp); measured 8 / 16 / 32 / 64 / 128:
    // C2 -- / 1.371 / 1.363 / 1.386 / 1.440 ms, C4 15.25 / 15.10 / 15.21 / 15.69 / 17.00 ms, C1 0.117 / 0.101 / 0.089 /
    // 0.103 / 0.143 ms.
    const uint32_t chunk = io.chunk;
    uint32_t w_next = 0, w_end = 0;
    bool exhausted = false;
    // per-lane ray state: ONE frame (world space at TLAS level, object space inside a transformed
    // instance; the world frame is re-loaded from the slot when such an instance is left)
    bool active = false;
    uint32_t slot = 0;                   // bit 31: a hit record has been written for this ray (slots are < 2^31)
#define PT_SLOT (slot & 0x7fffffffu)
    RayFrame f;
    RayPre rp;
    float tmin = 0.0f, tmax = 0.0f;
    uint32_t cur_inst = kInvalidId;
    float best_t = kMissT;               // distance of the closest hit; the hit itself is in the slot's record (test_triangle_rec)
    // G = node group: (child_base, inner hits << 24 | imask).  T = leaf group: (tri_base, pending << 16 | valid16): the pending
    // items as slot-strided bits (Bvh8Node::valid positions) above the node's valid bits, which turn a position into the dense
    // item index.  "Items pending" is T.y >= 0x10000; a stacked T always has a bit in 8..23 set, a stacked G never.
    uint2 G = make_uint2(0u, 0u), T = make_uint2(0u, 0u);
    uint2 stack[kStackSize];
#define PT_STACK_PUSH(v) do { if (sp < kStackSize) stack[sp++] = (v); else *sc.error_flag = 1u; } while (0)
#define PT_STACK_POP(e) do { (e) = stack[--sp]; } while (0)
    int sp = 0;
    f.o = v3s(0.0f); f.idir = v3s(1.0f); f.rs = 7u;
    rp.m0x = 1.0f; rp.m0y = 0.0f; rp.m0z = 0.0f; rp.m1x = 0.0f; rp.m1y = 1.0f; rp.m1z = 0.0f; rp.Sz = 1.0f; rp.kz = 2;

    for (;;) {
        // ---- dynamic fetch: refill idle lanes from the warp's chunk
        const uint32_t idle = __ballot_sync(0xffffffffu, !active);
        // refill once kRefillMin lanes are idle: the fetch path runs at a few lanes (5-7 % of K-trace's warp instructions when
        // every idle lane is refilled at once), waiting trades it against idle lanes in the node phase.  Measured on B200
        // (profiles/r2a_variants.txt, threshold 1 / 4 / 8 / 16): C2 1.334 / 1.290 / 1.302 / 1.354 ms, C4 14.98 / 14.51 / 14.71 /
        // 15.60 ms, one rank's share of C3 at 8 GPUs 116.7 / 114.1 / 115.3 / 122.7 ms.  Results do not depend on it.
        if (__popc(idle) >= kRefillMin && !exhausted) {
            if (w_next == w_end) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(io.cursor, chunk);
                base = __shfl_sync(0xffffffffu, base, 0);
                if (base >= n) exhausted = true;
                else { w_next = base; w_end = min(base + chunk, n); }
            }
            if (!exhausted) {
                const uint32_t want = __popc(idle);
                const uint32_t avail = w_end - w_next;
                if (!active) {
                    const uint32_t rank = __popc(idle & lt_mask);
                    if (rank < avail) {
                        const uint32_t qi = w_next + rank;
                        slot = io.queue ? io.queue[qi] : qi;
                        load_ray<kFlat>(io, slot, f, rp, tmin, tmax);
                        cur_inst = sc.start_inst;
                        best_t = kMissT;
                        sp = 0;
                        G = make_uint2(sc.start_root, 0x80000000u);
                        T = make_uint2(0u, 0u);
                        active = true;
                    }
                }
                w_next += min(want, avail);
            }
        }
        if (idle == 0xffffffffu && __ballot_sync(0xffffffffu, active) == 0u) {
            if (exhausted) break;
            continue;
        }
        // ---- phase 1: one leaf item.  (Postponing triangle groups while few lanes hold one — Ylitie et
        // al. 2017 — and repeating the phase while many do were both measured: neither pays here, and the
        // warp votes they need cost 3-4 % each; profiles/r1_experiments.md.)
        const bool has_t = active && T.y >= 0x10000u;
        if (has_t) {
            // next pending item = lowest pending bit, at position b < 16 of the strided layout; its dense index is the number of
            // valid bits below b.  Decrementing the pending half flips bit 16 + b and the pending zeros below it, so
            // (T.y ^ (T.y - 0x10000)) >> 17 == (1 << b) - 1 without a find-first-set (one POPC is the only slow-pipe instruction
            // ahead of the triangle loads; FLO + POPC measured C2 +0.6 %, profiles/r2w_variants.txt)
            const uint32_t dec = T.y - 0x10000u;
            const uint32_t item = T.x + (uint32_t)__popc(T.y & ((T.y ^ dec) >> 17));
            T.y &= dec;                                                         // clears the lowest pending bit
            if (kFlat || cur_inst != kInvalidId) {
                test_triangle_rec(sc.tris, item, f.o, rp, tmin, tmax, cur_inst, best_t, slot, io.hit_a + (size_t)PT_SLOT * io.hit_a_stride,
                                  io.hit_b + (size_t)PT_SLOT * io.hit_b_stride);
            } else {
                // TLAS leaf item = instance: keep the remaining work for later and enter the BLAS
                if (T.y >= 0x10000u) PT_STACK_PUSH(T);
                if (G.y & 0xFF000000u) PT_STACK_PUSH(G);
                cur_inst = __ldg(sc.tlas_leaf_instance + item);
                const uint4 ent = __ldg(reinterpret_cast<const uint4*>(&sc.instances[cur_inst].blas_root));   // blas_root, -, mesh | flags, -
                const bool identity = (ent.z & kInstIdentity) != 0u;
                // sentinel: y = 1 when the world frame must be re-loaded on the way out
                PT_STACK_PUSH(make_uint2(0xFFFFFFFFu, identity ? 0u : 1u));
                if (!identity) {
                    V3 oo, dd;
                    const float4 wd = io.ray_d[PT_SLOT];      // the world direction is not kept in registers
                    to_object(sc.instances, cur_inst, f.o, v3(wd.x, wd.y, wd.z), oo, dd);
                    frame_set(f, oo, dd);
                    rp = ray_precompute(dd);
                }
                G = make_uint2(ent.x, 0x80000000u);
                T = make_uint2(0u, 0u);
            }
        }
        // ---- pop when there is nothing else to do
        if (active && T.y < 0x10000u && (G.y & 0xFF000000u) == 0u) {
            for (;;) {
                if (sp == 0) {
                    if (!(slot & 0x80000000u)) {         // no hit was recorded: the miss record of hit_init()
                        io.hit_a[(size_t)slot * io.hit_a_stride] = make_float4(kMissT, 0.0f, 0.0f, __uint_as_float(kInvalidId));
                        io.hit_b[(size_t)slot * io.hit_b_stride] = make_uint2(kInvalidId, kInvalidId);
                    }
                    active = false;
                    break;
                }
                uint2 e;
                PT_STACK_POP(e);
                if (kFlat) { G = e; break; }
                if (e.x == 0xFFFFFFFFu) {
                    if (e.y) { float t0, t1; load_ray<kFlat>(io, PT_SLOT, f, rp, t0, t1); }
                    cur_inst = kInvalidId;
                    continue;
                }
                if (e.y & 0x00FFFF00u) { T = e; } else { G = e; }
                break;
            }
        }
        // ---- phase 2: one node test
        if (active && T.y < 0x10000u && (G.y & 0xFF000000u) != 0u) {
            const uint32_t k = 31u - (uint32_t)__clz(G.y);
            G.y &= ~(1u << k);
            const uint32_t s = (k - 24u) ^ f.rs;
            const uint32_t node = G.x + __popc(G.y & 0xffu & ((1u << s) - 1u));
            if (G.y & 0xFF000000u) PT_STACK_PUSH(G);
            uint32_t cb, tb, im, valid;
            const uint32_t hits = node_test_fast(sc.nodes, node, f, tmin, fminf(tmax, best_t), bias, perm_lut, cb, tb, im, valid);
            G = make_uint2(cb, (hits & 0xFF000000u) | im);
            T = make_uint2(tb, __byte_perm(valid, hits, 0x5410u));      // leaf hits << 16 | valid16
        }
    }
}

#undef PT_STACK_PUSH
#undef PT_STACK_POP
#undef PT_SLOT

}  // namespace pt
