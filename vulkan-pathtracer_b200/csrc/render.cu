// render.cu — the wavefront trace loop: K-raygen, K-trace, K-shade (+ fused queue compaction and
// pixel accumulation) and the image post-processing kernels.  Restates
// src/shaders/ray_gen.rgen:33-105 as a per-pixel state machine driven by ray queues.
#include <cstdlib>
#include <cstring>

#include "render.h"
#include "shade.cuh"
#include "trace.cuh"
#include "trace_kernel.cuh"

namespace pt {

namespace {

constexpr int kTraceThreads = 128;
constexpr int kShadeThreads = 128;

PT_D bool slot_to_pixel(const PixelMap& pm, uint32_t slot, uint32_t& x, uint32_t& y) {
    const uint32_t k = slot >> 8, within = slot & 255u;
    const uint32_t tile = pm.rank + k * pm.world;
    const uint32_t tx = tile % pm.tiles_x, ty = tile / pm.tiles_x;
    const uint32_t w = within >> 5, l = within & 31u;
    x = tx * 32u + (w & 3u) * 8u + (l & 7u);
    y = ty * 8u + (w >> 2) * 4u + (l >> 3);
    return ty < pm.tiles_y && x < pm.width && y < pm.height;
}

// ray_gen.rgen:34-36,48
PT_D V3 primary_direction(const PtCamera& cam, uint32_t x, uint32_t y, uint32_t W, uint32_t H) {
    const float px = (float)x + 0.5f, py = (float)y + 0.5f;
    const float u = px / (float)W, v = py / (float)H;
    const float dx = u * 2.0f - 1.0f, dy = v * 2.0f - 1.0f;
    const V3 hor = v3(cam.horizontal[0], cam.horizontal[1], cam.horizontal[2]);
    const V3 ver = v3(cam.vertical[0], cam.vertical[1], cam.vertical[2]);
    const V3 fwd = v3(cam.forward[0], cam.forward[1], cam.forward[2]);
    return normalize(fwd + hor * dx - ver * dy);
}

// warp-aggregated append of `slot` to a queue; must be reached by all 32 lanes of the warp
PT_D void queue_push(uint32_t* __restrict__ queue, uint32_t* __restrict__ count, bool pred, uint32_t slot) {
    const uint32_t active = 0xffffffffu;
    const uint32_t mask = __ballot_sync(active, pred);
    if (mask == 0u) return;
    const uint32_t lane = threadIdx.x & 31u;
    const int leader = __ffs((int)mask) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(count, (uint32_t)__popc(mask));
    base = __shfl_sync(active, base, leader);
    if (pred) queue[base + __popc(mask & ((1u << lane) - 1u))] = slot;
}

PT_D void write_pixel(const FrameParams& fp, uint32_t W, uint32_t x, uint32_t y, V3 final_color) {
    const float s = (float)fp.num_samples;
    const V3 c = final_color / s;                                       // ray_gen.rgen:102
    reinterpret_cast<float4*>(fp.image)[(size_t)y * W + x] = make_float4(c.x, c.y, c.z, 1.0f);  // :104
}

// ------------------------------------------------------------------ K-raygen  (ray_gen.rgen:34-48)
__global__ void __launch_bounds__(256) k_raygen(PixelMap pm, PathBuffers pb, FrameParams fp) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x = 0, y = 0;
    bool valid = slot < pm.num_slots && slot_to_pixel(pm, slot, x, y);
    bool live = valid;
    if (valid) {
        if (fp.num_samples == 0u || fp.num_bounces == 0u) {
            write_pixel(fp, pm.width, x, y, v3s(0.0f));
            const size_t pi = (size_t)y * pm.width + x;
            fp.id_inst[pi] = kInvalidId; fp.id_prim[pi] = kInvalidId; fp.id_tri[pi] = kInvalidId;
            live = false;
        } else {
            // the path state (seed :41, final :43, color :45, attenuation :46) is initialised by K-shade of
            // iteration 0, which saves 48 B written here and 48 B read there per pixel
            const V3 d = primary_direction(fp.cam, x, y, pm.width, pm.height);
            store_ray(pb.ray_o, pb.ray_d, pb.ray_e, pb.ray_s, pb.ray_m, slot, v3(fp.cam.position[0], fp.cam.position[1], fp.cam.position[2]), d,
                      0.001f, 10000.0f, fp.primary_lean);                                                        // :38-39,47-48
        }
    }
    // block-aggregated append: one atomic per 256 slots.  A full-HD frame is 64 800 warps; one same-address atomic per warp
    // (queue_push) serialises in L2 for about as long as the kernel needs to write its rays.
    __shared__ uint32_t s_off[9];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t mask = __ballot_sync(0xffffffffu, live);
    if (lane == 0u) s_off[warp] = (uint32_t)__popc(mask);
    __syncthreads();
    if (threadIdx.x == 0u) {
        uint32_t total = 0;
        for (int w = 0; w < 8; ++w) { const uint32_t c = s_off[w]; s_off[w] = total; total += c; }
        s_off[8] = total ? atomicAdd(pb.qcount + 0, total) : 0u;
    }
    __syncthreads();
    if (live) pb.queue[0][s_off[8] + s_off[warp] + (uint32_t)__popc(mask & ((1u << lane) - 1u))] = slot;
}

// ------------------------------------------------------------------ K-trace
// Persistent warps; every warp fetches batches of 32 queue entries through one atomic.
template <int MODE>
__global__ void __launch_bounds__(kTraceThreads) k_trace(DeviceScene sc, PathBuffers pb, uint32_t iter, uint32_t tlas2_count,
                                                         const uint32_t* __restrict__ tlas2_leaf_instance) {
    const uint32_t n = pb.qcount[iter];
    const uint32_t* __restrict__ queue = pb.queue[iter & 1u];
    uint32_t* cursor = pb.trace_next + iter;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t cnt_nodes = 0, cnt_tris = 0;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(cursor, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const uint32_t idx = base + lane;
        if (idx < n) {
            const uint32_t slot = queue[idx];
            const float4 ro = pb.ray_o[slot], rd = pb.ray_d[slot];
            Hit h;
            if (MODE == TRACE_BVH2) trace_bvh2(sc, tlas2_count, tlas2_leaf_instance, v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), ro.w, rd.w, h);
            else if (MODE == TRACE_BVH8_COUNT) trace_bvh8<true>(sc, v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), ro.w, rd.w, h, &cnt_nodes, &cnt_tris);
            else trace_bvh8<false>(sc, v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), ro.w, rd.w, h, nullptr, nullptr);
            pb.hit_a[slot] = make_float4(h.t, h.b1, h.b2, __uint_as_float(h.inst));
            pb.hit_b[slot] = make_uint2(h.prim, h.tri);
        }
    }
    if (MODE == TRACE_BVH8_COUNT) {
        atomicAdd(reinterpret_cast<unsigned long long*>(pb.trav_counters), (unsigned long long)cnt_nodes);
        atomicAdd(reinterpret_cast<unsigned long long*>(pb.trav_counters) + 1, (unsigned long long)cnt_tris);
    }
}

template <int MODE>
__global__ void __launch_bounds__(kTraceThreads) k_trace_rays(DeviceScene sc, const RayRec* __restrict__ rays, uint32_t n,
                                                              HitRec* __restrict__ hits, uint32_t tlas2_count,
                                                              const uint32_t* __restrict__ tlas2_leaf_instance, uint32_t* cursor) {
    const uint32_t lane = threadIdx.x & 31u;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(cursor, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const uint32_t idx = base + lane;
        if (idx < n) {
            const RayRec r = rays[idx];
            Hit h;
            if (MODE == TRACE_BVH2) trace_bvh2(sc, tlas2_count, tlas2_leaf_instance, v3(r.ox, r.oy, r.oz), v3(r.dx, r.dy, r.dz), r.tmin, r.tmax, h);
            else trace_bvh8<false>(sc, v3(r.ox, r.oy, r.oz), v3(r.dx, r.dy, r.dz), r.tmin, r.tmax, h, nullptr, nullptr);
            HitRec o;
            o.t = h.t; o.b1 = h.b1; o.b2 = h.b2; o.instance = h.inst; o.primitive_index = h.prim; o.triangle_index = h.tri;
            o._pad0 = 0; o._pad1 = 0;
            hits[idx] = o;
        }
    }
}

// standalone queries (pt_trace_rays): RayRec array -> the SoA ray records K-trace fetches
__global__ void __launch_bounds__(256) k_ray_setup(const RayRec* __restrict__ rays, uint32_t n, float4* __restrict__ ro, float4* __restrict__ rd,
                                                   float4* __restrict__ re, float4* __restrict__ rs, float4* __restrict__ rm) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const RayRec r = rays[i];
    store_ray(ro, rd, re, rs, rm, i, v3(r.ox, r.oy, r.oz), v3(r.dx, r.dy, r.dz), r.tmin, r.tmax);
}

// ------------------------------------------------------------------ material sort (optional)
// Counting sort of the live-path queue by material index (24 bit, triangle.glsl:107) so that the
// lanes of a K-shade warp read the same MatRec and the same textures.  Three small kernels: count,
// scan (one block), scatter.  Order inside a bin is arbitrary (atomics); results do not depend on it.
PT_D uint32_t path_material(const DeviceScene& sc, const PathBuffers& pb, uint32_t slot) {
    const float4 ha = pb.hit_a[slot];
    if (ha.x == kMissT) return sc.num_materials;           // sky paths: last bin
    const uint32_t prim = pb.hit_b[slot].x;
    return min(__ldg(&sc.prims[prim].info) & PT_MATERIAL_INDEX_MASK, sc.num_materials);
}

__global__ void __launch_bounds__(256) k_matsort_count(DeviceScene sc, PathBuffers pb, uint32_t iter) {
    const uint32_t n = pb.qcount[iter];
    const uint32_t* __restrict__ q = pb.queue[iter & 1u];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t key = path_material(sc, pb, q[i]);
        // warp-aggregated increment per distinct key
        const uint32_t peers = __match_any_sync(__activemask(), key);
        if ((threadIdx.x & 31u) == (uint32_t)(__ffs((int)peers) - 1)) atomicAdd(pb.sort_bins + key, (uint32_t)__popc(peers));
    }
}

__global__ void __launch_bounds__(1024) k_matsort_scan(PathBuffers pb, uint32_t bins) {
    // exclusive scan of the bins in place (bins <= a few thousand: one block, chunked)
    __shared__ uint32_t carry;
    __shared__ uint32_t warp_sums[32];
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < bins; base += blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < bins ? pb.sort_bins[i] : 0u;
        uint32_t x = v;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((int)(threadIdx.x & 31u) >= o) x += y; }
        if ((threadIdx.x & 31u) == 31u) warp_sums[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_sums[threadIdx.x];
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, w, o); if ((int)threadIdx.x >= o) w += y; }
            warp_sums[threadIdx.x] = w;
        }
        __syncthreads();
        const uint32_t before = carry + (threadIdx.x >= 32 ? warp_sums[(threadIdx.x >> 5) - 1] : 0u) + x - v;
        if (i < bins) pb.sort_bins[i] = before;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = before + v;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k_matsort_scatter(DeviceScene sc, PathBuffers pb, uint32_t iter) {
    const uint32_t n = pb.qcount[iter];
    const uint32_t* __restrict__ q = pb.queue[iter & 1u];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t slot = q[i];
        const uint32_t key = path_material(sc, pb, slot);
        const uint32_t peers = __match_any_sync(__activemask(), key);
        const uint32_t lane = threadIdx.x & 31u;
        const int leader = __ffs((int)peers) - 1;
        uint32_t base = 0;
        if ((int)lane == leader) base = atomicAdd(pb.sort_bins + key, (uint32_t)__popc(peers));
        base = __shfl_sync(peers, base, leader);
        pb.queue_sorted[base + __popc(peers & ((1u << lane) - 1u))] = slot;
    }
}

// ------------------------------------------------------------------ K-shade
// One thread per live path.  Consumes the hit of segment (sample s, bounce j), runs the body of the
// bounce loop (ray_gen.rgen:65-97), and either emits the next ray (appended to the next queue) or
// ends the sample; a finished sample restarts from the cached primary hit (every sample of a pixel
// shoots the same primary ray, quirk Q1) until NUM_SAMPLES are done, then the pixel is stored.
// Two instantiations.  kMulti = false (one sample per pixel, C2): no primary-surface cache, no restart-from-cache
// path; 80 registers / 6 CTAs per SM measured best (C2 shade 4 / 5 / 6 / 8 blocks: 0.362 / 0.363 / 0.339 / 0.346 ms).
// kMulti = true (S > 1): restarts read the cached primary surface; here spills are what hurts (C3 shade at
// 4 / 5 / 6 blocks: 656 / 798 / 846 ms), so 128 registers / 4 CTAs per SM.
#ifdef PT_SHADE_MIN_BLOCKS
#define PT_SHADE_BOUNDS __launch_bounds__(kShadeThreads, PT_SHADE_MIN_BLOCKS)
#else
#ifndef PT_SHADE_MULTI_MIN_BLOCKS
#define PT_SHADE_MULTI_MIN_BLOCKS 5     // 96 registers, no spills in the atlas-only instantiation; 4 -> 5 CTAs per SM measured on B200
#endif                                  // (profiles/r2e_variants.txt): K-shade C4 3.23 -> 3.13 ms, C3 share 63.7 -> 60.3 ms, C5 25.8 -> 23.7 ms
#define PT_SHADE_BOUNDS __launch_bounds__(kShadeThreads, kMulti ? PT_SHADE_MULTI_MIN_BLOCKS : 6)
#endif
// kTex (shade.cuh): how material maps are fetched — 0 general, 1 texture objects (PT_FLAG_HW_BILINEAR), 2 atlas-only scene.  K-shade is
// instruction-cache-bound on multi-sample workloads, so each mode is its own instantiation and carries only its own fetch code.
template <bool kMulti, int kTex = 0>
__global__ void PT_SHADE_BOUNDS k_shade(DeviceScene sc, PixelMap pm, PathBuffers pb, FrameParams fp, uint32_t iter) {
    const uint32_t n = pb.qcount[iter];
    const uint32_t* __restrict__ qin = (fp.flags & PT_FLAG_SORT_BY_MATERIAL) ? pb.queue_sorted : pb.queue[iter & 1u];
    uint32_t* __restrict__ qout = pb.queue[(iter + 1u) & 1u];
    uint32_t* qout_count = pb.qcount + iter + 1u;
    const uint32_t S = fp.num_samples, B = fp.num_bounces;
    // whole warps iterate together (queue_push is warp-collective over the active mask); 32 entries per atomic, taken
    // dynamically: the cost of an entry varies from a sky lookup to full material + BSDF, so a static split leaves
    // warps idle at the end of small wavefronts
    // unorm8 -> float table for the texel and factor conversions (48 per fully textured hit): x / 255.0f, IEEE
    __shared__ float s_unorm[256];
    for (uint32_t i = threadIdx.x; i < 256u; i += blockDim.x) s_unorm[i] = (float)i / 255.0f;
    __syncthreads();
    uint32_t* cursor = pb.shade_next + iter;
    uint32_t n_hit = 0;     // entries of this thread whose traced segment hit geometry (PtStats::rays_hit)
    for (;;) {
        uint32_t base = 0;
        if ((threadIdx.x & 31u) == 0u) base = atomicAdd(cursor, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const uint32_t idx = base + (threadIdx.x & 31u);
        bool emit = false;
        uint32_t slot = 0;
        if (idx < n) {
            slot = qin[idx];
            uint32_t x, y;
            slot_to_pixel(pm, slot, x, y);
            V3 atten = v3s(1.0f), color = v3s(0.0f), final_color = v3s(0.0f);            // :46, :45, :43
            U2 seed;
            uint32_t s = 0u, j = 0u;
            if (iter == 0u) {
                seed = prng_seed(x, y, fp.cam.frame_count);                                  // :41
            } else {
                const float4 a4 = pb.atten_sx[slot], c4 = pb.color_sy[slot];
                atten = v3(a4.x, a4.y, a4.z); color = v3(c4.x, c4.y, c4.z);
                seed.x = __float_as_uint(a4.w); seed.y = __float_as_uint(c4.w);
                if (kMulti) {
                    const float4 f4 = pb.final_sb[slot];
                    final_color = v3(f4.x, f4.y, f4.z);
                    s = __float_as_uint(f4.w) & 0xffffu; j = __float_as_uint(f4.w) >> 16;
                } else {
                    j = iter;        // S == 1: no restarts, so iteration k shades segment k of sample 0 and the sum over samples is still 0:
                }                    // the third state record is neither written nor read (16 B per path and iteration each way)
            }
            // S == 1, iteration 0 shades the primary hit: the ray is K-raygen's, restated here from the pixel (same function, same
            // bits — restarted samples rely on that already) instead of read back: 32 B per pixel less to fetch (C2 K-shade
            // 0.295 -> 0.281 ms together with the two-record path state; in the S > 1 kernel, where it touches one iteration of
            // hundreds, the extra code measured +1.5 % on C3 and is left out: profiles/r2y_variants.txt)
            V3 origin, direction;
            if (!kMulti && iter == 0u) {
                origin = v3(fp.cam.position[0], fp.cam.position[1], fp.cam.position[2]);
                direction = primary_direction(fp.cam, x, y, pm.width, pm.height);
            } else {
                const float4 rd = pb.ray_d[slot];
                direction = v3(rd.x, rd.y, rd.z);
                origin = v3s(0.0f);
                // the origin is only consumed by a segment that continues (:86): with S == 1 no path of the last iteration does;
                // primary rays (iteration 0) start at the camera, and K-raygen may not have written their origins (primary_lean)
                if (iter == 0u) origin = v3(fp.cam.position[0], fp.cam.position[1], fp.cam.position[2]);
                else if (kMulti || iter + 1u < B) { const float4 ro = pb.ray_o[slot]; origin = v3(ro.x, ro.y, ro.z); }
            }
            HitRec h;
            {
                const float4 ha = pb.hit_a[slot]; const uint2 hb = pb.hit_b[slot];
                h.t = ha.x; h.b1 = ha.y; h.b2 = ha.z; h.instance = __float_as_uint(ha.w);
                h.primitive_index = hb.x; h.triangle_index = hb.y;
            }
            // Every sample of a pixel shoots the same primary ray (quirk Q1), so the primary hit is traced once.
            // With S > 1 its *shading inputs* (64 B: normal, roughness, albedo, metallic, emission, hit distance,
            // geometric normal) are cached at iteration 0 and a restarted sample reads them back instead of
            // re-fetching vertices, material and texels: same values, hence the same bits, without the chain of
            // dependent loads (K-shade is latency-bound on multi-sample workloads).
            n_hit += h.t != kMissT ? 1u : 0u;
            const bool keep_primary = kMulti && iter == 0u && !(fp.flags & PT_FLAG_NO_PRIMARY_REUSE);
            if (iter == 0u) {
                const size_t pi = (size_t)y * pm.width + x;
                fp.id_inst[pi] = h.instance; fp.id_prim[pi] = h.primitive_index; fp.id_tri[pi] = h.triangle_index;
            }
            bool from_cache = false;
            for (;;) {
                bool end_sample = false;
                float t_hit;
                V3 sf_normal = v3s(0.0f), sf_albedo = v3s(0.0f), sf_emissive = v3s(0.0f), sf_gn = v3s(0.0f);
                float sf_roughness = 0.0f, sf_metallic = 0.0f;
                if (kMulti && from_cache) {
                    const float4 c2 = pb.prim_c[slot];
                    t_hit = c2.w; sf_emissive = v3(c2.x, c2.y, c2.z);
                    if (t_hit != kMissT && j + 1u < B) {
                        const float4 c0 = pb.prim_a[slot], c1 = pb.prim_b[slot], c3 = pb.prim_d[slot];
                        sf_normal = v3(c0.x, c0.y, c0.z); sf_roughness = c0.w;
                        sf_albedo = v3(c1.x, c1.y, c1.z); sf_metallic = c1.w;
                        sf_gn = v3(c3.x, c3.y, c3.z);
                    }
                } else {
                    t_hit = h.t;
                    if (t_hit != kMissT) {
                        if (j + 1u >= B && !keep_primary) {
                            // last segment: loop bound reached (:49,:77); only the emission of this hit is consumed (:75)
                            sf_emissive = get_emissive<kTex>(sc, h, s_unorm);
                        } else {
                            const TriangleData td = get_triangle_data(sc, h);                   // :72
                            const MaterialData md = get_material_data<kTex>(sc, td, s_unorm);         // :73
                            sf_normal = md.normal; sf_roughness = md.roughness;
                            sf_albedo = v3(md.albedo.x, md.albedo.y, md.albedo.z); sf_metallic = md.metallic;
                            sf_emissive = md.emissive; sf_gn = td.geometry_normal;
                        }
                    }
                    if (keep_primary) {
                        pb.prim_c[slot] = make_float4(sf_emissive.x, sf_emissive.y, sf_emissive.z, t_hit);
                        if (t_hit != kMissT) {
                            pb.prim_a[slot] = make_float4(sf_normal.x, sf_normal.y, sf_normal.z, sf_roughness);
                            pb.prim_b[slot] = make_float4(sf_albedo.x, sf_albedo.y, sf_albedo.z, sf_metallic);
                            pb.prim_d[slot] = make_float4(sf_gn.x, sf_gn.y, sf_gn.z, 0.0f);
                        }
                    }
                }
                if (t_hit == kMissT) {                                                      // :65-70
                    const float t = 0.5f * (direction.y + 1.0f);
                    color = color + atten * mix(v3(1.0f, 1.0f, 1.0f), v3(0.5f, 0.7f, 1.0f), t);
                    end_sample = true;
                } else {
                    color = color + sf_emissive * 10.0f * atten;                            // :75
                    if (j + 1u < B) {                                                       // :77
                        ShadingInfo si;
                        si.out_dir = -direction;                                            // :79
                        si.normal = sf_normal;                                              // :80
                        si.lambert_out = dot(si.normal, si.out_dir);                        // :81
                        si.fresnel_0 = mix(v3s(0.02f), sf_albedo, sf_metallic);             // :82
                        si.diffuse_albedo = sf_albedo - sf_metallic * sf_albedo;            // :83
                        si.roughness = gmax(0.006f, sf_roughness);                          // :84
                        origin = origin + t_hit * direction + sf_gn * 0.00001f;             // :86
                        direction = sample_brdf(si, pcg2d(seed));                           // :87
                        const float density = get_brdf_density(si, direction);              // :89
                        const float lambert_in = dot(si.normal, direction);                 // :90
                        if (lambert_in <= 0.0f) {                                           // :91-93
                            end_sample = true;
                        } else {
                            atten = atten * (brdf(si, direction) * lambert_in / density);   // :95
                            ++j;
                            emit = true;
                            // PT_FLAG_RUSSIAN_ROULETTE (extension, off by default; include/b200pt.h): same rule as the oracle's
                            if ((fp.flags & PT_FLAG_RUSSIAN_ROULETTE) && j >= 2u) {
                                const float p = gmin(1.0f, gmax(0.05f, gmax(atten.x, gmax(atten.y, atten.z))));
                                if (p < 1.0f) {
                                    const V2 xr = pcg2d(seed);
                                    if (!(xr.x < p)) { emit = false; end_sample = true; }
                                    else atten = atten / p;
                                }
                            }
                            // PT_FLAG_SKIP_DEAD_SEGMENTS: the last segment draws no random numbers and only adds
                            // atten * (emission | sky), both finite, so with a throughput of exactly zero (brdf() == 0: shading normal
                            // facing away from the viewer) it adds exactly nothing and is not traced (C2: 5 % of all rays)
                            if ((fp.flags & PT_FLAG_SKIP_DEAD_SEGMENTS) && emit && j + 1u >= B && atten.x == 0.0f && atten.y == 0.0f && atten.z == 0.0f) { emit = false; end_sample = true; }
                        }
                    } else {
                        end_sample = true;   // last segment: loop bound reached (:49)
                    }
                }
                if (!end_sample) break;      // next segment must be traced
                final_color = final_color + color;                                          // :99
                ++s;
                if (!kMulti || s == S) { write_pixel(fp, pm.width, x, y, final_color); break; }   // :102-104 (kMulti = false: S == 1)
                // next sample (:44-48)
                color = v3s(0.0f); atten = v3s(1.0f); j = 0u;
                origin = v3(fp.cam.position[0], fp.cam.position[1], fp.cam.position[2]);
                direction = primary_direction(fp.cam, x, y, pm.width, pm.height);
                if (fp.flags & PT_FLAG_NO_PRIMARY_REUSE) { emit = true; break; }
                from_cache = true;
            }
            if (emit) {
                store_ray(pb.ray_o, pb.ray_d, pb.ray_e, pb.ray_s, pb.ray_m, slot, origin, direction, 0.001f, 10000.0f);
                pb.atten_sx[slot] = make_float4(atten.x, atten.y, atten.z, __uint_as_float(seed.x));
                pb.color_sy[slot] = make_float4(color.x, color.y, color.z, __uint_as_float(seed.y));
                if (kMulti) pb.final_sb[slot] = make_float4(final_color.x, final_color.y, final_color.z, __uint_as_float(s | (j << 16)));
            }
        }
        queue_push(qout, qout_count, emit, slot);
    }
    n_hit = __reduce_add_sync(0xffffffffu, n_hit);
    if ((threadIdx.x & 31u) == 0u && n_hit) atomicAdd(pb.hit_count, (unsigned long long)n_hit);
}

__global__ void k_interleave_textures(const uint32_t* __restrict__ t0, const uint32_t* __restrict__ t1, const uint32_t* __restrict__ t2,
                                      const uint32_t* __restrict__ t3, size_t n, uint4* __restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = make_uint4(t0 ? t0[i] : 0xFFFFFFFFu, t1 ? t1[i] : 0xFFFFFFFFu, t2 ? t2[i] : 0xFFFFFFFFu, t3 ? t3[i] : 0xFFFFFFFFu);
}

// ------------------------------------------------------------------ post-processing
// fp32 RGBA -> b8g8r8a8_unorm as the reference's storage image does (Swapchain.zig:115-118):
// NaN -> 0, clamp to [0,1], round to nearest.
__global__ void k_to_bgra8(const float4* __restrict__ in, uchar4* __restrict__ out, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 v = in[i];
    auto q = [](float f) -> unsigned char {
        if (!(f == f)) f = 0.0f;
        f = fminf(fmaxf(f, 0.0f), 1.0f);
        return (unsigned char)(f * 255.0f + 0.5f);
    };
    out[i] = make_uchar4(q(v.z), q(v.y), q(v.x), q(v.w));
}

// nearest up-scale (the blit at RaytracingPass.zig:599-635)
__global__ void k_upscale(const float4* __restrict__ src, uint32_t sw, uint32_t sh, float4* __restrict__ dst, uint32_t dw, uint32_t dh) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= dw * dh) return;
    const uint32_t x = i % dw, y = i / dw;
    // vkCmdBlitImage with VK_FILTER_NEAREST samples the source at the destination texel's centre: floor((x + 0.5) * sw / dw)
    const uint32_t sx = min(sw - 1u, (uint32_t)(((2ull * x + 1ull) * sw) / (2ull * dw))), sy = min(sh - 1u, (uint32_t)(((2ull * y + 1ull) * sh) / (2ull * dh)));
    dst[i] = src[(size_t)sy * sw + sx];
}

__global__ void k_accumulate(const float4* __restrict__ frame, float4* __restrict__ accum, uint32_t n4) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 a = accum[i], f = frame[i];
    accum[i] = make_float4(a.x + f.x, a.y + f.y, a.z + f.z, a.w + f.w);
}

__global__ void k_scale(const float4* __restrict__ in, float4* __restrict__ out, uint32_t n4, float s) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 a = in[i];
    out[i] = make_float4(a.x * s, a.y * s, a.z * s, a.w * s);
}

__global__ void k_debug_bsdf(uint32_t n, const float* N, const float* V, const float* kd, const float* f0, const float* rough,
                             const float* xi, float* L, float* pdf, float* f) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    ShadingInfo s;
    s.normal = v3(N[3 * i], N[3 * i + 1], N[3 * i + 2]);
    s.out_dir = v3(V[3 * i], V[3 * i + 1], V[3 * i + 2]);
    s.lambert_out = dot(s.normal, s.out_dir);
    s.diffuse_albedo = v3(kd[3 * i], kd[3 * i + 1], kd[3 * i + 2]);
    s.fresnel_0 = v3(f0[3 * i], f0[3 * i + 1], f0[3 * i + 2]);
    s.roughness = rough[i];
    V2 r; r.x = xi[2 * i]; r.y = xi[2 * i + 1];
    const V3 l = sample_brdf(s, r);
    const float p = get_brdf_density(s, l);
    const V3 v = brdf(s, l);
    L[3 * i] = l.x; L[3 * i + 1] = l.y; L[3 * i + 2] = l.z;
    pdf[i] = p;
    f[3 * i] = v.x; f[3 * i + 1] = v.y; f[3 * i + 2] = v.z;
}

__global__ void k_debug_rng(uint32_t n, const uint32_t* pxf, uint32_t draws, uint32_t* state, float* xi) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    U2 s = prng_seed(pxf[3 * i], pxf[3 * i + 1], pxf[3 * i + 2]);
    V2 r; r.x = 0.0f; r.y = 0.0f;
    for (uint32_t k = 0; k < draws; ++k) r = pcg2d(s);
    state[2 * i] = s.x; state[2 * i + 1] = s.y;
    xi[2 * i] = r.x; xi[2 * i + 1] = r.y;
}

__global__ void k_debug_texture(TexRec tex, uint32_t n, const float* uv, float* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V2 c; c.x = uv[2 * i]; c.y = uv[2 * i + 1];
    const V4 t = sample_texture(tex, c);
    out[4 * i] = t.x; out[4 * i + 1] = t.y; out[4 * i + 2] = t.z; out[4 * i + 3] = t.w;
}

template <typename K>
int persistent_blocks(K kernel, int threads, int num_sms) {
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0);
    if (per_sm < 1) per_sm = 1;
    return per_sm * num_sms;
}

}  // namespace

// tuning knob (pt_set_option): default fetch granularity of renderers created afterwards
static uint32_t g_default_chunk = 32;
int set_option(const char* key, int value) {
    if (!strcmp(key, "chunk")) { g_default_chunk = (uint32_t)(value < 1 ? 1 : value); return 0; }   // K-trace: rays per queue atomic
    return -1;
}

LaunchDims compute_launch_dims(int num_sms) {
    LaunchDims d;
    d.trace_walker = persistent_blocks(k_trace<TRACE_BVH8>, kTraceThreads, num_sms);
    d.trace_bvh2 = persistent_blocks(k_trace<TRACE_BVH2>, kTraceThreads, num_sms);
    d.trace_count = persistent_blocks(k_trace<TRACE_BVH8_COUNT>, kTraceThreads, num_sms);
    d.trace_persistent = persistent_blocks(k_trace_persistent<false>, kTraceBlock, num_sms);
    d.trace_flat = persistent_blocks(k_trace_persistent<true>, kTraceBlock, num_sms);
    // the launch bounds fix the residency (6 / 4 CTAs per SM) for every fetch mode
    d.shade_single = persistent_blocks(k_shade<false, 0>, kShadeThreads, num_sms);
    d.shade_multi = persistent_blocks(k_shade<true, 0>, kShadeThreads, num_sms);
    d.chunk = g_default_chunk;
    return d;
}

void launch_raygen(const PixelMap& pm, const PathBuffers& pb, const FrameParams& fp, cudaStream_t st) {
    if (pm.num_slots == 0) return;
    k_raygen<<<(pm.num_slots + 255) / 256, 256, 0, st>>>(pm, pb, fp);
}

void launch_trace(const DeviceScene& sc, const PathBuffers& pb, uint32_t iter, TraceMode mode, const float* uniform_origin, uint32_t tlas2_count,
                  const uint32_t* tlas2_leaf_instance, const LaunchDims& dims, cudaStream_t st) {
    if (mode == TRACE_BVH8) {
        TraceIO io;
        io.queue = pb.queue[iter & 1u]; io.count_ptr = pb.qcount + iter; io.count = 0; io.cursor = pb.trace_next + iter;
        io.ray_o = pb.ray_o; io.ray_d = pb.ray_d; io.ray_e = pb.ray_e; io.ray_s = pb.ray_s; io.ray_m = pb.ray_m; io.hit_a = pb.hit_a; io.hit_b = pb.hit_b;
        io.hit_a_stride = 1; io.hit_b_stride = 1; io.bias_word = 0x4B000000u; io.chunk = dims.chunk;
        io.uniform_origin = uniform_origin ? 1u : 0u;
        io.origin = uniform_origin ? make_float4(uniform_origin[0], uniform_origin[1], uniform_origin[2], 0.001f) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (sc.start_inst != kInvalidId) k_trace_persistent<true><<<dims.trace_flat, kTraceBlock, 0, st>>>(sc, io);
        else k_trace_persistent<false><<<dims.trace_persistent, kTraceBlock, 0, st>>>(sc, io);
        return;
    }
    if (mode == TRACE_BVH2) k_trace<TRACE_BVH2><<<dims.trace_bvh2, kTraceThreads, 0, st>>>(sc, pb, iter, tlas2_count, tlas2_leaf_instance);
    else if (mode == TRACE_BVH8_COUNT) k_trace<TRACE_BVH8_COUNT><<<dims.trace_count, kTraceThreads, 0, st>>>(sc, pb, iter, tlas2_count, tlas2_leaf_instance);
    else k_trace<TRACE_BVH8><<<dims.trace_walker, kTraceThreads, 0, st>>>(sc, pb, iter, tlas2_count, tlas2_leaf_instance);
}

void launch_material_sort(const DeviceScene& sc, const PathBuffers& pb, uint32_t iter, int num_sms, cudaStream_t st) {
    const uint32_t bins = sc.num_materials + 1u;
    cudaMemsetAsync(pb.sort_bins, 0, sizeof(uint32_t) * (bins + 1u), st);
    k_matsort_count<<<num_sms * 4, 256, 0, st>>>(sc, pb, iter);
    k_matsort_scan<<<1, 1024, 0, st>>>(pb, bins);
    k_matsort_scatter<<<num_sms * 4, 256, 0, st>>>(sc, pb, iter);
}

void launch_shade(const DeviceScene& sc, const PixelMap& pm, const PathBuffers& pb, const FrameParams& fp, uint32_t iter,
                  const LaunchDims& dims, cudaStream_t st) {
    const int mode = sc.mat_tex_objs ? 1 : (sc.atlas_only ? 2 : 0);
    const int blocks = fp.num_samples > 1u ? dims.shade_multi : dims.shade_single;
    if (fp.num_samples > 1u) {
        if (mode == 1) k_shade<true, 1><<<blocks, kShadeThreads, 0, st>>>(sc, pm, pb, fp, iter);
        else if (mode == 2) k_shade<true, 2><<<blocks, kShadeThreads, 0, st>>>(sc, pm, pb, fp, iter);
        else k_shade<true, 0><<<blocks, kShadeThreads, 0, st>>>(sc, pm, pb, fp, iter);
    } else {
        if (mode == 1) k_shade<false, 1><<<blocks, kShadeThreads, 0, st>>>(sc, pm, pb, fp, iter);
        else if (mode == 2) k_shade<false, 2><<<blocks, kShadeThreads, 0, st>>>(sc, pm, pb, fp, iter);
        else k_shade<false, 0><<<blocks, kShadeThreads, 0, st>>>(sc, pm, pb, fp, iter);
    }
}

void launch_trace_rays(const DeviceScene& sc, const RayRec* rays, uint32_t n, HitRec* hits, TraceMode mode,
                       uint32_t tlas2_count, const uint32_t* tlas2_leaf_instance, uint32_t* cursor, float4* setup, int num_sms,
                       const LaunchDims& dims, cudaStream_t st) {
    const int blocks = num_sms * 4;
    if (mode == TRACE_BVH8) {
        static_assert(sizeof(RayRec) == 32 && sizeof(HitRec) == 32, "AoS strides");
        float4 *so = setup, *sd = setup + n, *se = setup + 2 * (size_t)n, *ss = setup + 3 * (size_t)n, *sm = setup + 4 * (size_t)n;
        k_ray_setup<<<(n + 255) / 256, 256, 0, st>>>(rays, n, so, sd, se, ss, sm);
        TraceIO io;
        io.queue = nullptr; io.count_ptr = nullptr; io.count = n; io.cursor = cursor;
        io.ray_o = so; io.ray_d = sd; io.ray_e = se; io.ray_s = ss; io.ray_m = sm;
        io.hit_a = reinterpret_cast<float4*>(hits); io.hit_a_stride = 2;
        io.hit_b = reinterpret_cast<uint2*>(hits) + 2; io.hit_b_stride = 4;
        io.bias_word = 0x4B000000u; io.chunk = dims.chunk; io.uniform_origin = 0u; io.origin = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (sc.start_inst != kInvalidId) k_trace_persistent<true><<<blocks, kTraceBlock, 0, st>>>(sc, io);
        else k_trace_persistent<false><<<blocks, kTraceBlock, 0, st>>>(sc, io);
        return;
    }
    if (mode == TRACE_BVH2) k_trace_rays<TRACE_BVH2><<<blocks, kTraceThreads, 0, st>>>(sc, rays, n, hits, tlas2_count, tlas2_leaf_instance, cursor);
    else k_trace_rays<TRACE_BVH8><<<blocks, kTraceThreads, 0, st>>>(sc, rays, n, hits, tlas2_count, tlas2_leaf_instance, cursor);
}

void launch_debug_bsdf(uint32_t n, const float* N, const float* V, const float* kd, const float* f0, const float* rough,
                       const float* xi, float* L, float* pdf, float* f, cudaStream_t st) {
    if (n) k_debug_bsdf<<<(n + 127) / 128, 128, 0, st>>>(n, N, V, kd, f0, rough, xi, L, pdf, f);
}
void launch_debug_rng(uint32_t n, const uint32_t* pxf, uint32_t draws, uint32_t* state, float* xi, cudaStream_t st) {
    if (n) k_debug_rng<<<(n + 127) / 128, 128, 0, st>>>(n, pxf, draws, state, xi);
}
void launch_debug_texture(TexRec tex, uint32_t n, const float* uv, float* out, cudaStream_t st) {
    if (n) k_debug_texture<<<(n + 127) / 128, 128, 0, st>>>(tex, n, uv, out);
}

void launch_interleave_textures(const uint32_t* t0, const uint32_t* t1, const uint32_t* t2, const uint32_t* t3, size_t n, uint4* out,
                                cudaStream_t st) {
    if (n) k_interleave_textures<<<1184, 256, 0, st>>>(t0, t1, t2, t3, n, out);
}
void launch_to_bgra8(const float* rgba, uint8_t* bgra, uint32_t n_pixels, cudaStream_t st) {
    if (n_pixels) k_to_bgra8<<<(n_pixels + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(rgba), reinterpret_cast<uchar4*>(bgra), n_pixels);
}
void launch_upscale(const float* src, uint32_t sw, uint32_t sh, float* dst, uint32_t dw, uint32_t dh, cudaStream_t st) {
    const uint32_t n = dw * dh;
    if (n) k_upscale<<<(n + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(src), sw, sh, reinterpret_cast<float4*>(dst), dw, dh);
}
void launch_accumulate(const float* frame, float* accum, uint32_t n_floats, cudaStream_t st) {
    const uint32_t n4 = n_floats / 4;
    if (n4) k_accumulate<<<(n4 + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(frame), reinterpret_cast<float4*>(accum), n4);
}
void launch_scale(const float* accum, float* out, uint32_t n_floats, float scale, cudaStream_t st) {
    const uint32_t n4 = n_floats / 4;
    if (n4) k_scale<<<(n4 + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(accum), reinterpret_cast<float4*>(out), n4, scale);
}

}  // namespace pt
