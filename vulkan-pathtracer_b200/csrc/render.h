// render.h — wavefront frame state and kernel launchers (render.cu).
#pragma once
#include "pt_internal.h"

namespace pt {

// Tile-ordered pixel slots: tile = 32x8 pixels, 8 warps per tile, each warp an 8x4 block (coherent
// rays per warp).  Slot i of this rank -> owned tile (rank + (i/256)*world).
struct PixelMap {
    uint32_t width, height;       // render extent
    uint32_t tiles_x, tiles_y;
    uint32_t rank, world;         // world == 1: every tile is owned
    uint32_t num_slots;           // owned tiles * 256
};

// Per-slot path state in HBM (SoA of 16-byte records).
struct PathBuffers {
    float4* ray_o;       // origin.xyz, tmin
    float4* ray_d;       // direction.xyz, tmax
    float4* ray_e;       // 1/direction, octant + shear axes (bits)     } derived by store_ray() where the ray is
    float4* ray_s;       // shear row m0.xyz, Sz                        } produced; K-trace only loads them
    float4* ray_m;       // shear row m1.xyz
    float4* hit_a;       // t, b1, b2, instance (bits)
    uint2* hit_b;        // primitive_index, triangle_index
    // cached shading inputs of the primary hit (S > 1; reused for every sample, quirk Q1):
    // a = normal, roughness | b = albedo, metallic | c = emission, hit distance (1e32 = miss) | d = geometric normal
    float4 *prim_a, *prim_b, *prim_c, *prim_d;
    float4* atten_sx;    // atten.rgb, seed.x (bits)
    float4* color_sy;    // color.rgb, seed.y (bits)
    float4* final_sb;    // final_color.rgb, sample | bounce << 16 (bits)
    uint32_t* queue[2];  // slot indices of live paths, double buffered
    uint32_t* queue_sorted;   // PT_FLAG_SORT_BY_MATERIAL: the current queue reordered by material for K-shade
    uint32_t* sort_bins;      // [num_materials + 2] counting-sort bins (zero between uses)
    uint32_t* qcount;    // [max_iters + 1] entries of queue k (zeroed per frame)
    uint32_t* trace_next;// [max_iters] dynamic-fetch cursors of the persistent trace kernel
    uint32_t* shade_next;// [max_iters] the same for K-shade (qcount, trace_next and shade_next are one allocation)
    uint32_t* trav_counters;  // [2] nodes, tris (counting walker)
    unsigned long long* hit_count;   // rays of this frame that hit geometry (PtStats::rays_hit); lives behind shade_next, zeroed per frame
    uint32_t frame_words;     // size of the qcount allocation in 32-bit words (counts, cursors, hit counter): one memset per frame
};

struct FrameParams {
    PtCamera cam;
    uint32_t num_samples, num_bounces;
    uint32_t flags;
    // which records of a primary ray nobody reads (bit 0: origin + tmin, bit 1: direction): the flat instantiation of the
    // persistent K-trace takes the camera position as a launch constant at iteration 0 and never reads directions, and K-shade
    // restates the primary ray from the pixel when S == 1.  Set by pt_render; 0 for two-level scenes and the walker / counting modes.
    uint32_t primary_lean;
    float* image;        // W*H*4 fp32 (full image; non-owned pixels untouched)
    uint32_t* id_inst;   // W*H each
    uint32_t* id_prim;
    uint32_t* id_tri;
};

// TRACE_BVH8: persistent dynamic-fetch kernel (trace_kernel.cuh); TRACE_BVH8_WALKER: the simple
// one-ray-per-thread walker of trace.cuh (TRACE_BVH8_COUNT: same, counting nodes/triangles);
// TRACE_BVH2: walker over the LBVH binary trees.
enum TraceMode { TRACE_BVH8 = 0, TRACE_BVH2 = 1, TRACE_BVH8_COUNT = 2, TRACE_BVH8_WALKER = 3 };

// Persistent grid sizes (resident CTAs per SM x SM count of the renderer's device) and the K-trace fetch granularity;
// computed once per renderer in pt_create, after cudaSetDevice (renderers on different devices do not share them).
struct LaunchDims {
    int trace_persistent = 0, trace_flat = 0, trace_walker = 0, trace_bvh2 = 0, trace_count = 0;
    int shade_single = 0, shade_multi = 0;
    uint32_t chunk = 32;          // rays a K-trace warp takes from the queue per global atomic
};
LaunchDims compute_launch_dims(int num_sms);

void launch_raygen(const PixelMap& pm, const PathBuffers& pb, const FrameParams& fp, cudaStream_t st);
// uniform_origin: non-null at iteration 0 when FrameParams::primary_lean bit 0 is set (camera position; tmin is K-raygen's 0.001)
void launch_trace(const DeviceScene& sc, const PathBuffers& pb, uint32_t iter, TraceMode mode, const float* uniform_origin, uint32_t tlas2_count,
                  const uint32_t* tlas2_leaf_instance, const LaunchDims& dims, cudaStream_t st);
// PT_FLAG_SORT_BY_MATERIAL: counting sort of queue `iter` by the material of each path's hit (misses last)
void launch_material_sort(const DeviceScene& sc, const PathBuffers& pb, uint32_t iter, int num_sms, cudaStream_t st);
void launch_shade(const DeviceScene& sc, const PixelMap& pm, const PathBuffers& pb, const FrameParams& fp, uint32_t iter,
                  const LaunchDims& dims, cudaStream_t st);
// standalone closest-hit queries (pt_trace_rays)
void launch_trace_rays(const DeviceScene& sc, const RayRec* rays, uint32_t n, HitRec* hits, TraceMode mode,
                       uint32_t tlas2_count, const uint32_t* tlas2_leaf_instance, uint32_t* cursor, float4* setup /* 5*n */,
                       int num_sms, const LaunchDims& dims, cudaStream_t st);
// material texture interleave (MatRec::atlas): out[i] = (t0[i], t1[i], t2[i], t3[i]); absent maps read as 0xFFFFFFFF
void launch_interleave_textures(const uint32_t* t0, const uint32_t* t1, const uint32_t* t2, const uint32_t* t3, size_t n, uint4* out,
                                cudaStream_t st);
// image post-processing
void launch_to_bgra8(const float* rgba, uint8_t* bgra, uint32_t n_pixels, cudaStream_t st);
void launch_upscale(const float* src, uint32_t sw, uint32_t sh, float* dst, uint32_t dw, uint32_t dh, cudaStream_t st);
void launch_accumulate(const float* frame, float* accum, uint32_t n_floats, cudaStream_t st);
void launch_scale(const float* accum, float* out, uint32_t n_floats, float scale, cudaStream_t st);

// kernel tuning knobs (pt_set_option)
int set_option(const char* key, int value);

// shading restatement on raw records (pt_debug_*)
void launch_debug_bsdf(uint32_t n, const float* N, const float* V, const float* kd, const float* f0, const float* rough,
                       const float* xi, float* L, float* pdf, float* f, cudaStream_t st);
void launch_debug_rng(uint32_t n, const uint32_t* pxf, uint32_t draws, uint32_t* state, float* xi, cudaStream_t st);
void launch_debug_texture(TexRec tex, uint32_t n, const float* uv, float* out, cudaStream_t st);

}  // namespace pt
