// noalloc.cu — pt_trace_rays_noalloc: the standalone closest-hit query of pt_trace_rays for callers that cannot
// allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The query needs scratch memory: the SoA ray records K-trace fetches (5 x float4 per ray), the 32-byte hit records it
// stores, and the dynamic-fetch cursor.  pt_trace_rays cudaMallocs them per call; here they live in ONE file-scope
// __device__ array of a fixed size (PT_NOALLOC_MAX_RAYS rays), which every GPU that loads the library holds until
// the process exits.  Consequences, also stated in include/b200pt.h:
//   * n > PT_NOALLOC_MAX_RAYS is rejected (PT_ERR_INVALID_ARGUMENT), nothing is launched;
//   * every call on a device shares the array: calls that overlap on the GPU (different streams, different renderers,
//     different host threads) race.  Calls on one stream are ordered and safe;
//   * the array keeps the last call's contents: the cursor is re-zeroed by the set-up kernel of every call, ray and hit
//     records of slots [0, n) are rewritten before they are read, slots >= n are never read.
// The host never uses the array's NAME as a pointer (in host code it is a stand-in symbol, not a device address):
// scratch_base() asks the runtime for the device address, once per device, with that device current.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "renderer.h"
#include "trace.cuh"
// This file's instantiation of the persistent traversal kernel gets a name of its own, so that render.cu's launches keep
// resolving to render.cu's copy (two same-named template kernels in one library would share one host stub).
#define k_trace_persistent k_trace_persistent_noalloc
#include "trace_kernel.cuh"

using namespace pt;

namespace {

constexpr size_t kMaxRays = PT_NOALLOC_MAX_RAYS;
constexpr size_t kCursorBytes = 256;                               // keeps the records 256-byte aligned
constexpr size_t kRayBytes = 5 * sizeof(float4) * kMaxRays;        // ray_o | ray_d | ray_e | ray_s | ray_m
constexpr size_t kHitBytes = sizeof(HitRec) * kMaxRays;
constexpr size_t kScratchBytes = kCursorBytes + kRayBytes + kHitBytes;   // 256 + 112 * 131072 = 14 680 320 B (14 MiB)

static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_base[kMaxDevices] = {};   // device address of g_noalloc_scratch on each device (nullptr = not asked yet)

// Device address of the scratch array on `device`, which must be current.  cudaGetSymbolAddress allocates nothing: the
// array is part of the module image.
int scratch_base(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxDevices) { set_error("pt_trace_rays_noalloc: device ordinal out of range"); return PT_ERR_INVALID_ARGUMENT; }
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_base[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            set_error(std::string("pt_trace_rays_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
            return PT_ERR_CUDA;
        }
        g_base[device] = static_cast<unsigned char*>(p);
    }
    *out = g_base[device];
    return PT_OK;
}

struct Scratch {
    uint32_t* cursor;
    float4 *ray_o, *ray_d, *ray_e, *ray_s, *ray_m;
    HitRec* hits;
};

Scratch carve(unsigned char* base) {
    Scratch s;
    s.cursor = reinterpret_cast<uint32_t*>(base);
    float4* rays = reinterpret_cast<float4*>(base + kCursorBytes);
    s.ray_o = rays; s.ray_d = rays + kMaxRays; s.ray_e = rays + 2 * kMaxRays; s.ray_s = rays + 3 * kMaxRays; s.ray_m = rays + 4 * kMaxRays;
    s.hits = reinterpret_cast<HitRec*>(base + kCursorBytes + kRayBytes);
    return s;
}

// caller's PtRay array -> the SoA ray records; also re-zeroes the cursor the last call left at >= its n
__global__ void __launch_bounds__(256) k_noalloc_setup(const RayRec* __restrict__ rays, uint32_t n, Scratch s) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) *s.cursor = 0u;
    if (i >= n) return;
    const RayRec r = rays[i];
    store_ray(s.ray_o, s.ray_d, s.ray_e, s.ray_s, s.ray_m, i, v3(r.ox, r.oy, r.oz), v3(r.dx, r.dy, r.dz), r.tmin, r.tmax);
}

// 32-byte hit records -> the caller's 24-byte PtHit array
__global__ void __launch_bounds__(256) k_noalloc_emit(const HitRec* __restrict__ recs, uint32_t n, PtHit* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const HitRec h = recs[i];
    PtHit o;
    o.t = h.t; o.b1 = h.b1; o.b2 = h.b2; o.instance = h.instance; o.primitive_index = h.primitive_index; o.triangle_index = h.triangle_index;
    out[i] = o;
}

}  // namespace

#define PT_REQUIRE(cond, msg) do { if (!(cond)) { set_error(msg); return PT_ERR_INVALID_ARGUMENT; } } while (0)

extern "C" int pt_trace_rays_noalloc(PtRenderer* r, const PtRay* device_rays, size_t n, PtHit* device_hits, void* cuda_stream) {
    PT_REQUIRE(r && (n == 0 || (device_rays && device_hits)), "pt_trace_rays_noalloc: null argument");
    PT_REQUIRE(n <= kMaxRays, "pt_trace_rays_noalloc: n exceeds PT_NOALLOC_MAX_RAYS (" + std::to_string(kMaxRays) +
                                  " rays): the scratch array is fixed-size; split the batch or use pt_trace_rays");
    PT_REQUIRE(r->trace_mode == TRACE_BVH8, "pt_trace_rays_noalloc: only the default traversal kernel (no PT_FLAG_TRACE_* flag) has a no-allocation variant");
    if (n == 0) return PT_OK;
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    unsigned char* base = nullptr;
    { const int rc = scratch_base(r->device, &base); if (rc != PT_OK) return rc; }
    const Scratch s = carve(base);
    const DeviceScene& sc = r->parent ? r->parent->scene : r->scene;
    cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : r->stream;
    const uint32_t n32 = (uint32_t)n;
    static_assert(sizeof(PtRay) == sizeof(RayRec), "PtRay layout");
    k_noalloc_setup<<<(n32 + 255u) / 256u, 256, 0, st>>>(reinterpret_cast<const RayRec*>(device_rays), n32, s);
    TraceIO io;
    io.queue = nullptr; io.count_ptr = nullptr; io.count = n32; io.cursor = s.cursor;
    io.ray_o = s.ray_o; io.ray_d = s.ray_d; io.ray_e = s.ray_e; io.ray_s = s.ray_s; io.ray_m = s.ray_m;
    io.hit_a = reinterpret_cast<float4*>(s.hits); io.hit_a_stride = 2;
    io.hit_b = reinterpret_cast<uint2*>(s.hits) + 2; io.hit_b_stride = 4;
    io.bias_word = 0x4B000000u; io.chunk = r->dims.chunk; io.uniform_origin = 0u; io.origin = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const int blocks = r->num_sms * 4;   // as launch_trace_rays (render.cu)
    if (sc.start_inst != kInvalidId) k_trace_persistent<true><<<blocks, kTraceBlock, 0, st>>>(sc, io);
    else k_trace_persistent<false><<<blocks, kTraceBlock, 0, st>>>(sc, io);
    k_noalloc_emit<<<(n32 + 255u) / 256u, 256, 0, st>>>(s.hits, n32, device_hits);
    PT_CUDA_CHECK(cudaGetLastError());
    return PT_OK;   // enqueued, not waited for: pt_sync(r) covers the renderer's own stream, a caller's stream is the caller's to wait on
}
