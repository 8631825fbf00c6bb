/* b200pt.h — C ABI of libb200pt.so, the B200-native (sm_100a CUDA) replacement for the
 * trace-loop hot path of antlilja/vulkan-pathtracer.
 *
 * The reference has no FFI; the seam is the public surface of `RaytracingPass`
 * (src/RaytracingPass.zig) as used by src/main.zig, with `Scene` (src/Scene.zig) and `Camera`
 * (src/Camera.zig) values crossing it.  Each entry point below cites the reference interface it
 * replaces.  Plain pointers and sizes only; no C++/torch types.  All functions return 0 on
 * success or a negative PtStatus; pt_last_error() gives a thread-local message.
 *
 * Ownership (mirrors `defer scene.deinit()` right after Scene.load, src/RaytracingPass.zig:102-103):
 * host scene arrays are borrowed for the duration of pt_create only; the renderer copies what it
 * needs to the device before returning and owns all device memory until pt_destroy.
 */
#ifndef B200PT_H
#define B200PT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ status codes
 * Negative values mirror the named Zig errors of the loader (src/Scene.zig:135,176-182,251,266,
 * 277,292,307,322,495-496) and generic failures. */
typedef enum PtStatus {
    PT_OK = 0,
    PT_ERR_INVALID_ARGUMENT = -1,
    PT_ERR_CUDA = -2,
    PT_ERR_OUT_OF_MEMORY = -3,
    PT_ERR_IO = -4,                       /* std.fs open/read errors, error.EndOfFile */
    PT_ERR_INVALID_FILE_EXTENSION = -10,  /* error.InvalidFileExtension   Scene.zig:135 */
    PT_ERR_GLTF_PARSE = -11,              /* zgltf parse errors            Scene.zig:92,109 */
    PT_ERR_GLTF_NOT_TRIANGLES = -12,      /* error.GltfNotATriangleTopology Scene.zig:176 */
    PT_ERR_GLTF_NO_INDICES = -13,         /* error.GltfNoIndices           Scene.zig:178 */
    PT_ERR_GLTF_NO_POSITIONS = -14,       /* error.GltfNoPositions         Scene.zig:179 */
    PT_ERR_GLTF_NO_NORMALS = -15,         /* error.GltfNoNormals           Scene.zig:180 */
    PT_ERR_GLTF_NO_TANGENTS = -16,        /* error.GltfNoTangents          Scene.zig:181 */
    PT_ERR_GLTF_NO_TEXCOORDS = -17,       /* error.GltfNoTextureCoordinates Scene.zig:182 */
    PT_ERR_NO_MATERIAL = -18,             /* error.NoMaterial              Scene.zig:251 */
    PT_ERR_NO_BUFFER_VIEW = -19,          /* error.NoBufferView            Scene.zig:266 */
    PT_ERR_INVALID_ATTRIBUTE_TYPE = -20,  /* error.Invalid{Position,Normal,Tangent,Texcoord}Type :277-322 */
    PT_ERR_NO_SCENE = -21,                /* error.NoScene                 Scene.zig:495 */
    PT_ERR_TOO_MANY_SCENES = -22,         /* error.TooManyScenes           Scene.zig:496 */
    PT_ERR_IMAGE_DECODE = -23,            /* stb_image failure             stb_image.zig:11-18 */
    PT_ERR_INVALID_INDEX_TYPE = -24,      /* `unreachable` at Scene.zig:252-256 (u8 indices) */
    PT_ERR_NCCL = -30,                    /* NCCL missing or a communicator error (multi-GPU combine) */
    PT_ERR_TRAVERSAL_STACK = -31          /* a ray needed more traversal stack than the kernels have (64 entries): the
                                             hierarchy is pathologically deep; reported by pt_sync / pt_get_stats instead
                                             of an image with missed hits */
} PtStatus;

/* ------------------------------------------------------------------ scene data contract
 * 1:1 with the host arrays of `Scene` (src/Scene.zig:14-54, 68-77). */

/* Scene.Primitive (src/Scene.zig:24-37).  Offsets are byte offsets into triangle_data.
 * info = material_index[0:24] | uint32_indices << 31  (device flags, src/RaytracingPass.zig:18-31,
 * src/shaders/triangle.glsl:4-5). */
typedef struct PtPrimitive {
    uint64_t indices_offset;
    uint64_t positions_offset;
    uint64_t normals_offset;
    uint64_t tangents_offset;
    uint64_t uvs_offset;
    uint32_t max_vertex;
    uint32_t triangle_count;
    uint32_t info;
    uint32_t _pad;
} PtPrimitive;

#define PT_MATERIAL_INDEX_MASK 0x00FFFFFFu
#define PT_UINT32_INDICES_MASK 0x80000000u
#define PT_INVALID_TEXTURE 0xFFFFFFFFu

/* Scene.Mesh (src/Scene.zig:19-22): primitive range [start, end) — one BLAS per mesh, one geometry
 * per primitive (src/RaytracingPass.zig:277-352). */
typedef struct PtMesh {
    uint32_t start;
    uint32_t end;
} PtMesh;

/* Scene.Material (src/Scene.zig:39-48) == shader Material (src/shaders/material.glsl:7-16). */
typedef struct PtMaterial {
    uint32_t albedo_factor;           /* r | g<<8 | b<<16, unorm8 */
    uint32_t metal_roughness_factor;  /* roughness<<8 | metallic<<16 */
    uint32_t emissive_factor;
    uint32_t albedo_texture_index;    /* PT_INVALID_TEXTURE = none */
    uint32_t metal_roughness_texture_index;
    uint32_t emissive_texture_index;
    uint32_t normal_texture_index;
} PtMaterial;

/* Scene.Texture (src/Scene.zig:50-54): tightly packed RGBA8, row 0 first, uploaded as
 * r8g8b8a8_unorm with a single mip (src/RaytracingPass.zig:500-512, src/Image.zig:32). */
typedef struct PtTexture {
    const uint8_t* rgba;
    uint32_t width;
    uint32_t height;
} PtTexture;

/* Scene.Instance (src/Scene.zig:14-17).  transform = 16 floats, rows of M (the Mat4 is stored
 * transposed, Scene.zig:529); the first 12 (3x4 row-major) are what the TLAS consumes
 * (src/Tlas.zig:33-41). */
typedef struct PtInstance {
    uint64_t mesh_index;
    float transform[16];
} PtInstance;

typedef struct PtSceneDesc {
    const uint8_t* triangle_data;
    size_t triangle_data_len;
    const PtPrimitive* primitives;
    size_t num_primitives;
    const PtMesh* meshes;
    size_t num_meshes;
    const PtMaterial* materials;
    size_t num_materials;
    const PtTexture* textures;
    size_t num_textures;
    const PtInstance* instances;
    size_t num_instances;
} PtSceneDesc;

/* ------------------------------------------------------------------ render configuration
 * Arguments of RaytracingPass.init (src/RaytracingPass.zig:87-97) and the CLI defaults
 * (src/main.zig:29-67). */
enum {
    PT_PARTITION_NONE = 0,   /* this renderer renders every pixel */
    PT_PARTITION_TILES = 1   /* interleaved 32x8 tiles: owner = tile_index % world */
};

enum {
    PT_FLAG_SORT_BY_MATERIAL = 1u << 0,   /* material-sorted shading queue */
    PT_FLAG_TRACE_BVH2 = 1u << 1,         /* use the simple BVH2 reference walker, not the BVH8 kernel */
    PT_FLAG_NO_PRIMARY_REUSE = 1u << 2,   /* re-trace the (identical) primary ray for every sample */
    PT_FLAG_TRACE_WALKER = 1u << 3,       /* simple one-ray-per-thread BVH8 walker instead of the persistent kernel */
    PT_FLAG_NO_TEXTURE_INTERLEAVE = 1u << 5,   /* keep every texture separate (A/B switch): by default the maps of a
                                             material that share one extent are interleaved per texel on the device */
    PT_FLAG_BVH_PLOC = 1u << 4,           /* build the binary hierarchy by locally-ordered clustering (PLOC) instead of
                                             the Morton radix tree (LBVH): slower build, better trees for irregular meshes */
    PT_FLAG_BVH_SAH = 1u << 6,            /* (default when no PT_FLAG_BVH_* is given: SAH | OPTIMAL_COLLAPSE)
                                             build the binary hierarchy top-down by binned SAH on the GPU (32 bins x 3 axes
                                             per segment, level-synchronous; csrc/bvh_sah.cuh).  Takes precedence over PLOC */
    PT_FLAG_BVH_OPTIMAL_COLLAPSE = 1u << 7, /* BVH2 -> BVH8 by the cost-optimal collapse (dynamic programme of Ylitie et al.
                                             2017) instead of the greedy surface-area expansion; LBVH and SAH hierarchies */
    PT_FLAG_BVH_LBVH = 1u << 9,           /* fast build: Morton radix tree (Karras 2012) + greedy collapse, 3-4x quicker to
                                             build and ~10 % slower to trace than the default (binned SAH + cost-optimal
                                             collapse, the reference's prefer_fast_trace: RaytracingPass.zig:353-358); with
                                             PT_FLAG_BVH_OPTIMAL_COLLAPSE: radix tree + optimal collapse */
    PT_FLAG_SKIP_DEAD_SEGMENTS = 1u << 10, /* do not trace the last segment of a sample whose throughput is exactly (0, 0, 0):
                                             it draws no random numbers and only adds atten * (emission | sky), i.e. nothing.
                                             Same image bit for bit, fewer executed queries than the reference's raygen
                                             issues (C2: -5 % rays); off by default so that ray counts stay the reference's */
    PT_FLAG_HW_BILINEAR = 1u << 11,       /* A/B switch (north_star (c) "CUDA texture objects"): material maps are
                                             sampled through cudaTextureObject_t (RGBA8 unorm, linear, wrap, LOD 0 — the
                                             sampler of RayTracingPipeline.zig:54-70 on the texture units) instead of four
                                             point fetches + fp32 lerp.  The hardware filter uses 1.8 fixed-point weights, so
                                             radiance is NOT bit-identical to the reference arithmetic: the stated tolerance is
                                             in tests/test_gpu_parity.py::test_hw_bilinear_...; off by default */
    PT_FLAG_RUSSIAN_ROULETTE = 1u << 8    /* EXTENSION, off by default and outside every parity claim against the reference
                                             (which has none, ray_gen.rgen:49-97): after the throughput update of a bounce
                                             (:95), from the second bounce on, the path survives with probability
                                             p = clamp(max(atten.rgb), 0.05, 1) decided by one more pcg2d draw (xi.x < p) and
                                             atten /= p — unbiased, fewer deep low-throughput segments.  The oracle implements
                                             the same rule (same flag), so GPU == oracle still holds bit for bit. */
};

typedef struct PtRenderConfig {
    uint32_t width;        /* --resolution-x (1920) */
    uint32_t height;       /* --resolution-y (1080) */
    uint32_t num_samples;  /* --num-samples  (1)  NUM_SAMPLES spec constant, ray_gen.rgen:17 */
    uint32_t num_bounces;  /* --num-bounces  (2)  NUM_BOUNCES = ray segments,  ray_gen.rgen:18 */
    uint32_t divider;      /* --render-resolution-divider (1), RaytracingPass.zig:149-152 */
    int32_t device;        /* CUDA device ordinal */
    int32_t rank;          /* tile owner id   (PT_PARTITION_TILES) */
    int32_t world;         /* number of ranks (PT_PARTITION_TILES) */
    uint32_t partition;    /* PT_PARTITION_* */
    uint32_t flags;        /* PT_FLAG_* */
} PtRenderConfig;

/* Per-frame input == PushConstants (src/RayTracingPipeline.zig:15-22, filled at
 * src/RaytracingPass.zig:563-569). */
typedef struct PtCamera {
    float position[4];
    float horizontal[4];
    float vertical[4];
    float forward[4];
    uint32_t frame_count;
} PtCamera;

/* Closest-hit record == Payload (src/shaders/common.glsl:4-11) without the matrices (they follow
 * from `instance`).  Miss: t == 1e32f and all ids 0xFFFFFFFF. */
typedef struct PtHit {
    float t;
    float b1;
    float b2;
    uint32_t instance;
    uint32_t primitive_index;  /* meshes[mesh].start + geometry, closest_hit.rchit:11 */
    uint32_t triangle_index;   /* gl_PrimitiveID,                closest_hit.rchit:12 */
} PtHit;

typedef struct PtRay {
    float origin[3];
    float tmin;
    float dir[3];
    float tmax;
} PtRay;

typedef struct PtStats {
    uint64_t rays_traced;        /* closest-hit queries executed by the last pt_render */
    uint64_t rays_hit;           /* ... of which hit geometry (the rest ended in the sky) */
    uint32_t wavefront_iterations;
    uint32_t kernel_launches;    /* kernels launched by the last pt_render */
    float ms_trace;              /* CUDA-event time of the traversal launches of the last frame
                                    (only when profiling is enabled with pt_set_profiling) */
    float ms_shade;
    float ms_other;
    float ms_total;
    uint64_t bvh_nodes;          /* BVH8 nodes (all BLAS + TLAS) */
    uint64_t bvh_bytes;
    uint64_t num_triangles;
    float ms_build;              /* device BVH build time in pt_create */
} PtStats;

typedef struct PtRenderer PtRenderer;
typedef struct PtScene PtScene;

/* ------------------------------------------------------------------ errors */
const char* pt_last_error(void);
const char* pt_status_name(int status);

/* ------------------------------------------------------------------ Scene.load  (src/Scene.zig:79-154)
 * glTF 2.0 (.gltf with external .bin/.png, or .glb) -> host arrays with the exact layout of the
 * reference loader.  `*out` owns the arrays; pt_scene_desc() borrows them. */
int pt_scene_load(const char* path, PtScene** out);
int pt_scene_desc(const PtScene* scene, PtSceneDesc* out);
void pt_scene_free(PtScene* scene);

/* ------------------------------------------------------------------ Camera  (src/Camera.zig:25-46,86-102)
 * Camera.new(fov, aspect, position) followed by updateOrientation/updateVectors with the given
 * yaw/pitch in degrees (0,0 = reference start-up state, src/main.zig:220-224). */
int pt_camera_new(float fov, float aspect_ratio, const float position[3], float yaw_deg,
                  float pitch_deg, uint32_t frame_count, PtCamera* out);

/* ------------------------------------------------------------------ interactive camera  (src/Camera.zig:10-102)
 * The whole Camera value, so that a recorded input/timer trace can be replayed headless.
 * Angles in degrees, rotation as (w, x, y, z). */
typedef struct PtCameraState {
    float fov, aspect_ratio;
    float position[3];
    float rotation[4];
    float pitch, yaw;
    float forward[3], left[3], up[3];
    float horizontal[3], vertical[3];
} PtCameraState;

enum PtKey {            /* the keys Camera.update reads (Camera.zig:65-78) */
    PT_KEY_W = 1u << 0, PT_KEY_S = 1u << 1, PT_KEY_D = 1u << 2, PT_KEY_A = 1u << 3,
    PT_KEY_SPACE = 1u << 4, PT_KEY_LEFT_CTRL = 1u << 5, PT_KEY_LEFT_SHIFT = 1u << 6
};

/* One frame of Input + Timer as Camera.update sees them (Input.zig:16-19, Timer.zig:6). */
typedef struct PtInput {
    uint32_t keys;                 /* PT_KEY_* currently held */
    uint32_t mouse_left;           /* isMouseButtonPressed(.left) */
    int32_t cursor_delta_x, cursor_delta_y;
    float delta_time;              /* Timer.delta_time, seconds */
} PtInput;

/* Camera.new (Camera.zig:25-46) */
int pt_camera_state_new(float fov, float aspect_ratio, const float position[3], PtCameraState* out);
/* Camera.update (Camera.zig:48-79): mouse look while the left button is held, WASD / space / ctrl motion */
int pt_camera_update(PtCameraState* camera, const PtInput* input);
/* Camera.updateAspectRatio (Camera.zig:81-84), called after a resize (main.zig:346) */
int pt_camera_update_aspect_ratio(PtCameraState* camera, float aspect_ratio);
/* the push-constant block of one frame (RaytracingPass.zig:563-569) */
int pt_camera_pack(const PtCameraState* camera, uint32_t frame_count, PtCamera* out);

/* ------------------------------------------------------------------ RaytracingPass
 * init    src/RaytracingPass.zig:87-200  (scene upload, BLAS/TLAS build -> GPU BVH8 builder) */
int pt_create(const PtSceneDesc* scene, const PtRenderConfig* config, PtRenderer** out);
/* RaytracingPass.init as the reference calls it, with a scene PATH (RaytracingPass.zig:87-103: Scene.load, upload,
 * `defer scene.deinit()`): pt_scene_load + pt_create + pt_scene_free in one call.  Loader errors keep their Zig names. */
int pt_create_from_gltf(const char* scene_path, const PtRenderConfig* config, PtRenderer** out);
/* frames in flight (the reference: one command buffer + fence per swapchain image, src/main.zig:209-217,255-259): a second
 * frame context over the SAME device scene and acceleration structures -- own stream, path state, image and
 * read-back pipeline.  `parent` must outlive it and owns the scene (pt_set_instance_transforms goes to the parent,
 * with no frame in flight on either; the sibling picks the new top level up at its next pt_render).  pt_resize
 * applies to one context at a time.  Alternate pt_render between the two to overlap consecutive frames. */
int pt_create_sibling(PtRenderer* parent, PtRenderer** out);
/* deinit  src/RaytracingPass.zig:202-222 */
void pt_destroy(PtRenderer* r);
/* record  src/RaytracingPass.zig:520-675: enqueue one frame on `cuda_stream` (a cudaStream_t, or
 * NULL for the renderer's own stream) and return immediately; like the reference the image is
 * overwritten every frame (ray_gen.rgen:21,104). */
int pt_render(PtRenderer* r, const PtCamera* camera, void* cuda_stream);
/* the frame fence  src/Swapchain.zig:60-62 */
int pt_sync(PtRenderer* r);
/* resize  src/RaytracingPass.zig:677-704.  Like the reference it re-creates the storage image at
 * the full new extent (the divider is dropped, quirk Q13). */
int pt_resize(PtRenderer* r, uint32_t width, uint32_t height);

/* Output image (storage image, ray_gen.rgen:104).  Render extent = resolution / divider. */
int pt_render_extent(const PtRenderer* r, uint32_t* width, uint32_t* height);
/* linear RGBA fp32, row 0 = top, alpha 1 */
int pt_read_rgba32f(PtRenderer* r, float* host_dst);
int pt_device_image(PtRenderer* r, float** device_ptr);
/* The reference's own pixel format: b8g8r8a8_unorm (src/Swapchain.zig:115-118): NaN->0, clamp,
 * round-to-nearest 8 bit.  Converted on the device, then copied. */
int pt_read_bgra8(PtRenderer* r, uint8_t* host_dst);
/* Asynchronous variants for pipelined hosts: convert/copy on the render stream into pinned host
 * memory supplied by the caller; complete after pt_sync or a sync of the stream. */
int pt_read_rgba32f_async(PtRenderer* r, float* pinned_host_dst, void* cuda_stream);
int pt_read_bgra8_async(PtRenderer* r, uint8_t* pinned_host_dst, void* cuda_stream);
/* Pipelined variant for frame loops (the analogue of presenting image i while frame i+1 is recorded,
 * main.zig:232-345): the copy is ordered after the last pt_render but runs on the renderer's own copy
 * stream; the next pt_render only waits for it before its first write to the image, so the copy overlaps
 * ray generation and primary traversal of the next frame.  Complete after pt_sync. */
int pt_read_rgba32f_overlapped(PtRenderer* r, float* pinned_host_dst);
/* Primary-hit id buffers of the last frame (parity: instance / primitive_index / triangle_index,
 * 0xFFFFFFFF on miss).  Any pointer may be NULL. */
int pt_read_ids(PtRenderer* r, uint32_t* instance, uint32_t* primitive, uint32_t* triangle);
/* Full-resolution output: nearest up-scale of the render extent when divider > 1 (the blit at
 * src/RaytracingPass.zig:599-635), plain copy otherwise. */
int pt_read_rgba32f_upscaled(PtRenderer* r, float* host_dst);

/* Progressive accumulation over frame_count (optional, off by default: the reference overwrites).
 * When enabled pt_render adds the frame into a running sum; pt_read_accum_mean divides. */
int pt_set_accumulate(PtRenderer* r, int enabled);
int pt_reset_accum(PtRenderer* r);
int pt_device_accum(PtRenderer* r, float** device_ptr, uint32_t* frames);
int pt_read_accum_mean(PtRenderer* r, float* host_dst, uint32_t total_frames);

/* Measurement. */
int pt_set_profiling(PtRenderer* r, int enabled);
int pt_get_stats(PtRenderer* r, PtStats* out);

/* Kernel-level access for parity tests: trace `n` arbitrary world-space rays through the same
 * BVH and traversal kernel pt_render uses (host pointers).  n < 2^31 (PT_ERR_INVALID_ARGUMENT beyond: the
 * kernel keeps a flag in bit 31 of a ray's slot index; the same bound holds for the pixel slots of a frame). */
int pt_trace_rays(PtRenderer* r, const PtRay* rays, size_t n, PtHit* hits);
/* The same query for callers that cannot allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * `device_rays` / `device_hits` are DEVICE pointers on the renderer's GPU
 * (n records each, caller-owned); the work is enqueued on `cuda_stream` (NULL = the renderer's own
 * stream, which pt_sync waits for) and not waited for.  The call makes no cudaMalloc / cudaFree or
 * other allocating call: the kernel's scratch (per-ray traversal records, hit records, the fetch
 * cursor) is one fixed 14 MiB __device__ array inside the library (held by every GPU that loads
 * it, until exit), sized for PT_NOALLOC_MAX_RAYS rays; a larger n returns
 * PT_ERR_INVALID_ARGUMENT and launches nothing (split the batch, or use pt_trace_rays).  Every call
 * on a device shares that array: calls on one stream are ordered, calls that overlap on the GPU
 * (other streams, renderers or threads) race.  Default traversal kernel only (a renderer created
 * with a PT_FLAG_TRACE_* flag is rejected).  Hits are bit-identical to pt_trace_rays'. */
#define PT_NOALLOC_MAX_RAYS 131072u
int pt_trace_rays_noalloc(PtRenderer* r, const PtRay* device_rays, size_t n, PtHit* device_hits, void* cuda_stream);
/* Independent, unoptimised one-thread-per-ray walker over the shipped BVH8: counts the nodes and
 * triangles an ordered closest-hit traversal must decode (SURVEY.md §8d algorithmic bytes =
 * 80*nodes + 48*tris).  Counts the ray set of the last pt_render. */
int pt_count_traversal(PtRenderer* r, uint64_t* nodes, uint64_t* tris, uint64_t* rays);

/* Device evaluation of the shading restatement on caller-supplied records (host pointers), for
 * pinning the CUDA BSDF / RNG / texture filter against the reference's shader code:
 *   sample_brdf -> L, get_brdf_density -> pdf, brdf -> f   (src/shaders/brdf.glsl:22-150)
 *   prng_seed + `draws` x pcg2d                            (src/shaders/rand.glsl:76-94)
 *   texture(): bilinear REPEAT RGBA8                        (src/RayTracingPipeline.zig:54-70) */
int pt_debug_bsdf(int device, size_t n, const float* normal3, const float* view3, const float* diffuse3,
                  const float* f0_3, const float* roughness, const float* xi2, float* L3, float* pdf, float* f3);
int pt_debug_rng(int device, size_t n, const uint32_t* pixel_xy_frame3, uint32_t draws, uint32_t* state2, float* xi2);
int pt_debug_texture(int device, const uint8_t* rgba, uint32_t width, uint32_t height, size_t n, const float* uv2,
                     float* rgba4);

/* Digest of the device acceleration structures (determinism probes): [0] Morton-ordered triangles, [1] binary
 * topology, [2] binary boxes, [3] BVH8 node count, [4] order-independent hash of the BVH8 nodes, [5] BLAS roots. */
int pt_debug_bvh_digest(PtRenderer* r, uint64_t digest[6]);

/* Animated instances: replace all instance transforms (num_instances x 16 floats, the
 * rows of M as in PtInstance.transform; mesh indices unchanged) and rebuild the top-level structure on the
 * device -- Tlas.init (Tlas.zig:15-182) run again.  Synchronises with frames in flight; *ms_rebuild (may be
 * NULL) receives the device time of the rebuild.  Resets the progressive accumulation. */
int pt_set_instance_transforms(PtRenderer* r, const float* transforms, size_t num_instances, float* ms_rebuild);

/* Kernel tuning knob (measurement only; results never depend on it), applied to renderers created afterwards:
 *   "chunk": rays a K-trace warp takes from the queue per global atomic (default 32). */
int pt_set_option(const char* key, int value);

/* ------------------------------------------------------------------ multi-GPU combine  (north_star: "per-rank
 * accumulation buffers are combined with an NCCL reduce over NVLink")
 * The reference renders on one device; this is the additive multi-GPU surface.  Every rank holds the whole scene and
 * renders its share — interleaved tiles of one frame (PtRenderConfig.partition = PT_PARTITION_TILES, rank, world) or its
 * own frames (frame_count = k * world + rank, pt_set_accumulate) — and one reduce(sum) to `root` combines the shares.
 * NCCL is opened at run time (libnccl.so.2; override with B200PT_NCCL_LIB); without it these calls fail with PT_ERR_NCCL.
 *
 *   one process per GPU:   rank 0: pt_comm_unique_id(id) -> ship the 128 bytes to the other ranks (any channel) ->
 *                          every rank: pt_comm_init(r, rank, world, id)
 *   one process, N GPUs:   pt_comm_init_all(renderers, n)   (renderer i = rank i; wrap the per-rank reduce calls of one
 *                          step in pt_comm_group_start / pt_comm_group_end when a single thread issues them)
 *
 * pt_reduce_image / pt_reduce_accum: in-place sum of the image / the accumulation buffer over all ranks into root's.
 * cuda_stream == NULL: the reduce runs on the library's communication stream, ordered after everything enqueued on the
 * renderer's render stream so far; the renderer's next frame waits for it only before it writes the image again, and
 * read-backs (pt_read_*) are ordered after it — so a reduce overlaps the ray generation and primary traversal of the next
 * frame.  Tile shares are disjoint and zero elsewhere, so the reduced image equals the one-GPU image bit for bit; for
 * frame shares read the mean with pt_read_accum_mean(r, dst, total_frames).  A renderer's siblings use its communicator. */
int pt_comm_available(int* nccl_version);                          /* 1 when NCCL could be opened (version code in *nccl_version) */
int pt_comm_unique_id(uint8_t id[128]);                            /* ncclGetUniqueId */
int pt_comm_init(PtRenderer* r, int rank, int world, const uint8_t id[128]);
int pt_comm_init_all(PtRenderer** renderers, int n);               /* ncclCommInitAll over the renderers' devices */
int pt_comm_rank(PtRenderer* r, int* rank, int* world);
int pt_comm_group_start(void);
int pt_comm_group_end(void);
int pt_reduce_image(PtRenderer* r, int root, void* cuda_stream);
int pt_reduce_accum(PtRenderer* r, int root, void* cuda_stream);
int pt_comm_destroy(PtRenderer* r);                                /* also done by pt_destroy */
/* dst.accum += src.accum, dst.frames += src.frames (same device and extent): folds the accumulation of a sibling context
 * (two frames in flight) into its parent's before the reduce. */
int pt_add_accum(PtRenderer* dst, PtRenderer* src);

/* Multi-GPU without a collective (one process per GPU).  Peer stores: rank 0 exports its image buffer as a CUDA IPC
 * handle, the other ranks open it and their accumulate kernel writes owned tiles straight into it
 * over NVLink.  `handle` is 64 bytes. */
int pt_ipc_export_image(PtRenderer* r, uint8_t handle[64]);
int pt_ipc_open_root_image(PtRenderer* r, const uint8_t handle[64]);

/* Headless image output (replaces the swapchain present). */
int pt_write_pfm(const char* path, const float* rgba, uint32_t width, uint32_t height);
int pt_write_exr(const char* path, const float* rgba, uint32_t width, uint32_t height);

/* Library/device introspection. */
/* measurement (SURVEY 8d "which roofline"): read bandwidth of a `bytes`-sized device buffer over `iters` coalesced passes;
 * a buffer that fits the L2 gives the L2 read bandwidth that bounds the traversal working set, a large one HBM's.  Also
 * returns the SM count (FP32 / issue peaks = SMs x 4 schedulers x clock). */
int pt_debug_read_bandwidth(int device, size_t bytes, int iters, float* gb_per_s, int* num_sms);
int pt_device_count(void);
const char* pt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200PT_H */
