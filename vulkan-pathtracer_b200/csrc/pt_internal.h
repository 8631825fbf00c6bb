// pt_internal.h — data layout in HBM shared by the builder, traversal and shading kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/b200pt.h"
#include "pt_math.cuh"

namespace pt {

// ------------------------------------------------------------------ error plumbing
void set_error(const std::string& msg);
#define PT_CUDA_CHECK(expr)                                                                     \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess) {                                                                \
            pt::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + \
                          ":" + std::to_string(__LINE__) + ")");                                \
            return PT_ERR_CUDA;                                                                 \
        }                                                                                       \
    } while (0)

// ------------------------------------------------------------------ compressed 8-wide BVH node
// 80 bytes = 5 x 16 B, 16-byte aligned (Ylitie, Karras, Laine 2017 layout):
//   n0: origin p.xyz (3 x f32) | ex, ey, ez (biased exponents, u8) | imask (u8: slots holding inner nodes)
//   n1: child_base (u32: index of the first inner child) | tri_base (u32: index of the first leaf
//       triangle / instance) | valid (u32: bit 24 + s = slot s holds an inner node (== imask); bits 2s, 2s + 1 = unary
//       count (01, 11) of the triangles / instances in leaf slot s; both clear = empty slot) | spare (0).
//       Leaf items are stored densely in slot order: item k of leaf slot s sits at tri_base + popc(valid & ((1 << 2s) - 1)) + k.
//       A node test ANDs its candidate hit bits with `valid`, so no per-slot table is loaded (round 2 until r2v kept the eight
//       pre-expanded hit-mask words of a node in a second array: two more LDG.128 and a second memory round trip per node test).
//   n2: qlo_x[8] | qlo_y[8]      n3: qlo_z[8] | qhi_x[8]      n4: qhi_y[8] | qhi_z[8]
// child box = p + q * 2^(e-127)   (lo rounded down, hi rounded up, with slack, see bvh_build.cu)
struct __align__(16) Bvh8Node {
    float px, py, pz;
    uint8_t ex, ey, ez, imask;
    uint32_t child_base, tri_base;
    uint32_t valid, spare;
    uint8_t qlo_x[8], qlo_y[8];
    uint8_t qlo_z[8], qhi_x[8];
    uint8_t qhi_y[8], qhi_z[8];
};
static_assert(sizeof(Bvh8Node) == 80, "Bvh8Node must be 80 bytes");
constexpr uint32_t kLeafStride = 2;      // bits of `valid` per leaf slot = triangles a leaf slot can hold

// Leaf triangle: 48 bytes = 3 x float4; ids ride in the w lanes.
//   v0.xyz, w = primitive_index (bits) | v1.xyz, w = triangle_index (bits) | v2.xyz, w = unused
struct __align__(16) Bvh8Tri { float4 v0, v1, v2; };
static_assert(sizeof(Bvh8Tri) == 48, "Bvh8Tri must be 48 bytes");

// Plain BVH2 produced by the LBVH stage (input of the collapse; also walked by the reference kernel).
// Nodes [0, n-1) are internal, leaf i is encoded as child index (i | 0x80000000).
struct __align__(16) Bvh2Node {
    float lo[3]; uint32_t left;
    float hi[3]; uint32_t right;
};

// Per instance: world->object for rays, object->world linear part for nothing but completeness,
// and the BLAS it references.  64 bytes.
struct __align__(16) InstanceRec {
    float inv[12];       // rows of [Ainv | -Ainv*T is NOT folded: o' = Ainv*(o - T)]: inv[0..8] = Ainv row-major, inv[9..11] = T
    uint32_t blas_root;  // BVH8 node index of the mesh's root
    uint32_t bvh2_root;  // BVH2 node index (reference walker)
    uint32_t mesh;       // mesh index | kInstIdentity
    uint32_t bvh2_tri_base;  // first sorted triangle of the mesh (BVH2 leaves)
};
static_assert(sizeof(InstanceRec) == 64, "InstanceRec must be 64 bytes");
// set in InstanceRec::mesh when Ainv is exactly the identity and T is zero: object space == world
// space bit for bit, so K-trace enters the BLAS without re-deriving the ray frame
constexpr uint32_t kInstIdentity = 0x80000000u;

// Device-side primitive record: byte offsets into the triangle_data blob (the reference keeps 5
// device addresses, RaytracingPass.zig:18-31; offsets suffice with one blob).  32 bytes.
struct __align__(16) PrimRec {
    uint32_t indices_offset, positions_offset, normals_offset, tangents_offset;
    uint32_t uvs_offset, info, triangle_count, _pad;
};

struct TexRec { const uint32_t* texels; uint32_t width, height; };  // RGBA8 packed, row 0 first

// Device material: the reference's 28-byte Material (material.glsl:7-16) with the four texture
// indices resolved to texel pointers + extents (texels == nullptr: no texture), so that shading
// needs one dependent load less per texture.  When at least two textures are present and all
// present ones have the same extent, `atlas` holds them interleaved per texel as one uint4
// (albedo, metal-roughness, emissive, normal RGBA8 words): the three or four bilinear taps that
// material.glsl takes at the same uv then cost four 16-byte loads and about a third of the DRAM
// sectors of separate textures.  `present` = bit k set when texture k exists.  96 bytes.
struct __align__(16) MatRec {
    uint32_t albedo_factor, metal_roughness_factor, emissive_factor, present;
    TexRec tex[4];   // albedo, metal-roughness, emissive, normal
    const uint4* atlas; uint32_t atlas_width, atlas_height;
};
static_assert(sizeof(MatRec) == 96, "MatRec must be 96 bytes");

struct DeviceScene {
    const uint8_t* triangle_data;
    const PrimRec* prims;
    const MatRec* materials;
    const TexRec* textures;
    const InstanceRec* instances;
    const Bvh8Node* nodes;       // all BLASes + TLAS
    const Bvh8Tri* tris;         // leaf triangles (BLAS) in collapse order
    const uint32_t* tlas_leaf_instance;  // TLAS leaf slots -> instance index
    const Bvh2Node* nodes2;      // LBVH trees (reference walker)
    const Bvh8Tri* tris2;        // triangles in Morton order per mesh (BVH2 leaves)
    const Bvh2Node* tlas2;       // BVH2 over instances
    uint32_t tlas_root;          // BVH8 node index of the TLAS root
    // where a ray starts: (tlas_root, kInvalidId) in general; when the scene is one instance whose transform
    // is exactly the identity the top level would be one node test with one child for every ray, so traversal
    // starts at that instance's BLAS root with the instance already entered
    uint32_t start_root, start_inst;
    uint32_t num_instances, num_prims, num_materials, num_textures;
    // set to 1 by a walker that had to drop a stack push (hierarchy deeper than the fixed traversal stack): host-mapped, so
    // pt_sync / pt_get_stats can report PT_ERR_TRAVERSAL_STACK instead of returning an image with missed hits
    uint32_t* error_flag;
    // PT_FLAG_HW_BILINEAR: 4 texture objects per material (albedo, metal-roughness, emissive, normal; 0 = absent), else null
    const unsigned long long* mat_tex_objs;
    uint32_t atlas_only;         // 1 when every material that has maps has them interleaved (MatRec::atlas): K-shade's atlas-only instantiation
};

// Hit record in HBM: 32 bytes (2 x 16 B)
struct __align__(16) HitRec {
    float t, b1, b2; uint32_t instance;
    uint32_t primitive_index, triangle_index, _pad0, _pad1;
};

// Ray in HBM: 32 bytes
struct __align__(16) RayRec { float ox, oy, oz, tmin, dx, dy, dz, tmax; };

constexpr float kMissT = 1e32f;
constexpr uint32_t kInvalidId = 0xFFFFFFFFu;

}  // namespace pt
