// bvh_build.h — host interface of the GPU BVH builder (bvh_build.cu).
#pragma once
#include <vector>

#include "pt_internal.h"

namespace pt {

struct SceneBvh {
    // BVH8 (shipped structure)
    Bvh8Node* nodes8 = nullptr;
    uint32_t num_nodes8 = 0;
    Bvh8Tri* tris8 = nullptr;
    uint32_t num_tris = 0;
    uint32_t* tlas_leaf_instance = nullptr;
    uint32_t tlas_root8 = kInvalidId;
    std::vector<uint32_t> blas_root8;   // per mesh, kInvalidId for empty meshes
    // BVH2 (LBVH or PLOC trees; input of the collapse, walked by the reference kernel)
    Bvh2Node* nodes2 = nullptr;         // per mesh at tri_off[m]
    Bvh8Tri* tris2 = nullptr;           // Morton-ordered triangles per mesh at tri_off[m]
    Bvh2Node* tlas2 = nullptr;
    uint32_t* tlas2_leaf_instance = nullptr;  // sorted TLAS leaf -> instance id
    uint32_t tlas2_count = 0;           // number of TLAS leaves (instances with a non-empty mesh)
    std::vector<uint32_t> tri_off;      // per mesh
    std::vector<uint32_t> tri_count;    // per mesh
    float* mesh_bounds = nullptr;       // device, 6 floats per mesh
    uint2* mesh_tri_range = nullptr;    // device, per mesh: (first triangle in tris2, count)
    uint32_t max_mesh_tris = 0;
    bool exact_instance_boxes = false;  // TLAS leaves from transformed vertices instead of transformed mesh-box corners
    // TLAS rebuild state (animated instances)
    uint32_t tlas_node_base = 0;        // first BVH8 node of the TLAS (everything below belongs to BLASes)
    std::vector<uint32_t> valid_inst;   // instances that reference a non-empty mesh
    uint32_t builder_flags = 0;         // PT_FLAG_BVH_* the scene was built with (a TLAS rebuild uses the same stages)
    float ms_build = 0.0f;
    uint32_t refit_repairs = 0;         // nodes corrected by the refit verification (expected 0)

    void release();
};

// d_* are device pointers, h_* their host copies.
int build_scene_bvh(const uint8_t* d_blob, const PrimRec* d_prims, const std::vector<PrimRec>& h_prims,
                    const std::vector<PtMesh>& meshes, const PtInstance* d_instances,
                    const std::vector<PtInstance>& h_instances, uint32_t builder_flags, cudaStream_t stream, SceneBvh* out);

// Rebuilds the top level in place after the transforms in d_instances changed (mesh indices unchanged).
int rebuild_tlas(SceneBvh* bvh, const PtInstance* d_instances, cudaStream_t stream, float* ms_out);

}  // namespace pt
