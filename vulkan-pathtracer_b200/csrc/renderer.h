// renderer.h — the renderer object behind the C ABI (RaytracingPass, src/RaytracingPass.zig:53-86): device scene, frame
// state, streams.  Shared by api.cu (the ABI) and comm.cpp (multi-GPU combine).
#pragma once
#include <cuda_runtime.h>

#include <vector>

#include "bvh_build.h"
#include "render.h"

struct PtComm;   // comm.cpp: NCCL communicator + communication stream of one rank

struct PtRenderer {
    PtRenderConfig cfg{};
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t last_stream = nullptr;   // stream of the most recent pt_render (caller's or our own)
    // pipelined read-back (pt_read_*_overlapped): the copy runs on its own stream, the next frame only
    // waits for it before its first image write (K-shade of iteration 0)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_frame_done = nullptr, ev_copy_done = nullptr;
    bool copy_pending = false;
    // scene on the device
    uint8_t* d_blob = nullptr;
    pt::PrimRec* d_prims = nullptr;
    pt::MatRec* d_materials = nullptr;
    pt::TexRec* d_textures = nullptr;
    std::vector<uint32_t*> d_texels;
    std::vector<uint4*> d_atlases;   // interleaved material textures (MatRec::atlas)
    // PT_FLAG_HW_BILINEAR: one CUDA array + texture object per glTF texture, and the per-material table of objects
    std::vector<cudaArray_t> tex_arrays;
    std::vector<cudaTextureObject_t> tex_objects;
    unsigned long long* d_mat_tex_objs = nullptr;
    PtInstance* d_instances_raw = nullptr;
    pt::InstanceRec* d_instances = nullptr;
    std::vector<PtInstance> h_instances;   // host copy (animated instances: pt_set_instance_transforms)
    pt::SceneBvh bvh;
    pt::DeviceScene scene{};
    uint64_t num_triangles = 0;
    // frame state
    pt::PixelMap pm{};
    pt::PathBuffers pb{};
    uint32_t max_iters = 0;
    uint32_t full_w = 0, full_h = 0;     // window extent (resolution-x/-y)
    float* d_image = nullptr;            // render extent, RGBA32F
    uint32_t* d_ids = nullptr;           // 3 * W*H
    uint8_t* d_bgra = nullptr;
    float* d_upscaled = nullptr;
    float* d_accum = nullptr;
    float* d_remote_image = nullptr;     // rank 0's image mapped through CUDA IPC (optional)
    bool accumulate = false;
    uint32_t accum_frames = 0;
    bool profiling = false;
    pt::TraceMode trace_mode = pt::TRACE_BVH8;
    std::vector<cudaEvent_t> events;     // profiling: 2 per kernel launch
    std::vector<int> event_kind;         // 0 trace, 1 shade, 2 other
    uint32_t last_iters = 0;
    uint32_t last_launches = 0;
    PtCamera last_cam{};
    bool have_last_cam = false;
    // standalone trace buffers
    uint32_t* d_cursor = nullptr;
    bool peer_root = false;   // pt_ipc_export_image was called: other ranks store their tiles into d_image
    bool owns_scene = true;   // false for a sibling (pt_create_sibling): scene + BVH belong to the parent
    PtRenderer* parent = nullptr;   // sibling: where the current DeviceScene is read from (the parent may rebuild its top level)
    uint32_t* h_error = nullptr;    // host-mapped traversal error flag (DeviceScene::error_flag); a sibling borrows its parent's
    // multi-GPU combine (comm.cpp): owned by the renderer pt_comm_init* was called on; siblings use their parent's
    PtComm* comm = nullptr;
    cudaEvent_t ev_reduce_done = nullptr;   // recorded on the communication stream after this renderer's last reduce
    bool reduce_pending = false;            // the next frame must wait for it before its first image / accumulation write
    // persistent grid sizes of K-trace / K-shade for THIS device (occupancy x SM count), filled by pt_create
    pt::LaunchDims dims{};
};

