// api.cu — the C ABI of include/b200pt.h: renderer object (RaytracingPass.init / record / resize /
// deinit, src/RaytracingPass.zig:87-222,520-704), image read-back, stats and test entry points.
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>
#include <string>
#include <array>
#include <map>
#include <vector>

#include "renderer.h"

namespace pt {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }

}  // namespace pt

using namespace pt;

void comm_release(PtRenderer* r);   // comm.cpp

namespace {

template <typename T>
int dev_alloc(T** p, size_t count) {
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(p), std::max<size_t>(count, 1) * sizeof(T));
    if (e != cudaSuccess) {
        set_error(std::string("cudaMalloc failed: ") + cudaGetErrorString(e));
        *p = nullptr;
        return e == cudaErrorMemoryAllocation ? PT_ERR_OUT_OF_MEMORY : PT_ERR_CUDA;
    }
    return PT_OK;
}

void free_frame(PtRenderer* r) {
    PathBuffers& b = r->pb;
    cudaFree(b.ray_o); cudaFree(b.ray_d); cudaFree(b.ray_e); cudaFree(b.ray_s); cudaFree(b.ray_m); cudaFree(b.hit_a); cudaFree(b.hit_b); cudaFree(b.prim_a); cudaFree(b.prim_b); cudaFree(b.prim_c); cudaFree(b.prim_d);
    cudaFree(b.atten_sx); cudaFree(b.color_sy); cudaFree(b.final_sb); cudaFree(b.queue[0]); cudaFree(b.queue[1]); cudaFree(b.queue_sorted); cudaFree(b.sort_bins);
    cudaFree(b.qcount); cudaFree(b.trav_counters);   // trace_next / shade_next live inside qcount's allocation
    b = PathBuffers{};
    cudaFree(r->d_image); cudaFree(r->d_ids); cudaFree(r->d_bgra); cudaFree(r->d_upscaled); cudaFree(r->d_accum);
    r->d_image = nullptr; r->d_ids = nullptr; r->d_bgra = nullptr; r->d_upscaled = nullptr; r->d_accum = nullptr;
}

// storage image = extent / divider (RaytracingPass.zig:149-152)
int alloc_frame(PtRenderer* r, uint32_t full_w, uint32_t full_h, uint32_t divider) {
    free_frame(r);
    const uint32_t div = divider ? divider : 1u;
    r->full_w = full_w; r->full_h = full_h;
    PixelMap& pm = r->pm;
    pm.width = full_w / div; pm.height = full_h / div;
    pm.tiles_x = (pm.width + 31u) / 32u; pm.tiles_y = (pm.height + 7u) / 8u;
    const bool tiles = r->cfg.partition == PT_PARTITION_TILES && r->cfg.world > 1;
    pm.rank = tiles ? (uint32_t)r->cfg.rank : 0u;
    pm.world = tiles ? (uint32_t)r->cfg.world : 1u;
    const uint32_t n_tiles = pm.tiles_x * pm.tiles_y;
    const uint32_t owned = n_tiles > pm.rank ? (n_tiles - pm.rank + pm.world - 1u) / pm.world : 0u;
    if ((uint64_t)owned * 256u >= 0x80000000ull) { set_error("extent too large (K-trace uses bit 31 of the slot index)"); return PT_ERR_INVALID_ARGUMENT; }
    pm.num_slots = owned * 256u;
    const size_t n = pm.num_slots;
    const uint32_t S = r->cfg.num_samples, B = r->cfg.num_bounces;
    // trace iterations: B for the first sample, B-1 for every further one (primary hit reused);
    // S*B when reuse is disabled.
    uint64_t iters = (S == 0 || B == 0) ? 0 : ((r->cfg.flags & PT_FLAG_NO_PRIMARY_REUSE) ? (uint64_t)S * B : (uint64_t)B + (uint64_t)(S - 1) * (B - 1));
    if (iters > (1u << 20)) { set_error("num_samples * num_bounces too large"); return PT_ERR_INVALID_ARGUMENT; }
    r->max_iters = (uint32_t)iters;
    PathBuffers& b = r->pb;
    int rc;
#define A(ptr, cnt) if ((rc = dev_alloc(&(ptr), (cnt))) != PT_OK) return rc
    A(b.ray_o, n); A(b.ray_d, n); A(b.ray_e, n); A(b.ray_s, n); A(b.ray_m, n); A(b.hit_a, n); A(b.hit_b, n); { const size_t np = r->cfg.num_samples > 1 ? n : 1; A(b.prim_a, np); A(b.prim_b, np); A(b.prim_c, np); A(b.prim_d, np); }
    A(b.atten_sx, n); A(b.color_sy, n); A(b.final_sb, n); A(b.queue[0], n); A(b.queue[1], n); A(b.queue_sorted, n); A(b.sort_bins, (size_t)r->scene.num_materials + 4);
    b.frame_words = 3u * (r->max_iters + 2u) + 4u;
    A(b.qcount, b.frame_words); A(b.trav_counters, 4);
    b.trace_next = b.qcount + ((size_t)r->max_iters + 2); b.shade_next = b.trace_next + ((size_t)r->max_iters + 2);
    b.hit_count = reinterpret_cast<unsigned long long*>(b.qcount + ((3u * (r->max_iters + 2u) + 1u) & ~1u));   // 8-byte aligned (cudaMalloc base is)
    const size_t px = (size_t)pm.width * pm.height;
    A(r->d_image, px * 4); A(r->d_ids, px * 3);
#undef A
    // on the renderer's stream: it is non-blocking, so work on the legacy default stream is not ordered with it
    PT_CUDA_CHECK(cudaMemsetAsync(r->d_image, 0, std::max<size_t>(px, 1) * 16, r->stream));
    PT_CUDA_CHECK(cudaMemsetAsync(r->d_ids, 0xff, std::max<size_t>(px, 1) * 12, r->stream));
    PT_CUDA_CHECK(cudaMemsetAsync(b.trav_counters, 0, 16, r->stream));
    PT_CUDA_CHECK(cudaStreamSynchronize(r->stream));
    r->accum_frames = 0;
    return PT_OK;
}

// instance records from r->h_instances: inverse of the 3x3 linear part (host, same formula as every other user)
static int upload_instance_records(PtRenderer* r) {
    const std::vector<PtInstance>& inst = r->h_instances;
    std::vector<InstanceRec> recs(inst.size());
    for (size_t i = 0; i < inst.size(); ++i) {
        const float* t = inst[i].transform;   // rows of M; first 12 = 3x4 (Tlas.zig:33-41)
        M3 A;
        for (int rr = 0; rr < 3; ++rr) for (int c = 0; c < 3; ++c) A.m[rr][c] = t[4 * rr + c];
        const M3 Ai = inverse(A);
        InstanceRec& ir = recs[i];
        for (int rr = 0; rr < 3; ++rr) for (int c = 0; c < 3; ++c) ir.inv[3 * rr + c] = Ai.m[rr][c];
        ir.inv[9] = t[3]; ir.inv[10] = t[7]; ir.inv[11] = t[11];
        const uint32_t m = (uint32_t)inst[i].mesh_index;
        bool identity = t[3] == 0.0f && t[7] == 0.0f && t[11] == 0.0f;
        for (int rr = 0; rr < 3; ++rr) for (int c = 0; c < 3; ++c) identity = identity && Ai.m[rr][c] == (rr == c ? 1.0f : 0.0f);
        ir.mesh = m | (identity ? kInstIdentity : 0u);
        ir.blas_root = r->bvh.blas_root8[m];
        ir.bvh2_root = r->bvh.tri_off[m];
        ir.bvh2_tri_base = r->bvh.tri_off[m];
    }
    // Uploads go through the renderer's stream.  A plain cudaMemcpy from pageable memory returns once the data
    // is staged, while the DMA itself runs on the legacy default stream -- which a non-blocking stream does
    // not wait for, so a kernel on r->stream could read the buffer before the copy has landed (seen twice in
    // this round as a BVH built from partly stale triangle data; profiles/r1_experiments.md).
    if (!recs.empty()) PT_CUDA_CHECK(cudaMemcpyAsync(r->d_instances, recs.data(), recs.size() * sizeof(InstanceRec), cudaMemcpyHostToDevice, r->stream));
    PT_CUDA_CHECK(cudaStreamSynchronize(r->stream));
    // traversal start (DeviceScene::start_root): skip the top level for a single identity instance
    r->scene.start_root = r->bvh.tlas_root8; r->scene.start_inst = kInvalidId;
    if (r->bvh.valid_inst.size() == 1) {
        const uint32_t i = r->bvh.valid_inst[0];
        if (recs[i].mesh & kInstIdentity) { r->scene.start_root = recs[i].blas_root; r->scene.start_inst = i; }
    }
    return PT_OK;
}

int upload_scene(PtRenderer* r, const PtSceneDesc* s) {
    int rc;
    if (s->triangle_data_len > 0xFFFFFFF0ull) { set_error("triangle_data larger than 4 GiB is not supported"); return PT_ERR_INVALID_ARGUMENT; }
    if ((rc = dev_alloc(&r->d_blob, s->triangle_data_len + 16)) != PT_OK) return rc;
    if (s->triangle_data_len) PT_CUDA_CHECK(cudaMemcpyAsync(r->d_blob, s->triangle_data, s->triangle_data_len, cudaMemcpyHostToDevice, r->stream));
    // primitives: validate every offset against the blob (the reference trusts the file; we must not fault)
    std::vector<PrimRec> prims(s->num_primitives);
    for (size_t i = 0; i < s->num_primitives; ++i) {
        const PtPrimitive& p = s->primitives[i];
        const uint64_t nv = (uint64_t)p.max_vertex + 1, isz = (p.info & PT_UINT32_INDICES_MASK) ? 4 : 2;
        const bool ok = p.indices_offset + 3ull * p.triangle_count * isz <= s->triangle_data_len && p.positions_offset + 12 * nv <= s->triangle_data_len &&
                        p.normals_offset + 12 * nv <= s->triangle_data_len && p.tangents_offset + 16 * nv <= s->triangle_data_len &&
                        p.uvs_offset + 8 * nv <= s->triangle_data_len && (p.indices_offset % isz) == 0 && (p.positions_offset % 4) == 0 &&
                        (p.normals_offset % 4) == 0 && (p.tangents_offset % 4) == 0 && (p.uvs_offset % 4) == 0;
        if (!ok) { set_error("primitive " + std::to_string(i) + ": offsets outside triangle_data or misaligned"); return PT_ERR_INVALID_ARGUMENT; }
        if ((p.info & PT_MATERIAL_INDEX_MASK) >= s->num_materials) { set_error("primitive " + std::to_string(i) + ": material index out of range"); return PT_ERR_INVALID_ARGUMENT; }
        // vertex indices must stay inside the attribute arrays
        if (p.triangle_count) {
            uint32_t mx = 0;
            if (isz == 2) { const uint16_t* q = reinterpret_cast<const uint16_t*>(s->triangle_data + p.indices_offset); for (uint64_t k = 0; k < 3ull * p.triangle_count; ++k) mx = std::max<uint32_t>(mx, q[k]); }
            else { const uint32_t* q = reinterpret_cast<const uint32_t*>(s->triangle_data + p.indices_offset); for (uint64_t k = 0; k < 3ull * p.triangle_count; ++k) mx = std::max(mx, q[k]); }
            if (mx > p.max_vertex) { set_error("primitive " + std::to_string(i) + ": vertex index beyond max_vertex"); return PT_ERR_INVALID_ARGUMENT; }
        }
        prims[i] = PrimRec{(uint32_t)p.indices_offset, (uint32_t)p.positions_offset, (uint32_t)p.normals_offset, (uint32_t)p.tangents_offset,
                           (uint32_t)p.uvs_offset, p.info, p.triangle_count, 0};
        r->num_triangles += p.triangle_count;
    }
    if ((rc = dev_alloc(&r->d_prims, prims.size())) != PT_OK) return rc;
    if (!prims.empty()) PT_CUDA_CHECK(cudaMemcpyAsync(r->d_prims, prims.data(), prims.size() * sizeof(PrimRec), cudaMemcpyHostToDevice, r->stream));
    // materials + textures (r8g8b8a8_unorm, single mip; RaytracingPass.zig:490-518)
    for (size_t i = 0; i < s->num_materials; ++i) {
        const uint32_t t[4] = {s->materials[i].albedo_texture_index, s->materials[i].metal_roughness_texture_index,
                               s->materials[i].emissive_texture_index, s->materials[i].normal_texture_index};
        for (uint32_t v : t)
            if (v != PT_INVALID_TEXTURE && v >= s->num_textures) { set_error("material " + std::to_string(i) + ": texture index out of range"); return PT_ERR_INVALID_ARGUMENT; }
    }
    std::vector<TexRec> texs(s->num_textures);
    r->d_texels.assign(s->num_textures, nullptr);
    for (size_t i = 0; i < s->num_textures; ++i) {
        const PtTexture& t = s->textures[i];
        if (t.width == 0 || t.height == 0 || !t.rgba) { set_error("texture " + std::to_string(i) + ": empty"); return PT_ERR_INVALID_ARGUMENT; }
        const size_t n = (size_t)t.width * t.height;
        if ((rc = dev_alloc(&r->d_texels[i], n)) != PT_OK) return rc;
        PT_CUDA_CHECK(cudaMemcpyAsync(r->d_texels[i], t.rgba, n * 4, cudaMemcpyHostToDevice, r->stream));
        texs[i] = TexRec{r->d_texels[i], t.width, t.height};
    }
    if ((rc = dev_alloc(&r->d_textures, texs.size())) != PT_OK) return rc;
    if (!texs.empty()) PT_CUDA_CHECK(cudaMemcpyAsync(r->d_textures, texs.data(), texs.size() * sizeof(TexRec), cudaMemcpyHostToDevice, r->stream));
    std::vector<MatRec> mats(s->num_materials);
    std::map<std::array<uint32_t, 4>, uint4*> atlas_of;
    for (size_t i = 0; i < s->num_materials; ++i) {
        const PtMaterial& m = s->materials[i];
        MatRec& d = mats[i];
        d.albedo_factor = m.albedo_factor; d.metal_roughness_factor = m.metal_roughness_factor; d.emissive_factor = m.emissive_factor; d.present = 0;
        const uint32_t t[4] = {m.albedo_texture_index, m.metal_roughness_texture_index, m.emissive_texture_index, m.normal_texture_index};
        for (int k = 0; k < 4; ++k) d.tex[k] = (t[k] == PT_INVALID_TEXTURE) ? TexRec{nullptr, 0, 0} : texs[t[k]];
        d.atlas = nullptr; d.atlas_width = 0; d.atlas_height = 0;
        // interleave the material's maps when they can share one set of texel coordinates
        uint32_t n_present = 0, w = 0, h = 0; bool same = true;
        for (int k = 0; k < 4; ++k) {
            if (!d.tex[k].texels) continue;
            d.present |= 1u << k;
            if (n_present++ == 0) { w = d.tex[k].width; h = d.tex[k].height; }
            else same = same && d.tex[k].width == w && d.tex[k].height == h;
        }
        if (n_present >= 2 && same && !(r->cfg.flags & PT_FLAG_NO_TEXTURE_INTERLEAVE)) {
            const std::array<uint32_t, 4> key = {t[0], t[1], t[2], t[3]};
            auto it = atlas_of.find(key);
            if (it == atlas_of.end()) {
                uint4* a = nullptr;
                if ((rc = dev_alloc(&a, (size_t)w * h)) != PT_OK) return rc;
                r->d_atlases.push_back(a);
                launch_interleave_textures(d.tex[0].texels, d.tex[1].texels, d.tex[2].texels, d.tex[3].texels, (size_t)w * h, a, r->stream);
                it = atlas_of.emplace(key, a).first;
            }
            d.atlas = it->second; d.atlas_width = w; d.atlas_height = h;
        }
    }
    PT_CUDA_CHECK(cudaStreamSynchronize(r->stream));
    r->scene.atlas_only = 1u;
    for (const MatRec& m : mats) if (m.present != 0u && !m.atlas) r->scene.atlas_only = 0u;
#ifdef PT_SHADE_NO_ATLAS_ONLY
    r->scene.atlas_only = 0u;       // A/B variant: always the general K-shade instantiation
#endif
    if (r->cfg.flags & PT_FLAG_HW_BILINEAR) {
        // the reference's sampler on the texture units: r8g8b8a8_unorm, linear mag/min, REPEAT, one mip (RayTracingPipeline.zig:54-70)
        r->tex_arrays.assign(s->num_textures, nullptr);
        r->tex_objects.assign(s->num_textures, 0);
        for (size_t i = 0; i < s->num_textures; ++i) {
            const PtTexture& t = s->textures[i];
            const cudaChannelFormatDesc fmt = cudaCreateChannelDesc<uchar4>();
            PT_CUDA_CHECK(cudaMallocArray(&r->tex_arrays[i], &fmt, t.width, t.height));
            PT_CUDA_CHECK(cudaMemcpy2DToArrayAsync(r->tex_arrays[i], 0, 0, t.rgba, (size_t)t.width * 4, (size_t)t.width * 4, t.height, cudaMemcpyHostToDevice, r->stream));
            cudaResourceDesc rd; std::memset(&rd, 0, sizeof(rd));
            rd.resType = cudaResourceTypeArray; rd.res.array.array = r->tex_arrays[i];
            cudaTextureDesc td; std::memset(&td, 0, sizeof(td));
            td.addressMode[0] = td.addressMode[1] = cudaAddressModeWrap;
            td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeNormalizedFloat; td.normalizedCoords = 1;
            PT_CUDA_CHECK(cudaCreateTextureObject(&r->tex_objects[i], &rd, &td, nullptr));
        }
        std::vector<unsigned long long> objs(4 * std::max<size_t>(s->num_materials, 1), 0ull);
        for (size_t i = 0; i < s->num_materials; ++i) {
            const PtMaterial& m = s->materials[i];
            const uint32_t t[4] = {m.albedo_texture_index, m.metal_roughness_texture_index, m.emissive_texture_index, m.normal_texture_index};
            for (int k = 0; k < 4; ++k) objs[4 * i + k] = t[k] == PT_INVALID_TEXTURE ? 0ull : (unsigned long long)r->tex_objects[t[k]];
        }
        if ((rc = dev_alloc(&r->d_mat_tex_objs, objs.size())) != PT_OK) return rc;
        PT_CUDA_CHECK(cudaMemcpyAsync(r->d_mat_tex_objs, objs.data(), objs.size() * 8, cudaMemcpyHostToDevice, r->stream));
        PT_CUDA_CHECK(cudaStreamSynchronize(r->stream));
    }
    if ((rc = dev_alloc(&r->d_materials, mats.size())) != PT_OK) return rc;
    if (!mats.empty()) PT_CUDA_CHECK(cudaMemcpyAsync(r->d_materials, mats.data(), mats.size() * sizeof(MatRec), cudaMemcpyHostToDevice, r->stream));
    // meshes / instances
    std::vector<PtMesh> meshes(s->meshes, s->meshes + s->num_meshes);
    for (size_t m = 0; m < meshes.size(); ++m)
        if (meshes[m].start > meshes[m].end || meshes[m].end > s->num_primitives) { set_error("mesh " + std::to_string(m) + ": primitive range out of bounds"); return PT_ERR_INVALID_ARGUMENT; }
    std::vector<PtInstance> inst(s->instances, s->instances + s->num_instances);
    for (size_t i = 0; i < inst.size(); ++i)
        if (inst[i].mesh_index >= meshes.size()) { set_error("instance " + std::to_string(i) + ": mesh index out of range"); return PT_ERR_INVALID_ARGUMENT; }
    if ((rc = dev_alloc(&r->d_instances_raw, inst.size())) != PT_OK) return rc;
    if (!inst.empty()) PT_CUDA_CHECK(cudaMemcpyAsync(r->d_instances_raw, inst.data(), inst.size() * sizeof(PtInstance), cudaMemcpyHostToDevice, r->stream));
    // acceleration structures (createBlases + Tlas.init)
    // builder stages: binned SAH + cost-optimal collapse unless the caller picks (measured on B200, profiles/r2a_builders.txt:
    // C2 trace 1.328 -> 1.202 ms, C5 52.8 -> 46.4 ms, C4 14.97 -> 14.41 ms against the radix tree + greedy collapse; the build
    // is 3-4x slower — 24 ms for 259 k triangles — and, like the reference's prefer_fast_trace build, off the frame path)
    uint32_t builder = r->cfg.flags & (PT_FLAG_BVH_PLOC | PT_FLAG_BVH_SAH | PT_FLAG_BVH_OPTIMAL_COLLAPSE | PT_FLAG_BVH_LBVH);
    if (builder == 0u) builder = PT_FLAG_BVH_SAH | PT_FLAG_BVH_OPTIMAL_COLLAPSE;
    builder &= ~(uint32_t)PT_FLAG_BVH_LBVH;
    if ((rc = build_scene_bvh(r->d_blob, r->d_prims, prims, meshes, r->d_instances_raw, inst, builder, r->stream, &r->bvh)) != PT_OK) return rc;
    r->h_instances = inst;
    if ((rc = dev_alloc(&r->d_instances, inst.size())) != PT_OK) return rc;
    if ((rc = upload_instance_records(r)) != PT_OK) return rc;
    DeviceScene& sc = r->scene;
    sc.triangle_data = r->d_blob; sc.prims = r->d_prims; sc.materials = r->d_materials; sc.textures = r->d_textures;
    sc.instances = r->d_instances; sc.nodes = r->bvh.nodes8; sc.tris = r->bvh.tris8; sc.tlas_leaf_instance = r->bvh.tlas_leaf_instance;
    sc.nodes2 = r->bvh.nodes2; sc.tris2 = r->bvh.tris2; sc.tlas2 = r->bvh.tlas2; sc.tlas_root = r->bvh.tlas_root8;
    PT_CUDA_CHECK(cudaHostGetDevicePointer(reinterpret_cast<void**>(&sc.error_flag), r->h_error, 0));
    sc.mat_tex_objs = r->d_mat_tex_objs;
    sc.num_instances = (uint32_t)inst.size(); sc.num_prims = (uint32_t)prims.size();
    sc.num_materials = (uint32_t)s->num_materials; sc.num_textures = (uint32_t)s->num_textures;
    return PT_OK;
}

void destroy_impl(PtRenderer* r) {
    if (!r) return;
    cudaSetDevice(r->device);
    if (r->stream) cudaStreamSynchronize(r->stream);
    free_frame(r);
    for (auto e : r->events) cudaEventDestroy(e);
    if (r->owns_scene) {
        for (auto p : r->d_texels) cudaFree(p);
        for (auto p : r->d_atlases) cudaFree(p);
        for (auto t : r->tex_objects) if (t) cudaDestroyTextureObject(t);
        for (auto a : r->tex_arrays) if (a) cudaFreeArray(a);
        cudaFree(r->d_mat_tex_objs);
        cudaFree(r->d_blob); cudaFree(r->d_prims); cudaFree(r->d_materials); cudaFree(r->d_textures);
        cudaFree(r->d_instances_raw); cudaFree(r->d_instances);
        r->bvh.release();
    }
    cudaFree(r->d_cursor);
    if (r->owns_scene && r->h_error) cudaFreeHost(r->h_error);
    comm_release(r);
    if (r->ev_reduce_done) cudaEventDestroy(r->ev_reduce_done);
    if (r->d_remote_image) cudaIpcCloseMemHandle(r->d_remote_image);
    if (r->copy_stream) { cudaStreamSynchronize(r->copy_stream); cudaStreamDestroy(r->copy_stream); }
    if (r->ev_frame_done) cudaEventDestroy(r->ev_frame_done);
    if (r->ev_copy_done) cudaEventDestroy(r->ev_copy_done);
    if (r->stream) cudaStreamDestroy(r->stream);
    delete r;
}

}  // namespace

// stream for follow-up work on the image: the caller's, else the stream of the last pt_render
static inline cudaStream_t follow_stream(PtRenderer* r, void* user) {
    if (user) return static_cast<cudaStream_t>(user);
    return r->last_stream ? r->last_stream : r->stream;
}

#define PT_REQUIRE(cond, msg) do { if (!(cond)) { set_error(msg); return PT_ERR_INVALID_ARGUMENT; } } while (0)

extern "C" {

const char* pt_last_error(void) { return g_last_error.c_str(); }

const char* pt_version(void) { return "b200pt 0.1 (sm_100a)"; }

int pt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

const char* pt_status_name(int s) {
    switch (s) {
        case PT_OK: return "Ok";
        case PT_ERR_INVALID_ARGUMENT: return "InvalidArgument";
        case PT_ERR_CUDA: return "CudaError";
        case PT_ERR_OUT_OF_MEMORY: return "OutOfMemory";
        case PT_ERR_IO: return "IoError";
        case PT_ERR_INVALID_FILE_EXTENSION: return "InvalidFileExtension";
        case PT_ERR_GLTF_PARSE: return "GltfParseError";
        case PT_ERR_GLTF_NOT_TRIANGLES: return "GltfNotATriangleTopology";
        case PT_ERR_GLTF_NO_INDICES: return "GltfNoIndices";
        case PT_ERR_GLTF_NO_POSITIONS: return "GltfNoPositions";
        case PT_ERR_GLTF_NO_NORMALS: return "GltfNoNormals";
        case PT_ERR_GLTF_NO_TANGENTS: return "GltfNoTangents";
        case PT_ERR_GLTF_NO_TEXCOORDS: return "GltfNoTextureCoordinates";
        case PT_ERR_NO_MATERIAL: return "NoMaterial";
        case PT_ERR_NO_BUFFER_VIEW: return "NoBufferView";
        case PT_ERR_INVALID_ATTRIBUTE_TYPE: return "InvalidAttributeType";
        case PT_ERR_NO_SCENE: return "NoScene";
        case PT_ERR_TOO_MANY_SCENES: return "TooManyScenes";
        case PT_ERR_IMAGE_DECODE: return "ImageDecodeError";
        case PT_ERR_INVALID_INDEX_TYPE: return "InvalidIndexType";
        case PT_ERR_NCCL: return "NcclError";
        case PT_ERR_TRAVERSAL_STACK: return "TraversalStackOverflow";
        default: return "Unknown";
    }
}

int pt_create(const PtSceneDesc* scene, const PtRenderConfig* config, PtRenderer** out) {
    PT_REQUIRE(scene && config && out, "pt_create: null argument");
    *out = nullptr;
    PT_REQUIRE(config->width > 0 && config->height > 0, "pt_create: zero resolution");
    PT_REQUIRE(config->num_samples < 65536 && config->num_bounces < 65536, "pt_create: num_samples / num_bounces must be < 65536");
    PT_REQUIRE(config->partition == PT_PARTITION_NONE || (config->world >= 1 && config->rank >= 0 && config->rank < config->world),
               "pt_create: invalid rank/world");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("pt_create: no CUDA device available (this library has no CPU fallback)");
        return PT_ERR_CUDA;
    }
    PT_REQUIRE(config->device >= 0 && config->device < ndev, "pt_create: device ordinal out of range");
    PtRenderer* r = new PtRenderer();
    r->cfg = *config;
    if (r->cfg.divider == 0) r->cfg.divider = 1;
    r->device = config->device;
    int rc = PT_OK;
    do {
        if (cudaSetDevice(r->device) != cudaSuccess) { set_error("cudaSetDevice failed"); rc = PT_ERR_CUDA; break; }
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, r->device) != cudaSuccess) { set_error("cudaGetDeviceProperties failed"); rc = PT_ERR_CUDA; break; }
        r->num_sms = prop.multiProcessorCount;
        r->dims = compute_launch_dims(r->num_sms);
        if (cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking) != cudaSuccess) { set_error("cudaStreamCreate failed"); rc = PT_ERR_CUDA; break; }
        if (cudaHostAlloc(reinterpret_cast<void**>(&r->h_error), 4, cudaHostAllocMapped) != cudaSuccess) { set_error("cudaHostAlloc failed"); rc = PT_ERR_CUDA; break; }
        *r->h_error = 0u;
        if ((rc = upload_scene(r, scene)) != PT_OK) break;
        if ((rc = alloc_frame(r, r->cfg.width, r->cfg.height, r->cfg.divider)) != PT_OK) break;
        if ((rc = dev_alloc(&r->d_cursor, 4)) != PT_OK) break;
        if (r->cfg.flags & PT_FLAG_TRACE_BVH2) r->trace_mode = TRACE_BVH2;
        else if (r->cfg.flags & PT_FLAG_TRACE_WALKER) r->trace_mode = TRACE_BVH8_WALKER;
    } while (0);
    if (rc != PT_OK) { std::string keep = g_last_error; destroy_impl(r); g_last_error = keep; return rc; }
    *out = r;
    return PT_OK;
}

// RaytracingPass.init(..., scene_path, ...) (RaytracingPass.zig:87-103): load, upload, drop the host scene.
int pt_create_from_gltf(const char* scene_path, const PtRenderConfig* config, PtRenderer** out) {
    PT_REQUIRE(scene_path && config && out, "pt_create_from_gltf: null argument");
    *out = nullptr;
    PtScene* scene = nullptr;
    int rc = pt_scene_load(scene_path, &scene);                 // Scene.load, :102
    if (rc != PT_OK) return rc;
    PtSceneDesc desc;
    rc = pt_scene_desc(scene, &desc);
    if (rc == PT_OK) rc = pt_create(&desc, config, out);
    const std::string keep = g_last_error;
    pt_scene_free(scene);                                       // defer scene.deinit(), :103
    if (rc != PT_OK) g_last_error = keep;
    return rc;
}

// A second frame context over the same device scene: its own stream, path state, image and read-back pipeline,
// sharing scene data and acceleration structures with `parent` (which must outlive it).  Alternating pt_render
// between a renderer and its sibling keeps two frames in flight -- the reference does the same with one command
// buffer and fence per swapchain image (main.zig:209-217,255-259) -- so the tail of one frame's last wavefront
// overlaps the next frame's ray generation and primary traversal.
static int create_sibling_impl(PtRenderer* parent, PtRenderer** out);

int pt_create_sibling(PtRenderer* parent, PtRenderer** out) {
    PT_REQUIRE(parent && out, "pt_create_sibling: null argument");
    return create_sibling_impl(parent, out);
}

static int create_sibling_impl(PtRenderer* parent, PtRenderer** out) {
    *out = nullptr;
    PT_REQUIRE(parent->owns_scene, "pt_create_sibling: the parent must be a renderer made by pt_create");
    PT_CUDA_CHECK(cudaSetDevice(parent->device));
    PtRenderer* r = new PtRenderer();
    r->cfg = parent->cfg;
    r->device = parent->device; r->num_sms = parent->num_sms; r->dims = parent->dims; r->h_error = parent->h_error; r->owns_scene = false; r->parent = parent;
    r->d_blob = parent->d_blob; r->d_prims = parent->d_prims; r->d_materials = parent->d_materials; r->d_textures = parent->d_textures;
    r->d_instances_raw = parent->d_instances_raw; r->d_instances = parent->d_instances;
    r->bvh = parent->bvh; r->scene = parent->scene; r->num_triangles = parent->num_triangles; r->trace_mode = parent->trace_mode;
    int rc = PT_OK;
    do {
        if (cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking) != cudaSuccess) { set_error("cudaStreamCreate failed"); rc = PT_ERR_CUDA; break; }
        if ((rc = alloc_frame(r, parent->full_w, parent->full_h, r->cfg.divider)) != PT_OK) break;
        if ((rc = dev_alloc(&r->d_cursor, 4)) != PT_OK) break;
    } while (0);
    if (rc != PT_OK) { std::string keep = g_last_error; destroy_impl(r); g_last_error = keep; return rc; }
    *out = r;
    return PT_OK;
}

void pt_destroy(PtRenderer* r) { destroy_impl(r); }

static int render_impl(PtRenderer* r, const PtCamera* camera, cudaStream_t st, TraceMode mode) {
    if (r->parent) { r->scene = r->parent->scene; r->bvh.tlas2_count = r->parent->bvh.tlas2_count; }   // follow pt_set_instance_transforms on the parent
    PathBuffers& pb = r->pb;
    FrameParams fp;
    fp.cam = *camera;
    fp.num_samples = r->cfg.num_samples; fp.num_bounces = r->cfg.num_bounces; fp.flags = r->cfg.flags;
    // records of the primary rays that nobody reads (render.h): origins when the flat instantiation of the persistent K-trace
    // runs, directions too when S == 1 (K-shade then restates the primary ray from the pixel).  C2: K-raygen 0.062 -> 0.056 ms
    fp.primary_lean = (mode != TRACE_BVH8 || r->scene.start_inst == kInvalidId) ? 0u : (1u | (fp.num_samples == 1u ? 2u : 0u));
    fp.image = r->d_remote_image ? r->d_remote_image : r->d_image;
    const size_t px = (size_t)r->pm.width * r->pm.height;
    fp.id_inst = r->d_ids; fp.id_prim = r->d_ids + px; fp.id_tri = r->d_ids + 2 * px;
    PT_CUDA_CHECK(cudaMemsetAsync(pb.qcount, 0, sizeof(uint32_t) * pb.frame_words, st));   // counts + both cursor arrays + hit counter
    bool waited_copy = !r->copy_pending && !r->reduce_pending;
    auto wait_copy = [&]() {   // a pipelined read-back of the previous frame must finish before the image is written again
        if (!waited_copy) { if (r->copy_pending) cudaStreamWaitEvent(st, r->ev_copy_done, 0); waited_copy = true; r->copy_pending = false; }
        // ... and so must a reduce of its image / accumulation buffer on the communication stream (comm.cpp)
        if (r->reduce_pending) { cudaStreamWaitEvent(st, r->ev_reduce_done, 0); r->reduce_pending = false; }
    };
    if (r->pm.world > 1 && !r->d_remote_image && !r->peer_root) wait_copy();
    // tile partition: non-owned pixels must read as 0 for the sum-reduce; not when peers store into this image
    if (r->pm.world > 1 && !r->d_remote_image && !r->peer_root) PT_CUDA_CHECK(cudaMemsetAsync(r->d_image, 0, px * 16, st));
    size_t ev = 0;
    auto mark = [&](int kind) {
        if (!r->profiling) return;
        if (ev + 2 > r->events.size()) {
            cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
            r->events.push_back(a); r->events.push_back(b); r->event_kind.push_back(kind);
        } else {
            r->event_kind[ev / 2] = kind;
        }
        cudaEventRecord(r->events[ev], st);
    };
    auto mark_end = [&]() { if (r->profiling) { cudaEventRecord(r->events[ev + 1], st); ev += 2; } };
    uint32_t launches = 0;
    if (r->max_iters == 0) wait_copy();   // S == 0 or B == 0: K-raygen itself writes the (black) image
    mark(2); launch_raygen(r->pm, pb, fp, st); mark_end(); ++launches;
    for (uint32_t it = 0; it < r->max_iters; ++it) {
        mark(0); launch_trace(r->scene, pb, it, mode, (it == 0 && (fp.primary_lean & 1u)) ? fp.cam.position : nullptr, r->bvh.tlas2_count, r->bvh.tlas2_leaf_instance, r->dims, st); mark_end();
        if (fp.flags & PT_FLAG_SORT_BY_MATERIAL) { mark(2); launch_material_sort(r->scene, pb, it, r->num_sms, st); mark_end(); launches += 3; }
        wait_copy();
        mark(1); launch_shade(r->scene, r->pm, pb, fp, it, r->dims, st); mark_end();
        launches += 2;
    }
    if (r->accumulate) {
        if (!r->d_accum) {
            int rc = dev_alloc(&r->d_accum, px * 4);
            if (rc != PT_OK) return rc;
            PT_CUDA_CHECK(cudaMemsetAsync(r->d_accum, 0, px * 16, st));
            r->accum_frames = 0;
        }
        mark(2); launch_accumulate(r->d_image, r->d_accum, (uint32_t)(px * 4), st); mark_end(); ++launches;
        r->accum_frames++;
    }
    PT_CUDA_CHECK(cudaGetLastError());
    r->last_iters = r->max_iters;
    r->last_launches = launches;
    r->last_cam = *camera;
    r->have_last_cam = true;
    r->last_stream = st;
    return PT_OK;
}

int pt_render(PtRenderer* r, const PtCamera* camera, void* cuda_stream) {
    PT_REQUIRE(r && camera, "pt_render: null argument");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : r->stream;
    return render_impl(r, camera, st, r->trace_mode);
}

int pt_sync(PtRenderer* r) {
    PT_REQUIRE(r, "pt_sync: null renderer");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    if (r->last_stream && r->last_stream != r->stream) PT_CUDA_CHECK(cudaStreamSynchronize(r->last_stream));
    PT_CUDA_CHECK(cudaStreamSynchronize(r->stream));
    if (r->copy_stream) PT_CUDA_CHECK(cudaStreamSynchronize(r->copy_stream));
    if (r->h_error && *reinterpret_cast<volatile uint32_t*>(r->h_error)) {
        set_error("a ray needed more than the 64 traversal stack entries of the kernels: the acceleration structure is too deep");
        return PT_ERR_TRAVERSAL_STACK;
    }
    return PT_OK;
}

int pt_resize(PtRenderer* r, uint32_t width, uint32_t height) {
    PT_REQUIRE(r && width > 0 && height > 0, "pt_resize: invalid argument");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    { int rc = pt_sync(r); if (rc != PT_OK) return rc; }
    // Like the reference, the storage image is re-created at the full new extent: the divider is
    // dropped (RaytracingPass.zig:684-697, quirk Q13).
    r->cfg.width = width; r->cfg.height = height; r->cfg.divider = 1;
    return alloc_frame(r, width, height, 1);
}

int pt_render_extent(const PtRenderer* r, uint32_t* width, uint32_t* height) {
    PT_REQUIRE(r, "pt_render_extent: null renderer");
    if (width) *width = r->pm.width;
    if (height) *height = r->pm.height;
    return PT_OK;
}

int pt_device_image(PtRenderer* r, float** device_ptr) {
    PT_REQUIRE(r && device_ptr, "pt_device_image: null argument");
    *device_ptr = r->d_image;
    return PT_OK;
}

int pt_read_rgba32f_async(PtRenderer* r, float* dst, void* cuda_stream) {
    PT_REQUIRE(r && dst, "pt_read_rgba32f_async: null argument");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    cudaStream_t st = follow_stream(r, cuda_stream);
    PT_CUDA_CHECK(cudaMemcpyAsync(dst, r->d_image, (size_t)r->pm.width * r->pm.height * 16, cudaMemcpyDeviceToHost, st));
    return PT_OK;
}

// Pipelined read-back: the device->host copy of the current image is ordered after everything
// enqueued so far on the render stream but runs on the renderer's copy stream, so it overlaps the
// next frame's ray generation and primary traversal (which do not touch the image).
static int overlapped_copy(PtRenderer* r, void* dst, const void* src, size_t bytes, cudaStream_t after) {
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    if (!r->copy_stream) {
        PT_CUDA_CHECK(cudaStreamCreateWithFlags(&r->copy_stream, cudaStreamNonBlocking));
        PT_CUDA_CHECK(cudaEventCreateWithFlags(&r->ev_frame_done, cudaEventDisableTiming));
        PT_CUDA_CHECK(cudaEventCreateWithFlags(&r->ev_copy_done, cudaEventDisableTiming));
    }
    PT_CUDA_CHECK(cudaEventRecord(r->ev_frame_done, after));
    PT_CUDA_CHECK(cudaStreamWaitEvent(r->copy_stream, r->ev_frame_done, 0));
    PT_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, r->copy_stream));
    PT_CUDA_CHECK(cudaEventRecord(r->ev_copy_done, r->copy_stream));
    r->copy_pending = true;
    return PT_OK;
}

int pt_read_rgba32f_overlapped(PtRenderer* r, float* pinned_host_dst) {
    PT_REQUIRE(r && pinned_host_dst, "pt_read_rgba32f_overlapped: null argument");
    return overlapped_copy(r, pinned_host_dst, r->d_image, (size_t)r->pm.width * r->pm.height * 16, follow_stream(r, nullptr));
}

int pt_read_rgba32f(PtRenderer* r, float* host_dst) {
    int rc = pt_read_rgba32f_async(r, host_dst, nullptr);
    if (rc != PT_OK) return rc;
    return pt_sync(r);
}

int pt_read_bgra8_async(PtRenderer* r, uint8_t* dst, void* cuda_stream) {
    PT_REQUIRE(r && dst, "pt_read_bgra8_async: null argument");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    cudaStream_t st = follow_stream(r, cuda_stream);
    const size_t px = (size_t)r->pm.width * r->pm.height;
    if (!r->d_bgra) { int rc = dev_alloc(&r->d_bgra, px * 4); if (rc != PT_OK) return rc; }
    launch_to_bgra8(r->d_image, r->d_bgra, (uint32_t)px, st);
    PT_CUDA_CHECK(cudaMemcpyAsync(dst, r->d_bgra, px * 4, cudaMemcpyDeviceToHost, st));
    return PT_OK;
}

int pt_read_bgra8(PtRenderer* r, uint8_t* host_dst) {
    int rc = pt_read_bgra8_async(r, host_dst, nullptr);
    if (rc != PT_OK) return rc;
    return pt_sync(r);
}

int pt_read_ids(PtRenderer* r, uint32_t* instance, uint32_t* primitive, uint32_t* triangle) {
    PT_REQUIRE(r, "pt_read_ids: null renderer");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    const size_t px = (size_t)r->pm.width * r->pm.height;
    { int rc = pt_sync(r); if (rc != PT_OK) return rc; }
    if (instance) PT_CUDA_CHECK(cudaMemcpy(instance, r->d_ids, px * 4, cudaMemcpyDeviceToHost));
    if (primitive) PT_CUDA_CHECK(cudaMemcpy(primitive, r->d_ids + px, px * 4, cudaMemcpyDeviceToHost));
    if (triangle) PT_CUDA_CHECK(cudaMemcpy(triangle, r->d_ids + 2 * px, px * 4, cudaMemcpyDeviceToHost));
    return PT_OK;
}

int pt_read_rgba32f_upscaled(PtRenderer* r, float* host_dst) {
    PT_REQUIRE(r && host_dst, "pt_read_rgba32f_upscaled: null argument");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    if (r->pm.width == r->full_w && r->pm.height == r->full_h) return pt_read_rgba32f(r, host_dst);
    const size_t px = (size_t)r->full_w * r->full_h;
    if (!r->d_upscaled) { int rc = dev_alloc(&r->d_upscaled, px * 4); if (rc != PT_OK) return rc; }
    cudaStream_t st = follow_stream(r, nullptr);
    launch_upscale(r->d_image, r->pm.width, r->pm.height, r->d_upscaled, r->full_w, r->full_h, st);
    PT_CUDA_CHECK(cudaMemcpyAsync(host_dst, r->d_upscaled, px * 16, cudaMemcpyDeviceToHost, st));
    return pt_sync(r);
}

int pt_set_accumulate(PtRenderer* r, int enabled) {
    PT_REQUIRE(r, "pt_set_accumulate: null renderer");
    r->accumulate = enabled != 0;
    return PT_OK;
}

int pt_reset_accum(PtRenderer* r) {
    PT_REQUIRE(r, "pt_reset_accum: null renderer");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    if (r->d_accum) PT_CUDA_CHECK(cudaMemsetAsync(r->d_accum, 0, (size_t)r->pm.width * r->pm.height * 16, follow_stream(r, nullptr)));
    r->accum_frames = 0;
    return PT_OK;
}

int pt_device_accum(PtRenderer* r, float** device_ptr, uint32_t* frames) {
    PT_REQUIRE(r && device_ptr, "pt_device_accum: null argument");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    if (!r->d_accum) {
        const size_t px = (size_t)r->pm.width * r->pm.height;
        int rc = dev_alloc(&r->d_accum, px * 4);
        if (rc != PT_OK) return rc;
        PT_CUDA_CHECK(cudaMemsetAsync(r->d_accum, 0, px * 16, r->stream));
    }
    *device_ptr = r->d_accum;
    if (frames) *frames = r->accum_frames;
    return PT_OK;
}

int pt_read_accum_mean(PtRenderer* r, float* host_dst, uint32_t total_frames) {
    PT_REQUIRE(r && host_dst && r->d_accum, "pt_read_accum_mean: accumulation is not enabled");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    const size_t px = (size_t)r->pm.width * r->pm.height;
    const uint32_t frames = total_frames ? total_frames : r->accum_frames;
    PT_REQUIRE(frames > 0, "pt_read_accum_mean: no frames accumulated");
    if (!r->d_upscaled || r->full_w * r->full_h < px) { cudaFree(r->d_upscaled); r->d_upscaled = nullptr; int rc = dev_alloc(&r->d_upscaled, std::max(px, (size_t)r->full_w * r->full_h) * 4); if (rc != PT_OK) return rc; }
    cudaStream_t st = follow_stream(r, nullptr);
    launch_scale(r->d_accum, r->d_upscaled, (uint32_t)(px * 4), 1.0f / (float)frames, st);
    PT_CUDA_CHECK(cudaMemcpyAsync(host_dst, r->d_upscaled, px * 16, cudaMemcpyDeviceToHost, st));
    return pt_sync(r);
}

int pt_add_accum(PtRenderer* dst, PtRenderer* src) {
    PT_REQUIRE(dst && src && dst != src, "pt_add_accum: invalid argument");
    PT_REQUIRE(dst->device == src->device && dst->pm.width == src->pm.width && dst->pm.height == src->pm.height, "pt_add_accum: device or extent differ");
    PT_REQUIRE(src->d_accum, "pt_add_accum: the source has no accumulation buffer");
    PT_CUDA_CHECK(cudaSetDevice(dst->device));
    const size_t px = (size_t)dst->pm.width * dst->pm.height;
    cudaStream_t st = follow_stream(dst, nullptr);
    if (!dst->d_accum) {
        int rc = dev_alloc(&dst->d_accum, px * 4);
        if (rc != PT_OK) return rc;
        PT_CUDA_CHECK(cudaMemsetAsync(dst->d_accum, 0, px * 16, st));
        dst->accum_frames = 0;
    }
    // ordered after the source's frames: an event on its stream, waited for on the destination's
    cudaEvent_t ev;
    PT_CUDA_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    PT_CUDA_CHECK(cudaEventRecord(ev, follow_stream(src, nullptr)));
    PT_CUDA_CHECK(cudaStreamWaitEvent(st, ev, 0));
    PT_CUDA_CHECK(cudaEventDestroy(ev));
    launch_accumulate(src->d_accum, dst->d_accum, (uint32_t)(px * 4), st);
    dst->accum_frames += src->accum_frames;
    return PT_OK;
}

int pt_set_profiling(PtRenderer* r, int enabled) {
    PT_REQUIRE(r, "pt_set_profiling: null renderer");
    r->profiling = enabled != 0;
    return PT_OK;
}

int pt_get_stats(PtRenderer* r, PtStats* out) {
    PT_REQUIRE(r && out, "pt_get_stats: null argument");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    std::memset(out, 0, sizeof(*out));
    { int rc = pt_sync(r); if (rc != PT_OK) return rc; }
    std::vector<uint32_t> q((size_t)r->max_iters + 2, 0);
    PT_CUDA_CHECK(cudaMemcpy(q.data(), r->pb.qcount, q.size() * 4, cudaMemcpyDeviceToHost));
    unsigned long long hits = 0;
    PT_CUDA_CHECK(cudaMemcpy(&hits, r->pb.hit_count, 8, cudaMemcpyDeviceToHost));
    out->rays_hit = hits;
    uint32_t its = 0;
    for (uint32_t i = 0; i < r->last_iters; ++i) { out->rays_traced += q[i]; if (q[i]) its = i + 1; }
    out->wavefront_iterations = its;
    out->kernel_launches = r->last_launches;
    if (r->profiling) {
        for (size_t e = 0; e + 1 < r->events.size() && e / 2 < r->event_kind.size(); e += 2) {
            float ms = 0.0f;
            if (cudaEventElapsedTime(&ms, r->events[e], r->events[e + 1]) != cudaSuccess) { cudaGetLastError(); continue; }
            const int k = r->event_kind[e / 2];
            if (k == 0) out->ms_trace += ms; else if (k == 1) out->ms_shade += ms; else out->ms_other += ms;
        }
        out->ms_total = out->ms_trace + out->ms_shade + out->ms_other;
    }
    out->bvh_nodes = r->bvh.num_nodes8;
    out->bvh_bytes = (uint64_t)r->bvh.num_nodes8 * sizeof(Bvh8Node) + (uint64_t)r->bvh.num_tris * sizeof(Bvh8Tri);
    out->num_triangles = r->num_triangles;
    out->ms_build = r->bvh.ms_build;
    return PT_OK;
}

int pt_trace_rays(PtRenderer* r, const PtRay* rays, size_t n, PtHit* hits) {
    PT_REQUIRE(r && (n == 0 || (rays && hits)), "pt_trace_rays: null argument");
    PT_REQUIRE(n < 0x80000000ull, "pt_trace_rays: too many rays (K-trace uses bit 31 of the slot index)");
    if (n == 0) return PT_OK;
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    if (r->parent) r->scene = r->parent->scene;
    RayRec* d_rays = nullptr; HitRec* d_hits = nullptr; float4* d_setup = nullptr;
    int rc;
    if ((rc = dev_alloc(&d_rays, n)) != PT_OK) return rc;
    if ((rc = dev_alloc(&d_hits, n)) != PT_OK) { cudaFree(d_rays); return rc; }
    if ((rc = dev_alloc(&d_setup, 5 * n)) != PT_OK) { cudaFree(d_rays); cudaFree(d_hits); return rc; }
    static_assert(sizeof(PtRay) == sizeof(RayRec), "PtRay layout");
    cudaError_t e = cudaMemcpyAsync(d_rays, rays, n * sizeof(RayRec), cudaMemcpyHostToDevice, r->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(r->d_cursor, 0, 4, r->stream);
    if (e == cudaSuccess) {
        launch_trace_rays(r->scene, d_rays, (uint32_t)n, d_hits, r->trace_mode, r->bvh.tlas2_count,
                          r->bvh.tlas2_leaf_instance, r->d_cursor, d_setup, r->num_sms, r->dims, r->stream);
        e = cudaGetLastError();
    }
    std::vector<HitRec> tmp(n);
    if (e == cudaSuccess) e = cudaMemcpyAsync(tmp.data(), d_hits, n * sizeof(HitRec), cudaMemcpyDeviceToHost, r->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(r->stream);
    cudaFree(d_rays); cudaFree(d_hits); cudaFree(d_setup);
    if (e != cudaSuccess) { set_error(std::string("pt_trace_rays: ") + cudaGetErrorString(e)); return PT_ERR_CUDA; }
    for (size_t i = 0; i < n; ++i) {
        hits[i].t = tmp[i].t; hits[i].b1 = tmp[i].b1; hits[i].b2 = tmp[i].b2; hits[i].instance = tmp[i].instance;
        hits[i].primitive_index = tmp[i].primitive_index; hits[i].triangle_index = tmp[i].triangle_index;
    }
    return PT_OK;
}

int pt_count_traversal(PtRenderer* r, uint64_t* nodes, uint64_t* tris, uint64_t* rays) {
    PT_REQUIRE(r && r->have_last_cam, "pt_count_traversal: call pt_render first");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    { int rc = pt_sync(r); if (rc != PT_OK) return rc; }
    PT_CUDA_CHECK(cudaMemsetAsync(r->pb.trav_counters, 0, 16, r->stream));
    const bool prof = r->profiling, acc = r->accumulate;
    r->profiling = false; r->accumulate = false;
    const PtCamera cam = r->last_cam;
    int rc = render_impl(r, &cam, r->stream, TRACE_BVH8_COUNT);
    r->profiling = prof; r->accumulate = acc;
    if (rc != PT_OK) return rc;
    unsigned long long c[2] = {0, 0};
    PT_CUDA_CHECK(cudaMemcpyAsync(c, r->pb.trav_counters, 16, cudaMemcpyDeviceToHost, r->stream));
    PtStats st;
    rc = pt_get_stats(r, &st);
    if (rc != PT_OK) return rc;
    if (nodes) *nodes = c[0];
    if (tris) *tris = c[1];
    if (rays) *rays = st.rays_traced;
    return PT_OK;
}

// Digest of the acceleration structures, for determinism probes (tools/gpu_determinism.py):
//   [0] Morton-ordered triangles of every mesh   [1] BVH2 topology (left/right)   [2] BVH2 boxes
//   [3] number of BVH8 nodes   [4] BVH8 nodes as an order-independent sum of per-node hashes of the geometry
//   fields (frame, exponents, quantised planes, metas' top bits)   [5] per-mesh BVH8 roots
int pt_debug_bvh_digest(PtRenderer* r, uint64_t digest[6]) {
    PT_REQUIRE(r && digest, "pt_debug_bvh_digest: null argument");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    { int rc = pt_sync(r); if (rc != PT_OK) return rc; }
    auto fnv = [](uint64_t h, const void* p, size_t n) { const uint8_t* b = static_cast<const uint8_t*>(p); for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; } return h; };
    const uint64_t seed = 1469598103934665603ull;
    for (int k = 0; k < 6; ++k) digest[k] = seed;
    const SceneBvh& b = r->bvh;
    std::vector<Bvh8Tri> tris(std::max<uint32_t>(b.num_tris, 1));
    std::vector<Bvh2Node> n2(std::max<uint32_t>(b.num_tris, 1));
    if (b.num_tris) {
        PT_CUDA_CHECK(cudaMemcpy(tris.data(), b.tris2, sizeof(Bvh8Tri) * b.num_tris, cudaMemcpyDeviceToHost));
        PT_CUDA_CHECK(cudaMemcpy(n2.data(), b.nodes2, sizeof(Bvh2Node) * b.num_tris, cudaMemcpyDeviceToHost));
    }
    for (size_t m = 0; m < b.tri_count.size(); ++m) {
        const uint32_t n = b.tri_count[m], off = b.tri_off[m];
        for (uint32_t i = 0; i < n; ++i) {
            digest[0] = fnv(digest[0], &tris[off + i].v0, 12); digest[0] = fnv(digest[0], &tris[off + i].v1, 12); digest[0] = fnv(digest[0], &tris[off + i].v2, 12);
            digest[0] = fnv(digest[0], &tris[off + i].v0.w, 4); digest[0] = fnv(digest[0], &tris[off + i].v1.w, 4);
        }
        for (uint32_t i = 0; i + 1 < std::max(n, 2u) && n > 0; ++i) {
            const Bvh2Node& q = n2[off + i];
            digest[1] = fnv(digest[1], &q.left, 4); digest[1] = fnv(digest[1], &q.right, 4);
            digest[2] = fnv(digest[2], q.lo, 12); digest[2] = fnv(digest[2], q.hi, 12);
        }
    }
    digest[3] = b.num_nodes8;
    std::vector<Bvh8Node> n8(std::max<uint32_t>(b.num_nodes8, 1));
    if (b.num_nodes8) PT_CUDA_CHECK(cudaMemcpy(n8.data(), b.nodes8, sizeof(Bvh8Node) * b.num_nodes8, cudaMemcpyDeviceToHost));
    uint64_t sum = 0;
    for (uint32_t i = 0; i < b.num_nodes8; ++i) {
        uint64_t h = fnv(seed, &n8[i], 16);                 // frame + exponents + imask
        uint8_t kinds[8];
        for (int k = 0; k < 8; ++k) kinds[k] = (uint8_t)(((n8[i].valid >> (24 + k)) & 1u) ? 4u : ((n8[i].valid >> (kLeafStride * k)) & 3u));
        h = fnv(h, kinds, 8);
        h = fnv(h, n8[i].qlo_x, 48);
        sum += h;
    }
    digest[4] = sum;
    digest[5] = fnv(seed, b.blas_root8.data(), b.blas_root8.size() * 4);
    return PT_OK;
}

// Animated instances.  The reference fixes the transforms at Tlas.init (Tlas.zig:32-52);
// this is "Tlas.init again" with new matrices: instance records re-derived, top level rebuilt in place.
int pt_set_instance_transforms(PtRenderer* r, const float* transforms, size_t num_instances, float* ms_rebuild) {
    PT_REQUIRE(r && (transforms || num_instances == 0), "pt_set_instance_transforms: null argument");
    PT_REQUIRE(r->owns_scene, "pt_set_instance_transforms: call it on the renderer that owns the scene (not a sibling)");
    PT_REQUIRE(num_instances == r->h_instances.size(), "pt_set_instance_transforms: instance count differs from the scene's");
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    { int rc = pt_sync(r); if (rc != PT_OK) return rc; }       // frames in flight still read the old top level
    for (size_t i = 0; i < num_instances; ++i) std::memcpy(r->h_instances[i].transform, transforms + 16 * i, sizeof(float) * 16);
    if (num_instances) PT_CUDA_CHECK(cudaMemcpyAsync(r->d_instances_raw, r->h_instances.data(), num_instances * sizeof(PtInstance), cudaMemcpyHostToDevice, r->stream));
    int rc = rebuild_tlas(&r->bvh, r->d_instances_raw, r->stream, ms_rebuild);
    if (rc != PT_OK) return rc;
    r->scene.tlas_root = r->bvh.tlas_root8;
    if ((rc = upload_instance_records(r)) != PT_OK) return rc;
    r->accum_frames = 0;
    return PT_OK;
}

}  // extern "C"

namespace {
// scoped device buffers for the pt_debug_* entry points
struct DevBufs {
    std::vector<void*> ptrs;
    ~DevBufs() { for (void* p : ptrs) cudaFree(p); }
    template <typename T> T* in(const T* host, size_t count) {
        void* d = nullptr;
        if (cudaMalloc(&d, std::max<size_t>(count, 1) * sizeof(T)) != cudaSuccess) return nullptr;
        ptrs.push_back(d);
        if (host && count && cudaMemcpy(d, host, count * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) return nullptr;
        return static_cast<T*>(d);
    }
};
int debug_begin(int device) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { set_error("no CUDA device available (no CPU fallback)"); return PT_ERR_CUDA; }
    if (device < 0 || device >= ndev) { set_error("device ordinal out of range"); return PT_ERR_INVALID_ARGUMENT; }
    PT_CUDA_CHECK(cudaSetDevice(device));
    return PT_OK;
}
}  // namespace

extern "C" {

int pt_debug_bsdf(int device, size_t n, const float* N, const float* V, const float* kd, const float* f0, const float* rough,
                  const float* xi, float* L, float* pdf, float* f) {
    PT_REQUIRE(n == 0 || (N && V && kd && f0 && rough && xi && L && pdf && f), "pt_debug_bsdf: null argument");
    int rc = debug_begin(device); if (rc != PT_OK) return rc;
    if (n == 0) return PT_OK;
    DevBufs b;
    float *dN = b.in(N, 3 * n), *dV = b.in(V, 3 * n), *dk = b.in(kd, 3 * n), *d0 = b.in(f0, 3 * n), *dr = b.in(rough, n), *dx = b.in(xi, 2 * n);
    float *dL = b.in<float>(nullptr, 3 * n), *dp = b.in<float>(nullptr, n), *df = b.in<float>(nullptr, 3 * n);
    if (!dN || !dV || !dk || !d0 || !dr || !dx || !dL || !dp || !df) { set_error("pt_debug_bsdf: device allocation/copy failed"); return PT_ERR_CUDA; }
    launch_debug_bsdf((uint32_t)n, dN, dV, dk, d0, dr, dx, dL, dp, df, nullptr);
    PT_CUDA_CHECK(cudaMemcpy(L, dL, 12 * n, cudaMemcpyDeviceToHost));
    PT_CUDA_CHECK(cudaMemcpy(pdf, dp, 4 * n, cudaMemcpyDeviceToHost));
    PT_CUDA_CHECK(cudaMemcpy(f, df, 12 * n, cudaMemcpyDeviceToHost));
    return PT_OK;
}

int pt_debug_rng(int device, size_t n, const uint32_t* pxf, uint32_t draws, uint32_t* state, float* xi) {
    PT_REQUIRE(n == 0 || (pxf && state && xi), "pt_debug_rng: null argument");
    int rc = debug_begin(device); if (rc != PT_OK) return rc;
    if (n == 0) return PT_OK;
    DevBufs b;
    uint32_t* dp = b.in(pxf, 3 * n); uint32_t* ds = b.in<uint32_t>(nullptr, 2 * n); float* dx = b.in<float>(nullptr, 2 * n);
    if (!dp || !ds || !dx) { set_error("pt_debug_rng: device allocation/copy failed"); return PT_ERR_CUDA; }
    launch_debug_rng((uint32_t)n, dp, draws, ds, dx, nullptr);
    PT_CUDA_CHECK(cudaMemcpy(state, ds, 8 * n, cudaMemcpyDeviceToHost));
    PT_CUDA_CHECK(cudaMemcpy(xi, dx, 8 * n, cudaMemcpyDeviceToHost));
    return PT_OK;
}

int pt_debug_texture(int device, const uint8_t* rgba, uint32_t width, uint32_t height, size_t n, const float* uv, float* out) {
    PT_REQUIRE(rgba && width && height && (n == 0 || (uv && out)), "pt_debug_texture: invalid argument");
    int rc = debug_begin(device); if (rc != PT_OK) return rc;
    if (n == 0) return PT_OK;
    DevBufs b;
    uint32_t* dt = b.in(reinterpret_cast<const uint32_t*>(rgba), (size_t)width * height);
    float* du = b.in(uv, 2 * n); float* dout = b.in<float>(nullptr, 4 * n);
    if (!dt || !du || !dout) { set_error("pt_debug_texture: device allocation/copy failed"); return PT_ERR_CUDA; }
    launch_debug_texture(TexRec{dt, width, height}, (uint32_t)n, du, dout, nullptr);
    PT_CUDA_CHECK(cudaMemcpy(out, dout, 16 * n, cudaMemcpyDeviceToHost));
    return PT_OK;
}

// ------------------------------------------------------------------ roofline denominators measured on the box
// Coalesced 16-byte reads of a buffer, `iters` passes by a grid of 8 CTAs per SM: a buffer that fits the L2 (B200: 126 MB)
// gives the L2 read bandwidth the traversal working set can at best be served at, a large one the HBM read
// bandwidth.  The first pass of a small buffer comes from HBM: use iters >= 50.
__global__ void __launch_bounds__(256) k_read_bandwidth(const uint4* __restrict__ buf, size_t n16, int iters, uint32_t* sink) {
    uint4 acc = make_uint4(0u, 0u, 0u, 0u);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int it = 0; it < iters; ++it) {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
            const uint4 v = __ldcg(buf + i);     // L2 only: the L1 must not serve repeated passes
            acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
        }
    }
    if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x9e3779b9u) *sink = 1u;   // keeps the loads alive
}

int pt_debug_read_bandwidth(int device, size_t bytes, int iters, float* gb_per_s, int* num_sms) {
    PT_REQUIRE(gb_per_s && bytes >= 4096 && iters >= 1, "pt_debug_read_bandwidth: invalid argument");
    int rc = debug_begin(device); if (rc != PT_OK) return rc;
    cudaDeviceProp prop;
    PT_CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
    if (num_sms) *num_sms = prop.multiProcessorCount;
    const size_t n16 = bytes / 16;
    uint4* buf = nullptr; uint32_t* sink = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaError_t e = cudaMalloc(&buf, n16 * 16);
    if (e == cudaSuccess) e = cudaMalloc(&sink, 4);
    if (e == cudaSuccess) e = cudaMemset(buf, 0x5a, n16 * 16);
    if (e == cudaSuccess) e = cudaMemset(sink, 0, 4);
    if (e == cudaSuccess) e = cudaEventCreate(&e0);
    if (e == cudaSuccess) e = cudaEventCreate(&e1);
    float ms = 0.0f;
    if (e == cudaSuccess) {
        const int grid = prop.multiProcessorCount * 8;
        k_read_bandwidth<<<grid, 256>>>(buf, n16, 1, sink);            // warm-up: brings a small buffer into the L2
        cudaEventRecord(e0);
        k_read_bandwidth<<<grid, 256>>>(buf, n16, iters, sink);
        cudaEventRecord(e1);
        e = cudaEventSynchronize(e1);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
        if (e == cudaSuccess) e = cudaGetLastError();
    }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(buf); cudaFree(sink);
    if (e != cudaSuccess) { set_error(std::string("pt_debug_read_bandwidth: ") + cudaGetErrorString(e)); return PT_ERR_CUDA; }
    *gb_per_s = ms > 0.0f ? (float)((double)n16 * 16.0 * iters / (ms * 1e-3) / 1e9) : 0.0f;
    return PT_OK;
}

int pt_set_option(const char* key, int value) {
    PT_REQUIRE(key, "pt_set_option: null key");
    if (set_option(key, value) != 0) { set_error(std::string("pt_set_option: unknown key ") + key); return PT_ERR_INVALID_ARGUMENT; }
    return PT_OK;
}

int pt_ipc_export_image(PtRenderer* r, uint8_t handle[64]) {
    PT_REQUIRE(r && handle, "pt_ipc_export_image: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t size");
    cudaIpcMemHandle_t h;
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    PT_CUDA_CHECK(cudaIpcGetMemHandle(&h, r->d_image));
    std::memcpy(handle, &h, 64);
    r->peer_root = true;   // from now on the image is written by every rank: never cleared per frame
    return PT_OK;
}

int pt_ipc_open_root_image(PtRenderer* r, const uint8_t handle[64]) {
    PT_REQUIRE(r && handle, "pt_ipc_open_root_image: null argument");
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, 64);
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    void* p = nullptr;
    PT_CUDA_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    r->d_remote_image = static_cast<float*>(p);
    return PT_OK;
}

}  // extern "C"
