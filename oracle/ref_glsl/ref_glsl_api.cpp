// ref_glsl_api.cpp — C entry points over the REFERENCE's own shader sources compiled as C++:
//   rand.glsl, common.glsl, triangle.glsl, material.glsl, brdf.glsl, ray_gen.rgen, closest_hit.rchit, miss.rmiss
// (oracle/_ref/<same names> are generated from /root/reference/src/shaders by rewrite.py; see its header for
// the rewrite rules).  Test infrastructure: pins oracle.cpp's restatement of the whole raygen loop
// (ray_gen.rgen:33-105), the hit attribute fetch (triangle.glsl:46-112), the material fetch
// (material.glsl:31-69), the BSDF and the RNG to the reference code.
//
// What this file supplies is the *pipeline around* the shaders — what Vulkan supplies in the reference:
//   * descriptor bindings: primitives (RaytracingPass.zig:316-349: device addresses = blob + offsets,
//     flags = Scene.Primitive.info), materials, textures, the storage image, push constants, spec constants;
//   * traceRayEXT: the acceleration-structure query is the driver's (no source): a callback returns the
//     closest hit (the oracle's canonical rule), then the built-ins of the hit are set the way
//     Tlas.zig:32-52 configures them (custom index = meshes[mesh].start, 3x4 transform = rows of M) and the
//     reference's closest_hit.rchit main() — or miss.rmiss main() — runs and fills the payload;
//   * the launch grid (gl_LaunchIDEXT / gl_LaunchSizeEXT), one raygen main() per pixel.
#include "glsl_compat.h"

#include <atomic>
#include <cstddef>
#include <thread>
#include <vector>

#include "../../include/b200pt.h"
#undef M_PI

namespace glsl {

// ---- built-in variables of the ray-tracing stages
struct LaunchXY {
    uint x, y;
    operator uvec2() const { return uvec2(x, y); }
    operator vec2() const { return vec2((float)x, (float)y); }      // GLSL implicit conversion uvec2 -> vec2
};
struct LaunchVar { LaunchXY xy; };
thread_local LaunchVar gl_LaunchIDEXT, gl_LaunchSizeEXT;
thread_local int gl_InstanceCustomIndexEXT, gl_GeometryIndexEXT, gl_PrimitiveID;
thread_local float gl_HitTEXT;
thread_local mat4x3 gl_ObjectToWorldEXT, gl_WorldToObjectEXT;

void traceRayEXT(accelerationStructureEXT as, uint ray_flags, uint cull_mask, uint sbt_offset, uint sbt_stride, uint miss_index,
                 vec3 origin, float tmin, vec3 direction, float tmax, int payload_location);

// ---- the reference's shaders, verbatim apart from rewrite.py's rules
#define main raygen_main
#include "ray_gen.rgen"
#undef main
namespace rchit {
#include "closest_hit.rchit"
}
namespace rmiss {
#include "miss.rmiss"
}

// ---- the pipeline state behind traceRayEXT
typedef int (*TraceFn)(const void* ctx, const float o[3], float tmin, const float d[3], float tmax, int use_bvh, PtHit* out);
struct Pipeline {
    TraceFn trace = nullptr;
    const void* trace_ctx = nullptr;
    int use_bvh = 1;
    const PtSceneDesc* scene = nullptr;
    std::atomic<uint64_t> rays{0};
};
static Pipeline g_pipe;
static thread_local uint64_t t_rays = 0;      // traceRayEXT calls of this thread (summed into g_pipe.rays per worker)
static std::vector<Primitive> g_primitives;
static std::vector<Material> g_materials;
static std::vector<sampler2D> g_textures;

static void set_hit_builtins(const PtHit& h) {
    const PtSceneDesc& d = *g_pipe.scene;
    const PtInstance& inst = d.instances[h.instance];
    const uint32_t custom = d.meshes[inst.mesh_index].start;                    // Tlas.zig:43-46
    gl_InstanceCustomIndexEXT = (int)custom;
    gl_GeometryIndexEXT = (int)(h.primitive_index - custom);                    // RaytracingPass.zig:316-349: geometry g <-> primitive start + g
    gl_PrimitiveID = (int)h.triangle_index;
    gl_HitTEXT = h.t;
    for (int c = 0; c < 4; ++c)                                                 // Tlas.zig:33-41: 3x4 row-major = first three rows of M
        gl_ObjectToWorldEXT.c[c] = vec3(inst.transform[0 * 4 + c], inst.transform[1 * 4 + c], inst.transform[2 * 4 + c]);
    const mat3 ai = inverse(mat3(gl_ObjectToWorldEXT));                          // not read by any reference shader
    gl_WorldToObjectEXT.c[0] = ai.c[0]; gl_WorldToObjectEXT.c[1] = ai.c[1]; gl_WorldToObjectEXT.c[2] = ai.c[2];
    gl_WorldToObjectEXT.c[3] = -(ai * gl_ObjectToWorldEXT.c[3]);
}

void traceRayEXT(accelerationStructureEXT, uint, uint, uint, uint, uint, vec3 origin, float tmin, vec3 direction, float tmax, int) {
    const float o[3] = {origin.x, origin.y, origin.z}, d[3] = {direction.x, direction.y, direction.z};
    PtHit h;
    ++t_rays;
    if (g_pipe.trace(g_pipe.trace_ctx, o, tmin, d, tmax, g_pipe.use_bvh, &h)) {
        set_hit_builtins(h);
        rchit::attribs = vec2(h.b1, h.b2);
        rchit::main();                                                          // closest_hit.rchit:10-17
    } else {
        rmiss::main();                                                          // miss.rmiss:9 (payload untouched)
    }
}

static void bind_scene(const PtSceneDesc* d) {
    g_pipe.scene = d;
    g_primitives.resize(d->num_primitives);
    for (size_t i = 0; i < d->num_primitives; ++i) {                             // RaytracingPass.zig:338-348
        const PtPrimitive& p = d->primitives[i];
        const uint64_t base = (uint64_t)(uintptr_t)d->triangle_data;
        Primitive q;
        q.index_address = base + p.indices_offset; q.position_address = base + p.positions_offset;
        q.normal_address = base + p.normals_offset; q.tangent_address = base + p.tangents_offset;
        q.uv_address = base + p.uvs_offset; q.info = p.info;
        g_primitives[i] = q;
    }
    g_materials.resize(d->num_materials);
    for (size_t i = 0; i < d->num_materials; ++i) {
        const PtMaterial& m = d->materials[i];
        Material q;
        q.albedo_factor = m.albedo_factor; q.metal_roughness_factor = m.metal_roughness_factor; q.emissive_factor = m.emissive_factor;
        q.albedo_texture_index = m.albedo_texture_index; q.metal_roughness_texture_index = m.metal_roughness_texture_index;
        q.emissive_texture_index = m.emissive_texture_index; q.normal_texture_index = m.normal_texture_index;
        g_materials[i] = q;
    }
    g_textures.resize(d->num_textures);
    for (size_t i = 0; i < d->num_textures; ++i) g_textures[i] = sampler2D{d->textures[i].rgba, d->textures[i].width, d->textures[i].height};
    primitives.i = g_primitives.data();
    materials.i = g_materials.data();
    textures = g_textures.data();
}

}  // namespace glsl

using namespace glsl;

extern "C" {
// ---- the whole frame: one reference raygen main() per pixel (vkCmdTraceRaysKHR(width, height, 1))
// Tile partitioning as in orc_render / PT_PARTITION_TILES; unowned pixels are not launched.
uint64_t ref_render(const PtSceneDesc* scene, const PtRenderConfig* cfg, const PtCamera* cam, void* trace_fn, const void* trace_ctx,
                    int use_bvh, int nthreads, float* rgba) {
    bind_scene(scene);
    g_pipe.trace = (TraceFn)trace_fn; g_pipe.trace_ctx = trace_ctx; g_pipe.use_bvh = use_bvh; g_pipe.rays = 0;
    const uint32_t div = cfg->divider ? cfg->divider : 1;
    const uint32_t W = cfg->width / div, H = cfg->height / div;
    NUM_SAMPLES = cfg->num_samples; NUM_BOUNCES = cfg->num_bounces;             // RayTracingPipeline.zig:286-322
    pcs.position = vec4(cam->position[0], cam->position[1], cam->position[2], cam->position[3]);   // RaytracingPass.zig:563-569
    pcs.horizontal = vec4(cam->horizontal[0], cam->horizontal[1], cam->horizontal[2], cam->horizontal[3]);
    pcs.vertical = vec4(cam->vertical[0], cam->vertical[1], cam->vertical[2], cam->vertical[3]);
    pcs.forward = vec4(cam->forward[0], cam->forward[1], cam->forward[2], cam->forward[3]);
    pcs.frame_count = cam->frame_count;
    image = image2D{rgba, W, H};
    if (nthreads < 1) nthreads = 1;
    std::atomic<uint32_t> next_row{0};
    const uint32_t tiles_x = (W + 31) / 32;
    auto worker = [&]() {
        t_rays = 0;
        for (;;) {
            const uint32_t y = next_row.fetch_add(1);
            if (y >= H) { g_pipe.rays.fetch_add(t_rays); break; }
            for (uint32_t x = 0; x < W; ++x) {
                if (cfg->partition == PT_PARTITION_TILES && cfg->world > 1) {
                    const uint32_t tile = (y / 8) * tiles_x + (x / 32);
                    if ((int)(tile % (uint32_t)cfg->world) != cfg->rank) continue;
                }
                gl_LaunchIDEXT.xy.x = x; gl_LaunchIDEXT.xy.y = y;
                gl_LaunchSizeEXT.xy.x = W; gl_LaunchSizeEXT.xy.y = H;
                raygen_main();
            }
        }
    };
    std::vector<std::thread> th;
    for (int i = 1; i < nthreads; ++i) th.emplace_back(worker);
    worker();
    for (auto& t : th) t.join();
    return g_pipe.rays.load();
}

// ---- getTriangleData + getMaterialData of n hits through closest_hit.rchit (record layout of orc_shade_hits)
void ref_shade_hits(const PtSceneDesc* scene, const PtHit* hits, size_t n, float* out) {
    bind_scene(scene);
    for (size_t i = 0; i < n; ++i) {
        set_hit_builtins(hits[i]);
        rchit::attribs = vec2(hits[i].b1, hits[i].b2);
        rchit::main();
        const TriangleData td = getTriangleData(payload);
        const MaterialData md = getMaterialData(td);
        float* o = out + 25 * i;
        o[0] = td.uv.x; o[1] = td.uv.y; o[2] = td.normal.x; o[3] = td.normal.y; o[4] = td.normal.z;
        o[5] = td.tangent.x; o[6] = td.tangent.y; o[7] = td.tangent.z; o[8] = td.tangent.w;
        o[9] = td.geometry_normal.x; o[10] = td.geometry_normal.y; o[11] = td.geometry_normal.z;
        std::memcpy(o + 12, &td.material_index, 4);
        o[13] = md.albedo.x; o[14] = md.albedo.y; o[15] = md.albedo.z; o[16] = md.albedo.w;
        o[17] = md.normal.x; o[18] = md.normal.y; o[19] = md.normal.z;
        o[20] = md.emissive.x; o[21] = md.emissive.y; o[22] = md.emissive.z; o[23] = md.roughness; o[24] = md.metallic;
    }
}

// ---- scalar / batched entry points for the known-answer tests
void ref_prng_seed(uint32_t x, uint32_t y, uint32_t frame, uint32_t out[2]) {
    uvec2 s = prng_seed(vec2((float)x, (float)y), frame);
    out[0] = s.x; out[1] = s.y;
}
void ref_pcg2d(uint32_t seed[2], float out[2]) {
    uvec2 s(seed[0], seed[1]);
    vec2 r = pcg2d(s);
    seed[0] = s.x; seed[1] = s.y; out[0] = r.x; out[1] = r.y;
}
void ref_bsdf_batch(size_t n, const float* N, const float* V, const float* kd, const float* f0, const float* rough, const float* xi,
                    float* L_out, float* pdf_out, float* f_out) {
    for (size_t i = 0; i < n; ++i) {
        ShadingInfo s;
        s.normal = vec3(N[3 * i], N[3 * i + 1], N[3 * i + 2]);
        s.out_dir = vec3(V[3 * i], V[3 * i + 1], V[3 * i + 2]);
        s.lambert_out = dot(s.normal, s.out_dir);
        s.diffuse_albedo = vec3(kd[3 * i], kd[3 * i + 1], kd[3 * i + 2]);
        s.fresnel_0 = vec3(f0[3 * i], f0[3 * i + 1], f0[3 * i + 2]);
        s.roughness = rough[i];
        vec3 L = sample_brdf(s, vec2(xi[2 * i], xi[2 * i + 1]));
        float pdf = get_brdf_density(s, L);
        vec3 f = brdf(s, L);
        L_out[3 * i] = L.x; L_out[3 * i + 1] = L.y; L_out[3 * i + 2] = L.z;
        pdf_out[i] = pdf;
        f_out[3 * i] = f.x; f_out[3 * i + 1] = f.y; f_out[3 * i + 2] = f.z;
    }
}
void ref_fresnel_schlick(const float f0[3], const float f90[3], float c, float out[3]) {
    vec3 r = fresnel_schlick(vec3(f0[0], f0[1], f0[2]), vec3(f90[0], f90[1], f90[2]), c);
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void ref_shading_space(const float n[3], float out[9]) {
    mat3 m = get_shading_space(vec3(n[0], n[1], n[2]));
    for (int k = 0; k < 3; ++k) { out[3 * k] = m.c[k].x; out[3 * k + 1] = m.c[k].y; out[3 * k + 2] = m.c[k].z; }
}
}
