// ref_glsl_api.cpp — C entry points over the REFERENCE's own brdf.glsl / rand.glsl compiled as C++
// (oracle/_ref/brdf.inc, rand.inc are generated from /root/reference by rewrite.sh).
// Test infrastructure: pins oracle.cpp's restatement of the BSDF and RNG to the reference code.
#include "glsl_compat.h"
#include <cstddef>
#undef M_PI

namespace glsl {
#include "rand.inc"
#include "brdf.inc"
}  // namespace glsl

using namespace glsl;

extern "C" {
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
