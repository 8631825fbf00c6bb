// shade.cuh — device restatement of the reference's shading code: rand.glsl (pcg2d, prng_seed),
// triangle.glsl (getTriangleData), material.glsl (getMaterialData) and brdf.glsl.  Each function
// cites the reference lines it follows (paths relative to /root/reference/src/shaders).
#pragma once
#include "pt_internal.h"

namespace pt {

constexpr float kPi = 3.141592653589793238462643f;  // brdf.glsl:4 (GLSL literal is fp32)

// ------------------------------------------------------------------ rand.glsl
// rand.glsl:89-94
PT_D U2 prng_seed(uint32_t px, uint32_t py, uint32_t frame) {
    U2 s; s.x = px; s.y = py;
    s.x = s.x * 0x9e3779b9u + frame;
    s.y = s.y * 0x9e3779b9u + frame;
    s.x ^= s.x >> 16;
    s.y ^= s.y >> 16;
    s.x *= 0x9e3779b9u;
    s.y *= 0x9e3779b9u;
    return s;
}
// rand.glsl:76-87 (2.32830643654e-10 == 2^-32 in fp32; uint->float rounds to nearest even)
PT_D V2 pcg2d(U2& seed) {
    seed.x = 1664525u * seed.x + 1013904223u;
    seed.y = 1664525u * seed.y + 1013904223u;
    seed.x += 1664525u * seed.y;
    seed.y += 1664525u * seed.x;
    seed.x ^= seed.x >> 16;
    seed.y ^= seed.y >> 16;
    seed.x += 1664525u * seed.y;
    seed.y += 1664525u * seed.x;
    seed.x ^= seed.x >> 16;
    seed.y ^= seed.y >> 16;
    V2 r;
    r.x = __uint2float_rn(seed.x) * 2.32830643654e-10f;
    r.y = __uint2float_rn(seed.y) * 2.32830643654e-10f;
    return r;
}

// ------------------------------------------------------------------ brdf.glsl
struct ShadingInfo {  // brdf.glsl:6-13
    V3 normal, out_dir;
    float lambert_out;
    V3 diffuse_albedo, fresnel_0;
    float roughness;
};

// brdf.glsl:15-20
PT_D V3 fresnel_schlick(V3 f0, V3 f90, float cosine_theta) {
    float factor = 1.0f - cosine_theta;
    float factor_squared = factor * factor;
    float factor_fifth = factor_squared * factor_squared * factor;
    return mix(f0, f90, factor_fifth);
}

// brdf.glsl:22-50  (the GGX "distribution" is copied literally)
PT_D V3 brdf(const ShadingInfo& s, V3 light_dir) {
    float n_dot_light = dot(s.normal, light_dir);
    float n_dot_view = s.lambert_out;
    if (gmin(n_dot_light, n_dot_view) < 0.0f) return v3s(0.0f);
    V3 half_vector = normalize(light_dir + s.out_dir);
    float half_dot_view = dot(half_vector, s.out_dir);
    float f90 = (half_dot_view * half_dot_view) * (2.0f * s.roughness) + 0.5f;
    float diffuse_fresnel = fresnel_schlick(v3s(1.0f), v3s(f90), n_dot_view).x *
                            fresnel_schlick(v3s(1.0f), v3s(f90), n_dot_light).x;
    V3 result = diffuse_fresnel * s.diffuse_albedo;
    float half_dot_normal = dot(half_vector, s.normal);
    float roughness_sq = s.roughness * s.roughness;
    float denominator = half_dot_normal * (roughness_sq - 1.0f) + 1.0f;
    float distribution = roughness_sq / (denominator * denominator);
    float masking = n_dot_light * sqrtf((n_dot_view - roughness_sq * n_dot_view) * n_dot_view + roughness_sq);
    float shadowing = n_dot_view * sqrtf((n_dot_light - roughness_sq * n_dot_light) * n_dot_light + roughness_sq);
    float geometry = 0.5f / (masking + shadowing);
    V3 specular_fresnel = fresnel_schlick(s.fresnel_0, v3s(1.0f), gmax(0.0f, half_dot_view));
    result = result + (distribution * geometry) * specular_fresnel;
    return result / kPi;
}

// brdf.glsl:52-66
PT_D V3 sample_ggx_vndf(V3 view_dir, float rx, float ry, V2 rs) {
    V3 transformed_view = normalize(v3(view_dir.x * rx, view_dir.y * ry, view_dir.z));
    float phi = (2.0f * kPi) * rs.x;
    float z = 1.0f - rs.y * (1.0f + transformed_view.z);
    float sin_theta = sqrtf(gmax(0.0f, 1.0f - z * z));
    float sp, cp;
    sincos_cr(phi, sp, cp);
    V3 hemisphere_sample = v3(sin_theta * cp, sin_theta * sp, z);
    V3 sum = hemisphere_sample + transformed_view;
    return normalize(v3(sum.x * rx, sum.y * ry, sum.z));
}

// brdf.glsl:68-79
PT_D float get_ggx_vndf_density(float n_dot_view, float half_dot_normal, float half_dot_view, float roughness) {
    if (half_dot_normal < 0.0f) return 0.0f;
    float roughness_sq = roughness * roughness;
    float inv_roughness_sq = 1.0f - roughness_sq;
    float denominator = n_dot_view + sqrtf(roughness_sq + inv_roughness_sq * n_dot_view * n_dot_view);
    float d_vis = gmax(0.0f, half_dot_view) * (2.0f / kPi) / denominator;
    float m_sq_term = 1.0f - inv_roughness_sq * half_dot_normal * half_dot_normal;
    return d_vis * roughness_sq / (m_sq_term * m_sq_term);
}

// brdf.glsl:81-84; reflect(I,N) = I - 2*dot(N,I)*N
PT_D V3 sample_ggx_in_dir(V3 view_dir, float roughness, V2 rs) {
    V3 h = sample_ggx_vndf(view_dir, roughness, roughness, rs);
    V3 refl = view_dir - (2.0f * dot(h, view_dir)) * h;
    return -refl;
}

// brdf.glsl:86-93
PT_D float get_ggx_in_dir_density(float n_dot_view, V3 view_dir, V3 light_dir, V3 normal, float roughness) {
    V3 half_vector = normalize(light_dir + view_dir);
    float half_dot_view = dot(half_vector, view_dir);
    float half_dot_normal = dot(half_vector, normal);
    float density = get_ggx_vndf_density(n_dot_view, half_dot_normal, half_dot_view, roughness);
    return density / (4.0f * half_dot_view);
}

// brdf.glsl:95-105: columns of the shading-space matrix
struct Onb { V3 tx, ty, tz; };
PT_D Onb get_shading_space(V3 n) {
    float sign = n.z > 0.0f ? 1.0f : -1.0f;
    float a = -1.0f / (sign + n.z);
    float b = n.x * n.y * a;
    Onb o;
    o.tx = v3(1.0f + sign * n.x * n.x * a, sign * b, -sign * n.x);
    o.ty = v3(b, sign + n.y * n.y * a, -n.y);
    o.tz = n;
    return o;
}
PT_D V3 onb_to_world(const Onb& o, V3 v) { return o.tx * v.x + o.ty * v.y + o.tz * v.z; }
PT_D V3 onb_to_local(const Onb& o, V3 w) { return v3(dot(o.tx, w), dot(o.ty, w), dot(o.tz, w)); }

// brdf.glsl:107-113
PT_D V3 sample_hemisphere_psa(V2 rs) {
    float phi = (2.0f * kPi) * rs.x;
    float radius = sqrtf(rs.y);
    float z = sqrtf(1.0f - radius * radius);
    float sp, cp;
    sincos_cr(phi, sp, cp);
    return v3(radius * cp, radius * sp, z);
}
// brdf.glsl:115-117
PT_D float get_hemisphere_psa_density(float sampled_z) { return gmax(0.0f, sampled_z) / kPi; }
// brdf.glsl:119-122
PT_D float get_diffuse_sampling_probability(const ShadingInfo& s) {
    float luminance = dot(s.diffuse_albedo, v3(0.2126f, 0.7152f, 0.0722f));
    return gmin(0.5f, luminance);
}

// brdf.glsl:124-140
PT_D V3 sample_brdf(const ShadingInfo& s, V2 rs) {
    Onb t2w = get_shading_space(s.normal);
    float diffuse_prob = get_diffuse_sampling_probability(s);
    V3 sampled_dir;
    if (rs.x < diffuse_prob) {
        rs.x /= diffuse_prob;
        sampled_dir = onb_to_world(t2w, sample_hemisphere_psa(rs));
    } else {
        rs.x = (rs.x - diffuse_prob) / (1.0f - diffuse_prob);
        V3 local_view = onb_to_local(t2w, s.out_dir);
        V3 local_light = sample_ggx_in_dir(local_view, s.roughness, rs);
        sampled_dir = onb_to_world(t2w, local_light);
    }
    return sampled_dir;
}

// brdf.glsl:142-150
PT_D float get_brdf_density(const ShadingInfo& s, V3 sampled_dir) {
    float diffuse_prob = get_diffuse_sampling_probability(s);
    float specular_density = get_ggx_in_dir_density(s.lambert_out, s.out_dir, sampled_dir, s.normal, s.roughness);
    float diffuse_density = get_hemisphere_psa_density(dot(s.normal, sampled_dir));
    return mixf(specular_density, diffuse_density, diffuse_prob);
}

// ------------------------------------------------------------------ triangle.glsl
struct TriangleData { V2 uv; V3 normal; V4 tangent; V3 geometry_normal; uint32_t material_index; };

PT_D float ldf(const uint8_t* base, uint32_t off) { return __ldg(reinterpret_cast<const float*>(base + off)); }
PT_D V3 ld3(const uint8_t* base, uint32_t off) { return v3(ldf(base, off), ldf(base, off + 4), ldf(base, off + 8)); }

PT_D void tri_indices(const DeviceScene& sc, const PrimRec& p, uint32_t tri, uint32_t id[3]) {
    // triangle.glsl:56-69
    if ((p.info & PT_UINT32_INDICES_MASK) == 0) {
        const uint16_t* q = reinterpret_cast<const uint16_t*>(sc.triangle_data + p.indices_offset) + 3u * tri;
        id[0] = __ldg(q); id[1] = __ldg(q + 1); id[2] = __ldg(q + 2);
    } else {
        const uint32_t* q = reinterpret_cast<const uint32_t*>(sc.triangle_data + p.indices_offset) + 3u * tri;
        id[0] = __ldg(q); id[1] = __ldg(q + 1); id[2] = __ldg(q + 2);
    }
}

// triangle.glsl:46-112.  `v * transpose(inverse(A))` == inverse(A) * v (quirk Q3/Q4, literal).
PT_D TriangleData get_triangle_data(const DeviceScene& sc, const HitRec& h) {
    const PrimRec p = sc.prims[h.primitive_index];
    TriangleData td;
    const V3 c = v3(1.0f - h.b1 - h.b2, h.b1, h.b2);                                   // :50-54
    uint32_t id[3];
    tri_indices(sc, p, h.triangle_index, id);
    const uint8_t* base = sc.triangle_data;
    {                                                                                   // :71-77
        float u0 = ldf(base, p.uvs_offset + 8u * id[0]), v0 = ldf(base, p.uvs_offset + 8u * id[0] + 4);
        float u1 = ldf(base, p.uvs_offset + 8u * id[1]), v1 = ldf(base, p.uvs_offset + 8u * id[1] + 4);
        float u2 = ldf(base, p.uvs_offset + 8u * id[2]), v2 = ldf(base, p.uvs_offset + 8u * id[2] + 4);
        td.uv.x = u0 * c.x + u1 * c.y + u2 * c.z;
        td.uv.y = v0 * c.x + v1 * c.y + v2 * c.z;
    }
    M3 Ainv;                                                                            // :79
    {
        const float* iv = sc.instances[h.instance].inv;
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int k = 0; k < 3; ++k) Ainv.m[r][k] = __ldg(iv + 3 * r + k);
    }
    {                                                                                   // :80-87
        V3 n0 = ld3(base, p.normals_offset + 12u * id[0]);
        V3 n1 = ld3(base, p.normals_offset + 12u * id[1]);
        V3 n2 = ld3(base, p.normals_offset + 12u * id[2]);
        V3 nn = n0 * c.x + n1 * c.y + n2 * c.z;
        td.normal = normalize(mul(Ainv, nn));
    }
    {                                                                                   // :89-97
        V3 t0 = ld3(base, p.tangents_offset + 16u * id[0]); float w0 = ldf(base, p.tangents_offset + 16u * id[0] + 12);
        V3 t1 = ld3(base, p.tangents_offset + 16u * id[1]); float w1 = ldf(base, p.tangents_offset + 16u * id[1] + 12);
        V3 t2 = ld3(base, p.tangents_offset + 16u * id[2]); float w2 = ldf(base, p.tangents_offset + 16u * id[2] + 12);
        V3 tt = t0 * c.x + t1 * c.y + t2 * c.z;
        float tw = w0 * c.x + w1 * c.y + w2 * c.z;
        V3 txyz = normalize(mul(Ainv, tt));
        td.tangent.x = txyz.x; td.tangent.y = txyz.y; td.tangent.z = txyz.z; td.tangent.w = tw;
    }
    {                                                                                   // :99-107
        V3 p0 = ld3(base, p.positions_offset + 12u * id[0]);
        V3 p1 = ld3(base, p.positions_offset + 12u * id[1]);
        V3 p2 = ld3(base, p.positions_offset + 12u * id[2]);
        td.geometry_normal = normalize(mul(Ainv, cross(p1 - p0, p2 - p0)));
    }
    td.material_index = p.info & PT_MATERIAL_INDEX_MASK;                                // :109
    return td;
}

// ------------------------------------------------------------------ material.glsl
// x / 255.0f, correctly rounded, for x in 0..255 without the IEEE division sequence: q = x * RN(1/255)
// followed by one Newton correction with exact remainders (explicit FMAs).  Exhaustively equal to
// (float)x / 255.0f for all 256 inputs (tests/test_gpu_parity.py::test_unorm8_conversion_is_exact).
PT_D float unorm8(uint32_t x) {
    const float xf = __uint_as_float(0x4B000000u | x) - 8388608.0f;   // exact int -> float without I2F
    const float r = 1.0f / 255.0f;
    const float q = xf * r;
    const float rem = __fmaf_rn(-q, 255.0f, xf);
    return __fmaf_rn(rem, r, q);
}
// `lut` (optional): 256 floats, lut[x] = (float)x / 255.0f, in shared memory (K-shade builds it per CTA): one LDS
// per channel instead of the five-instruction sequence above; the same values by construction.
PT_D V4 unpack_unorm4x8(uint32_t p, const float* lut = nullptr) {
    V4 r;
    if (lut) {
        r.x = lut[p & 0xff]; r.y = lut[(p >> 8) & 0xff]; r.z = lut[(p >> 16) & 0xff]; r.w = lut[p >> 24];
    } else {
        r.x = unorm8(p & 0xff); r.y = unorm8((p >> 8) & 0xff);
        r.z = unorm8((p >> 16) & 0xff); r.w = unorm8(p >> 24);
    }
    return r;
}

// texture(sampler2D, uv): bilinear, REPEAT, LOD 0, r8g8b8a8_unorm, one mip
// (RayTracingPipeline.zig:54-70, Image.zig:32, RaytracingPass.zig:505).  Exact fp32 filtering from
// four point fetches (the hardware bilinear path uses 1.8 fixed-point weights).
PT_D uint32_t wrap_texel(float f, uint32_t n) {
    if (!(f >= -2147483648.0f && f < 2147483648.0f)) return 0u;
    const int i = (int)f;
    if ((n & (n - 1u)) == 0u) return (uint32_t)i & (n - 1u);   // power of two: same result as the modulo
    int m = i % (int)n;
    if (m < 0) m += (int)n;
    return (uint32_t)m;
}
// The four texel fetches of one bilinear tap, separated from the filter arithmetic so that the loads
// of all textures of a material are in flight together (the kernel is latency bound).
struct Taps { uint32_t t00, t10, t01, t11; float a, b; };
PT_D Taps fetch_taps(const TexRec& tex, V2 uv) {
    const float W = (float)tex.width, H = (float)tex.height;
    const float u = uv.x * W - 0.5f, v = uv.y * H - 0.5f;
    const float fu = floorf(u), fv = floorf(v);
    Taps t;
    t.a = u - fu; t.b = v - fv;
    const uint32_t x0 = wrap_texel(fu, tex.width), x1 = (x0 + 1 == tex.width) ? 0u : x0 + 1;
    const uint32_t y0 = wrap_texel(fv, tex.height), y1 = (y0 + 1 == tex.height) ? 0u : y0 + 1;
    const uint32_t* r0 = tex.texels + (size_t)y0 * tex.width;
    const uint32_t* r1 = tex.texels + (size_t)y1 * tex.width;
    t.t00 = __ldg(r0 + x0); t.t10 = __ldg(r0 + x1); t.t01 = __ldg(r1 + x0); t.t11 = __ldg(r1 + x1);
    return t;
}
PT_D V4 filter_taps(const Taps& t, const float* lut = nullptr) {
    const V4 t00 = unpack_unorm4x8(t.t00, lut), t10 = unpack_unorm4x8(t.t10, lut), t01 = unpack_unorm4x8(t.t01, lut), t11 = unpack_unorm4x8(t.t11, lut);
    const float a = t.a, b = t.b;
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    V4 r;
    r.x = t00.x * w00 + t10.x * w10 + t01.x * w01 + t11.x * w11;
    r.y = t00.y * w00 + t10.y * w10 + t01.y * w01 + t11.y * w11;
    r.z = t00.z * w00 + t10.z * w10 + t01.z * w01 + t11.z * w11;
    r.w = t00.w * w00 + t10.w * w10 + t01.w * w01 + t11.w * w11;
    return r;
}
PT_D V4 sample_texture(const TexRec& tex, V2 uv, const float* lut = nullptr) { return filter_taps(fetch_taps(tex, uv), lut); }

struct MaterialData { V4 albedo; V3 normal, emissive; float roughness, metallic; };

// material.glsl:31-69
// kTex: 0 = general (interleaved atlas or separate maps per material), 1 = texture objects (PT_FLAG_HW_BILINEAR), 2 = every
// textured material of the scene has an atlas (DeviceScene::atlas_only), so the separate-map path is not compiled in
template <int kTex = 0>
PT_D MaterialData get_material_data(const DeviceScene& sc, const TriangleData& td, const float* lut = nullptr) {
    const MatRec* mp = sc.materials + td.material_index;
    const uint4 fac = __ldg(reinterpret_cast<const uint4*>(mp));     // factors + presence bits
    if (kTex == 1) {
        MaterialData md;
        // PT_FLAG_HW_BILINEAR (A/B): the same material.glsl statements with texture() on the texture units (1.8 fixed-point
        // filter weights: not bit-identical to the fp32 lerp below)
        const ulonglong2 o01 = __ldg(reinterpret_cast<const ulonglong2*>(sc.mat_tex_objs + 4u * td.material_index));
        const ulonglong2 o23 = __ldg(reinterpret_cast<const ulonglong2*>(sc.mat_tex_objs + 4u * td.material_index) + 1);
        float4 ta = make_float4(1.0f, 1.0f, 1.0f, 1.0f), tm = ta, te = ta, tn = ta;
        if (o01.x) ta = tex2D<float4>((cudaTextureObject_t)o01.x, td.uv.x, td.uv.y);
        if (o01.y) tm = tex2D<float4>((cudaTextureObject_t)o01.y, td.uv.x, td.uv.y);
        if (o23.x) te = tex2D<float4>((cudaTextureObject_t)o23.x, td.uv.x, td.uv.y);
        if (o23.y) tn = tex2D<float4>((cudaTextureObject_t)o23.y, td.uv.x, td.uv.y);
        md.albedo = unpack_unorm4x8(fac.x, lut);
        if (o01.x) { md.albedo.x *= ta.x; md.albedo.y *= ta.y; md.albedo.z *= ta.z; md.albedo.w *= ta.w; }
        const V4 mr = unpack_unorm4x8(fac.y, lut);
        md.roughness = mr.y; md.metallic = mr.z;
        if (o01.y) { md.roughness *= tm.y; md.metallic *= tm.z; }
        const V4 e = unpack_unorm4x8(fac.z, lut);
        md.emissive = v3(e.x, e.y, e.z);
        if (o23.x) md.emissive = md.emissive * v3(te.x, te.y, te.z);
        md.normal = td.normal;
        if (o23.y) {
            const V3 T = v3(td.tangent.x, td.tangent.y, td.tangent.z);
            const V3 bitangent = cross(td.normal, T) * td.tangent.w;
            const V3 n = normalize(2.0f * v3(tn.x, tn.y, tn.z) - v3s(1.0f));
            md.normal = T * n.x + bitangent * n.y + td.normal * n.z;
        }
        return md;
    }
    const uint4 aw = __ldg(reinterpret_cast<const uint4*>(mp) + 5);  // interleaved texels, if any
    const uint4* atlas = reinterpret_cast<const uint4*>(((uint64_t)aw.y << 32) | aw.x);
    TexRec tex[4];
    Taps taps[4];
    if (kTex == 2 && !atlas) {
#pragma unroll
        for (int k = 0; k < 4; ++k) tex[k].texels = nullptr;       // atlas-only scene: no atlas means no maps
    } else if (kTex == 2 || atlas) {
        // one set of texel coordinates, four 16-byte fetches: every present map's 2x2 footprint at once
        TexRec at; at.texels = nullptr; at.width = aw.z; at.height = aw.w;
        const float W = (float)at.width, H = (float)at.height;
        const float u = td.uv.x * W - 0.5f, v = td.uv.y * H - 0.5f;
        const float fu = floorf(u), fv = floorf(v);
        const float a = u - fu, b = v - fv;
        const uint32_t x0 = wrap_texel(fu, at.width), x1 = (x0 + 1 == at.width) ? 0u : x0 + 1;
        const uint32_t y0 = wrap_texel(fv, at.height), y1 = (y0 + 1 == at.height) ? 0u : y0 + 1;
        const uint4* r0 = atlas + (size_t)y0 * at.width;
        const uint4* r1 = atlas + (size_t)y1 * at.width;
        const uint4 q00 = __ldg(r0 + x0), q10 = __ldg(r0 + x1), q01 = __ldg(r1 + x0), q11 = __ldg(r1 + x1);
        taps[0] = Taps{q00.x, q10.x, q01.x, q11.x, a, b};
        taps[1] = Taps{q00.y, q10.y, q01.y, q11.y, a, b};
        taps[2] = Taps{q00.z, q10.z, q01.z, q11.z, a, b};
        taps[3] = Taps{q00.w, q10.w, q01.w, q11.w, a, b};
#pragma unroll
        for (int k = 0; k < 4; ++k) tex[k].texels = ((fac.w >> k) & 1u) ? reinterpret_cast<const uint32_t*>(atlas) : nullptr;   // presence only
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint4 w = __ldg(reinterpret_cast<const uint4*>(mp) + 1 + k);
            tex[k].texels = reinterpret_cast<const uint32_t*>(((uint64_t)w.y << 32) | w.x);
            tex[k].width = w.z; tex[k].height = w.w;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (tex[k].texels) taps[k] = fetch_taps(tex[k], td.uv);
    }
    MaterialData md;
    md.albedo = unpack_unorm4x8(fac.x, lut);                                                // :36
    if (tex[0].texels) {                                                                 // :37-39
        const V4 t = filter_taps(taps[0], lut);
        md.albedo.x *= t.x; md.albedo.y *= t.y; md.albedo.z *= t.z; md.albedo.w *= t.w;
    }
    const V4 mr = unpack_unorm4x8(fac.y, lut);                                              // :41-43
    md.roughness = mr.y; md.metallic = mr.z;
    if (tex[1].texels) {                                                                 // :44-48
        const V4 t = filter_taps(taps[1], lut);
        md.roughness *= t.y; md.metallic *= t.z;
    }
    const V4 e = unpack_unorm4x8(fac.z, lut);                                            // :50
    md.emissive = v3(e.x, e.y, e.z);
    if (tex[2].texels) {                                                                 // :51-53
        const V4 t = filter_taps(taps[2], lut);
        md.emissive = md.emissive * v3(t.x, t.y, t.z);
    }
    md.normal = td.normal;                                                               // :55
    if (tex[3].texels) {                                                                 // :56-66
        const V4 t = filter_taps(taps[3], lut);
        const V3 nts = v3(t.x, t.y, t.z);
        const V3 T = v3(td.tangent.x, td.tangent.y, td.tangent.z);
        const V3 bitangent = cross(td.normal, T) * td.tangent.w;
        const V3 n = normalize(2.0f * nts - v3s(1.0f));
        md.normal = T * n.x + bitangent * n.y + td.normal * n.z;
    }
    return md;
}

// The last segment of a sample (ray_gen.rgen:75-77 with b == NUM_BOUNCES - 1) only adds the hit's emission:
// of getTriangleData / getMaterialData nothing but `emissive` is consumed, so only the emissive factor and --
// when the material has an emissive map -- the interpolated uv and that map's four texels are fetched.
// Same expressions as above for everything that is computed, hence the same bits.
template <int kTex = 0>
PT_D V3 get_emissive(const DeviceScene& sc, const HitRec& h, const float* lut = nullptr) {
    const PrimRec p = sc.prims[h.primitive_index];
    const MatRec* mp = sc.materials + (p.info & PT_MATERIAL_INDEX_MASK);
    const uint4 fac = __ldg(reinterpret_cast<const uint4*>(mp));
    const V4 e = unpack_unorm4x8(fac.z, lut);                                            // material.glsl:50
    V3 emissive = v3(e.x, e.y, e.z);
    if ((fac.w >> 2) & 1u) {                                                             // :51-53
        const V3 c = v3(1.0f - h.b1 - h.b2, h.b1, h.b2);
        uint32_t id[3];
        tri_indices(sc, p, h.triangle_index, id);
        const uint8_t* base = sc.triangle_data;
        const float u0 = ldf(base, p.uvs_offset + 8u * id[0]), v0 = ldf(base, p.uvs_offset + 8u * id[0] + 4);
        const float u1 = ldf(base, p.uvs_offset + 8u * id[1]), v1 = ldf(base, p.uvs_offset + 8u * id[1] + 4);
        const float u2 = ldf(base, p.uvs_offset + 8u * id[2]), v2 = ldf(base, p.uvs_offset + 8u * id[2] + 4);
        V2 uv;
        uv.x = u0 * c.x + u1 * c.y + u2 * c.z;
        uv.y = v0 * c.x + v1 * c.y + v2 * c.z;
        if (kTex == 1) {      // PT_FLAG_HW_BILINEAR
            const unsigned long long o = __ldg(sc.mat_tex_objs + 4u * (p.info & PT_MATERIAL_INDEX_MASK) + 2u);
            const float4 te = tex2D<float4>((cudaTextureObject_t)o, uv.x, uv.y);
            return emissive * v3(te.x, te.y, te.z);
        }
        if (kTex == 2) {
            // atlas-only scene: the emissive map is word z of the material's interleaved atlas
            const uint4 aw = __ldg(reinterpret_cast<const uint4*>(mp) + 5);
            const uint4* atlas = reinterpret_cast<const uint4*>(((uint64_t)aw.y << 32) | aw.x);
            const float W = (float)aw.z, H = (float)aw.w;
            const float u = uv.x * W - 0.5f, v = uv.y * H - 0.5f;
            const float fu = floorf(u), fv = floorf(v);
            const uint32_t x0 = wrap_texel(fu, aw.z), x1 = (x0 + 1 == aw.z) ? 0u : x0 + 1;
            const uint32_t y0 = wrap_texel(fv, aw.w), y1 = (y0 + 1 == aw.w) ? 0u : y0 + 1;
            const uint4* r0 = atlas + (size_t)y0 * aw.z;
            const uint4* r1 = atlas + (size_t)y1 * aw.z;
            const Taps tp{__ldg(&r0[x0].z), __ldg(&r0[x1].z), __ldg(&r1[x0].z), __ldg(&r1[x1].z), u - fu, v - fv};
            const V4 t = filter_taps(tp, lut);
            return emissive * v3(t.x, t.y, t.z);
        }
        const uint4 w = __ldg(reinterpret_cast<const uint4*>(mp) + 3);                   // TexRec of the emissive map
        TexRec tex;
        tex.texels = reinterpret_cast<const uint32_t*>(((uint64_t)w.y << 32) | w.x);
        tex.width = w.z; tex.height = w.w;
        const V4 t = sample_texture(tex, uv, lut);
        emissive = emissive * v3(t.x, t.y, t.z);
    }
    return emissive;
}

}  // namespace pt
