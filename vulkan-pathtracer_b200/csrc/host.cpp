// host.cpp — host-side pieces of the C ABI that need no CUDA: the camera model
// (src/Camera.zig:25-46,86-102) and headless image output (PFM / EXR instead of the swapchain).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/b200pt.h"

namespace pt { void set_error(const std::string& msg); }

namespace {
struct Q { float w, x, y, z; };
struct V { float x, y, z; };

// zalgebra Quat.fromAxis(degrees, axis)
Q from_axis(float degrees, V axis) {
    const float radians = degrees * (3.14159265358979323846f / 180.0f);
    const float s = std::sin(radians * 0.5f), c = std::cos(radians * 0.5f);
    return Q{c, axis.x * s, axis.y * s, axis.z * s};
}
Q qmul(Q l, Q r) {
    return Q{l.w * r.w - l.x * r.x - l.y * r.y - l.z * r.z,
             l.w * r.x + l.x * r.w + l.y * r.z - l.z * r.y,
             l.w * r.y - l.x * r.z + l.y * r.w + l.z * r.x,
             l.w * r.z + l.x * r.y - l.y * r.x + l.z * r.w};
}
V cross(V a, V b) { return V{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
float dot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
V scale(V a, float s) { return V{a.x * s, a.y * s, a.z * s}; }
V add(V a, V b) { return V{a.x + b.x, a.y + b.y, a.z + b.z}; }
// zalgebra Quat.rotateVec
V rotate(Q q, V v) {
    const float n = std::sqrt(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
    q = Q{q.w / n, q.x / n, q.y / n, q.z / n};
    const V b{q.x, q.y, q.z};
    const float b2 = dot(b, b);
    return add(add(scale(v, q.w * q.w - b2), scale(b, dot(v, b) * 2.0f)), scale(cross(b, v), q.w * 2.0f));
}
}  // namespace

extern "C" {

// Camera.new + updateOrientation + updateVectors, packed like the push constants
// (src/RaytracingPass.zig:563-569).
int pt_camera_new(float fov, float aspect_ratio, const float position[3], float yaw_deg, float pitch_deg,
                  uint32_t frame_count, PtCamera* out) {
    if (!position || !out) { pt::set_error("pt_camera_new: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    const Q rotation = qmul(from_axis(yaw_deg, V{0.0f, 1.0f, 0.0f}), from_axis(pitch_deg, V{1.0f, 0.0f, 0.0f}));   // Camera.zig:87-89
    const V forward = rotate(rotation, V{0.0f, 0.0f, 1.0f});    // :91 Vec3.forward()
    const V left = rotate(rotation, V{-1.0f, 0.0f, 0.0f});      // :92 Vec3.left()
    const V up = cross(left, forward);                          // :93
    const float viewport_height = 2.0f * std::tan(fov * 0.5f);  // :97
    const float viewport_width = viewport_height * aspect_ratio;  // :98
    const V hor = scale(left, viewport_width), ver = scale(up, viewport_height);   // :100-101
    out->position[0] = position[0]; out->position[1] = position[1]; out->position[2] = position[2]; out->position[3] = 1.0f;
    out->horizontal[0] = hor.x; out->horizontal[1] = hor.y; out->horizontal[2] = hor.z; out->horizontal[3] = 0.0f;
    out->vertical[0] = ver.x; out->vertical[1] = ver.y; out->vertical[2] = ver.z; out->vertical[3] = 0.0f;
    out->forward[0] = forward.x; out->forward[1] = forward.y; out->forward[2] = forward.z; out->forward[3] = 0.0f;
    out->frame_count = frame_count;
    return PT_OK;
}

// PFM (Portable FloatMap): "PF", bottom row first, little endian; RGB only.
int pt_write_pfm(const char* path, const float* rgba, uint32_t width, uint32_t height) {
    if (!path || !rgba) { pt::set_error("pt_write_pfm: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    FILE* f = std::fopen(path, "wb");
    if (!f) { pt::set_error(std::string("pt_write_pfm: cannot open ") + path); return PT_ERR_IO; }
    std::fprintf(f, "PF\n%u %u\n-1.0\n", width, height);
    std::vector<float> row((size_t)width * 3);
    for (uint32_t y = 0; y < height; ++y) {
        const float* src = rgba + (size_t)(height - 1 - y) * width * 4;
        for (uint32_t x = 0; x < width; ++x) { row[3 * x] = src[4 * x]; row[3 * x + 1] = src[4 * x + 1]; row[3 * x + 2] = src[4 * x + 2]; }
        if (std::fwrite(row.data(), sizeof(float), row.size(), f) != row.size()) { std::fclose(f); pt::set_error("pt_write_pfm: short write"); return PT_ERR_IO; }
    }
    std::fclose(f);
    return PT_OK;
}

// Minimal OpenEXR 2.0 writer: single part, scanline, uncompressed, FLOAT channels A B G R.
int pt_write_exr(const char* path, const float* rgba, uint32_t width, uint32_t height) {
    if (!path || !rgba) { pt::set_error("pt_write_exr: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    std::vector<uint8_t> hdr;
    auto put = [&](const void* p, size_t n) { const uint8_t* b = static_cast<const uint8_t*>(p); hdr.insert(hdr.end(), b, b + n); };
    auto put_str = [&](const char* s) { put(s, std::strlen(s) + 1); };
    auto put_i32 = [&](int32_t v) { put(&v, 4); };
    auto put_f32 = [&](float v) { put(&v, 4); };
    const uint32_t magic = 20000630u, version = 2u;
    put(&magic, 4); put(&version, 4);
    // channels
    put_str("channels"); put_str("chlist");
    put_i32(4 * 18 + 1);
    for (const char* ch : {"A", "B", "G", "R"}) { put_str(ch); put_i32(2 /*FLOAT*/); const uint8_t plin[4] = {0, 0, 0, 0}; put(plin, 4); put_i32(1); put_i32(1); }
    { const uint8_t z = 0; put(&z, 1); }
    put_str("compression"); put_str("compression"); put_i32(1); { const uint8_t c = 0; put(&c, 1); }
    put_str("dataWindow"); put_str("box2i"); put_i32(16); put_i32(0); put_i32(0); put_i32((int32_t)width - 1); put_i32((int32_t)height - 1);
    put_str("displayWindow"); put_str("box2i"); put_i32(16); put_i32(0); put_i32(0); put_i32((int32_t)width - 1); put_i32((int32_t)height - 1);
    put_str("lineOrder"); put_str("lineOrder"); put_i32(1); { const uint8_t c = 0; put(&c, 1); }
    put_str("pixelAspectRatio"); put_str("float"); put_i32(4); put_f32(1.0f);
    put_str("screenWindowCenter"); put_str("v2f"); put_i32(8); put_f32(0.0f); put_f32(0.0f);
    put_str("screenWindowWidth"); put_str("float"); put_i32(4); put_f32(1.0f);
    { const uint8_t z = 0; put(&z, 1); }
    FILE* f = std::fopen(path, "wb");
    if (!f) { pt::set_error(std::string("pt_write_exr: cannot open ") + path); return PT_ERR_IO; }
    const uint64_t row_bytes = (uint64_t)width * 16, block = 8 + row_bytes;
    const uint64_t table_start = hdr.size(), data_start = table_start + 8ull * height;
    bool ok = std::fwrite(hdr.data(), 1, hdr.size(), f) == hdr.size();
    for (uint32_t y = 0; y < height && ok; ++y) { const uint64_t off = data_start + block * y; ok = std::fwrite(&off, 8, 1, f) == 1; }
    std::vector<float> row((size_t)width * 4);
    for (uint32_t y = 0; y < height && ok; ++y) {
        const int32_t yy = (int32_t)y, sz = (int32_t)row_bytes;
        const float* src = rgba + (size_t)y * width * 4;
        for (uint32_t x = 0; x < width; ++x) {
            row[x] = src[4 * x + 3]; row[width + x] = src[4 * x + 2]; row[2 * (size_t)width + x] = src[4 * x + 1]; row[3 * (size_t)width + x] = src[4 * x];
        }
        ok = std::fwrite(&yy, 4, 1, f) == 1 && std::fwrite(&sz, 4, 1, f) == 1 && std::fwrite(row.data(), 4, row.size(), f) == row.size();
    }
    std::fclose(f);
    if (!ok) { pt::set_error("pt_write_exr: short write"); return PT_ERR_IO; }
    return PT_OK;
}

}  // extern "C"
