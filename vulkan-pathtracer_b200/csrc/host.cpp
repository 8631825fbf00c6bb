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

}  // extern "C"

namespace {
V v3(const float* p) { return V{p[0], p[1], p[2]}; }
void put3(float* p, V v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

// Camera.updateOrientation, Camera.zig:86-94
void update_orientation(PtCameraState& c) {
    const Q rotation = qmul(from_axis(c.yaw, V{0.0f, 1.0f, 0.0f}), from_axis(c.pitch, V{1.0f, 0.0f, 0.0f}));
    c.rotation[0] = rotation.w; c.rotation[1] = rotation.x; c.rotation[2] = rotation.y; c.rotation[3] = rotation.z;
    const V forward = rotate(rotation, V{0.0f, 0.0f, 1.0f}), left = rotate(rotation, V{-1.0f, 0.0f, 0.0f});
    put3(c.forward, forward); put3(c.left, left); put3(c.up, cross(left, forward));
}
// Camera.updateVectors, Camera.zig:96-102
void update_vectors(PtCameraState& c) {
    const float viewport_height = 2.0f * std::tan(c.fov * 0.5f);
    const float viewport_width = viewport_height * c.aspect_ratio;
    put3(c.horizontal, scale(v3(c.left), viewport_width));
    put3(c.vertical, scale(v3(c.up), viewport_height));
}
// zalgebra Vec3.norm (pinned 4a85a71, un-vendored): the zero vector is returned unchanged
V norm(V v) {
    const float l = std::sqrt(dot(v, v));
    return l == 0.0f ? v : V{v.x / l, v.y / l, v.z / l};
}
}  // namespace

extern "C" {

int pt_camera_state_new(float fov, float aspect_ratio, const float position[3], PtCameraState* out) {
    if (!position || !out) { pt::set_error("pt_camera_state_new: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    std::memset(out, 0, sizeof(*out));
    out->fov = fov; out->aspect_ratio = aspect_ratio;
    out->position[0] = position[0]; out->position[1] = position[1]; out->position[2] = position[2];
    out->pitch = 0.0f; out->yaw = 0.0f;                    // Camera.zig:32-34
    update_orientation(*out); update_vectors(*out);        // :42-43
    return PT_OK;
}

int pt_camera_update(PtCameraState* c, const PtInput* in) {
    if (!c || !in) { pt::set_error("pt_camera_update: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    if (in->mouse_left) {                                                          // Camera.zig:49
        const float rotate_speed = 0.25f;
        if (in->cursor_delta_x != 0 || in->cursor_delta_y != 0) {                  // :51
            c->yaw -= (float)in->cursor_delta_x * rotate_speed;                   // :54
            c->pitch += (float)in->cursor_delta_y * rotate_speed;                 // :55
            if (c->yaw < 0.0f) c->yaw += 360.0f;                                   // :57
            if (c->yaw >= 360.0f) c->yaw -= 360.0f;                                // :58
            c->pitch = std::fmin(std::fmax(c->pitch, -90.0f), 90.0f);              // :60
            update_orientation(*c); update_vectors(*c);                            // :62-63
        }
    }
    const float move_speed = (in->keys & PT_KEY_LEFT_SHIFT) ? 10.0f : 5.0f;        // :67
    V dir{0.0f, 0.0f, 0.0f};
    const V fwd = v3(c->forward), left = v3(c->left);
    if (in->keys & PT_KEY_W) dir = add(dir, fwd);                                   // :69
    if (in->keys & PT_KEY_S) dir = add(dir, scale(fwd, -1.0f));                     // :70  (sub)
    if (in->keys & PT_KEY_D) dir = add(dir, left);                                  // :71
    if (in->keys & PT_KEY_A) dir = add(dir, scale(left, -1.0f));                    // :72
    const float step = move_speed * in->delta_time;
    const V velocity = scale(norm(dir), step);                                      // :74
    c->position[0] += velocity.x; c->position[1] += velocity.y; c->position[2] += velocity.z;   // :75
    if (in->keys & PT_KEY_SPACE) c->position[1] += step;                            // :77
    if (in->keys & PT_KEY_LEFT_CTRL) c->position[1] -= step;                        // :78
    return PT_OK;
}

int pt_camera_update_aspect_ratio(PtCameraState* c, float aspect_ratio) {
    if (!c) { pt::set_error("pt_camera_update_aspect_ratio: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    c->aspect_ratio = aspect_ratio;
    update_vectors(*c);
    return PT_OK;
}

int pt_camera_pack(const PtCameraState* c, uint32_t frame_count, PtCamera* out) {
    if (!c || !out) { pt::set_error("pt_camera_pack: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    for (int k = 0; k < 3; ++k) {
        out->position[k] = c->position[k]; out->horizontal[k] = c->horizontal[k];
        out->vertical[k] = c->vertical[k]; out->forward[k] = c->forward[k];
    }
    out->position[3] = 1.0f; out->horizontal[3] = 0.0f; out->vertical[3] = 0.0f; out->forward[3] = 0.0f;
    out->frame_count = frame_count;
    return PT_OK;
}

}  // extern "C"

extern "C" {

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
