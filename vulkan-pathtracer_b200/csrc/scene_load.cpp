// scene_load.cpp — Scene.load (src/Scene.zig:79-546) behind the C ABI: glTF 2.0 (.gltf with external
// buffers/images, or .glb) -> the host arrays `Scene` holds, in the reference's exact layout.
//
// What the reference delegates to un-vendored dependencies is restated here from the published
// formats: zgltf (JSON / GLB container, glTF-2.0 defaults), stb_image (PNG: RFC 2083 + zlib inflate;
// baseline JPEG: ITU T.81, see image_decode.cpp) and zalgebra (Mat4 translate/mul/scale, Quat.toMat4 —
// order and arithmetic as tools/scene.py::SceneBuilder.pack).
// Reference quirks kept on purpose: URIs resolve against the cwd, not the file (Scene.zig:125,414);
// accessors are copied tightly packed and byteStride is ignored (:262-331); TRS only on nodes that
// carry a mesh and node.matrix ignored (:507-530); LIFO DFS instance order (:501-541).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200pt.h"
#include "image_decode.h"

namespace pt { void set_error(const std::string& msg); }

namespace {

struct LoadError {
    int status;
    std::string msg;
};
[[noreturn]] void fail(int status, const std::string& msg) { throw LoadError{status, msg}; }

// ------------------------------------------------------------------ JSON (RFC 8259) DOM
struct Json {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    const Json* get(const char* key) const {
        if (kind != Obj) return nullptr;
        for (const auto& kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
    size_t size() const { return kind == Arr ? arr.size() : 0; }
};

class JsonParser {
  public:
    JsonParser(const char* p, size_t n) : p_(p), end_(p + n) {}
    Json parse() {
        Json v = value(0);
        ws();
        if (p_ != end_) err("trailing characters");
        return v;
    }

  private:
    const char *p_, *end_;
    [[noreturn]] void err(const char* what) { fail(PT_ERR_GLTF_PARSE, std::string("glTF JSON: ") + what); }
    void ws() { while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) ++p_; }
    bool lit(const char* s) {
        const size_t n = std::strlen(s);
        if ((size_t)(end_ - p_) >= n && !std::memcmp(p_, s, n)) { p_ += n; return true; }
        return false;
    }
    static void utf8(std::string& out, uint32_t c) {
        if (c < 0x80) out += (char)c;
        else if (c < 0x800) { out += (char)(0xC0 | (c >> 6)); out += (char)(0x80 | (c & 63)); }
        else if (c < 0x10000) { out += (char)(0xE0 | (c >> 12)); out += (char)(0x80 | ((c >> 6) & 63)); out += (char)(0x80 | (c & 63)); }
        else { out += (char)(0xF0 | (c >> 18)); out += (char)(0x80 | ((c >> 12) & 63)); out += (char)(0x80 | ((c >> 6) & 63)); out += (char)(0x80 | (c & 63)); }
    }
    uint32_t hex4() {
        if (end_ - p_ < 4) err("bad \\u escape");
        uint32_t v = 0;
        for (int i = 0; i < 4; ++i) {
            const char c = *p_++;
            v <<= 4;
            if (c >= '0' && c <= '9') v |= (uint32_t)(c - '0');
            else if (c >= 'a' && c <= 'f') v |= (uint32_t)(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') v |= (uint32_t)(c - 'A' + 10);
            else err("bad \\u escape");
        }
        return v;
    }
    std::string string() {
        std::string out;
        ++p_;  // opening quote
        for (;;) {
            if (p_ >= end_) err("unterminated string");
            const char c = *p_++;
            if (c == '"') break;
            if (c != '\\') { out += c; continue; }
            if (p_ >= end_) err("unterminated escape");
            const char e = *p_++;
            switch (e) {
                case '"': out += '"'; break;
                case '\\': out += '\\'; break;
                case '/': out += '/'; break;
                case 'b': out += '\b'; break;
                case 'f': out += '\f'; break;
                case 'n': out += '\n'; break;
                case 'r': out += '\r'; break;
                case 't': out += '\t'; break;
                case 'u': {
                    uint32_t c1 = hex4();
                    if (c1 >= 0xD800 && c1 < 0xDC00 && end_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
                        p_ += 2;
                        const uint32_t c2 = hex4();
                        c1 = 0x10000 + ((c1 - 0xD800) << 10) + (c2 - 0xDC00);
                    }
                    utf8(out, c1);
                    break;
                }
                default: err("bad escape");
            }
        }
        return out;
    }
    Json value(int depth) {
        if (depth > 256) err("nesting too deep");
        ws();
        if (p_ >= end_) err("unexpected end");
        Json v;
        const char c = *p_;
        if (c == '{') {
            v.kind = Json::Obj;
            ++p_; ws();
            if (p_ < end_ && *p_ == '}') { ++p_; return v; }
            for (;;) {
                ws();
                if (p_ >= end_ || *p_ != '"') err("expected a key");
                std::string k = string();
                ws();
                if (p_ >= end_ || *p_ != ':') err("expected ':'");
                ++p_;
                v.obj.emplace_back(std::move(k), value(depth + 1));
                ws();
                if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                if (p_ < end_ && *p_ == '}') { ++p_; break; }
                err("expected ',' or '}'");
            }
        } else if (c == '[') {
            v.kind = Json::Arr;
            ++p_; ws();
            if (p_ < end_ && *p_ == ']') { ++p_; return v; }
            for (;;) {
                v.arr.push_back(value(depth + 1));
                ws();
                if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                if (p_ < end_ && *p_ == ']') { ++p_; break; }
                err("expected ',' or ']'");
            }
        } else if (c == '"') {
            v.kind = Json::Str;
            v.str = string();
        } else if (lit("true")) { v.kind = Json::Bool; v.b = true; }
        else if (lit("false")) { v.kind = Json::Bool; v.b = false; }
        else if (lit("null")) { v.kind = Json::Null; }
        else {
            const char* s = p_;
            while (p_ < end_ && (std::strchr("+-.eE", *p_) || (*p_ >= '0' && *p_ <= '9'))) ++p_;
            if (p_ == s) err("unexpected character");
            const std::string t(s, p_);
            char* e = nullptr;
            v.kind = Json::Num;
            v.num = std::strtod(t.c_str(), &e);
            if (!e || *e) err("bad number");
        }
        return v;
    }
};

// ------------------------------------------------------------------ glTF document model (what Scene.zig:6-8 asks zgltf for)
constexpr uint32_t kNone = 0xFFFFFFFFu;

struct Accessor {
    uint32_t buffer_view = kNone;
    uint64_t byte_offset = 0;
    uint32_t component_type = 0;  // 5120..5126
    uint64_t count = 0;
    uint32_t type_count = 0;      // SCALAR 1, VEC2 2, VEC3 3, VEC4 4, MAT2 4, MAT3 9, MAT4 16
    uint64_t component_size() const {
        switch (component_type) {
            case 5120: case 5121: return 1;
            case 5122: case 5123: return 2;
            case 5125: case 5126: return 4;
            default: return 0;
        }
    }
    uint64_t byte_size() const { return count * component_size() * type_count; }
};
struct BufferView { uint32_t buffer = 0; uint64_t byte_offset = 0, byte_length = 0; };
struct GBuffer { std::string uri; uint64_t byte_length = 0; };
struct GImage { std::string uri; uint32_t buffer_view = kNone; };
struct GPrimitive { uint32_t mode = 4, indices = kNone, position = kNone, normal = kNone, tangent = kNone, texcoord_0 = kNone, material = kNone; };
struct GMesh { std::vector<GPrimitive> primitives; };
struct GMaterial {
    float base_color[4] = {1.0f, 1.0f, 1.0f, 1.0f};
    float metallic = 1.0f, roughness = 1.0f;
    float emissive[3] = {0.0f, 0.0f, 0.0f};
    uint32_t base_color_texture = kNone, metallic_roughness_texture = kNone, emissive_texture = kNone, normal_texture = kNone;
};
struct GNode {
    uint32_t mesh = kNone;
    float translation[3] = {0.0f, 0.0f, 0.0f};
    float rotation[4] = {0.0f, 0.0f, 0.0f, 1.0f};  // x y z w
    float scale[3] = {1.0f, 1.0f, 1.0f};
    std::vector<uint32_t> children;
};
struct Gltf {
    std::vector<Accessor> accessors;
    std::vector<BufferView> buffer_views;
    std::vector<GBuffer> buffers;
    std::vector<GImage> images;
    std::vector<uint32_t> textures;  // source image per texture
    std::vector<GMaterial> materials;
    std::vector<GMesh> meshes;
    std::vector<GNode> nodes;
    std::vector<std::vector<uint32_t>> scenes;
};

uint64_t j_u64(const Json* v, uint64_t def) {
    if (!v || v->kind != Json::Num || v->num < 0.0) return def;
    return (uint64_t)v->num;
}
uint32_t j_index(const Json* v) { return (!v || v->kind != Json::Num || v->num < 0.0) ? kNone : (uint32_t)v->num; }
float j_f32(const Json* v, float def) { return (v && v->kind == Json::Num) ? (float)v->num : def; }
void j_floats(const Json* v, float* out, size_t n) {
    if (!v || v->kind != Json::Arr) return;
    for (size_t i = 0; i < n && i < v->arr.size(); ++i) out[i] = j_f32(&v->arr[i], out[i]);
}
uint32_t j_texture(const Json* info) { return info ? j_index(info->get("index")) : kNone; }
const std::vector<Json>& j_array(const Json& doc, const char* key) {
    static const std::vector<Json> empty;
    const Json* v = doc.get(key);
    return (v && v->kind == Json::Arr) ? v->arr : empty;
}

Gltf build_model(const Json& doc) {
    if (doc.kind != Json::Obj) fail(PT_ERR_GLTF_PARSE, "glTF JSON: root is not an object");
    Gltf g;
    for (const Json& a : j_array(doc, "accessors")) {
        Accessor acc;
        acc.buffer_view = j_index(a.get("bufferView"));
        acc.byte_offset = j_u64(a.get("byteOffset"), 0);
        acc.component_type = (uint32_t)j_u64(a.get("componentType"), 0);
        acc.count = j_u64(a.get("count"), 0);
        const Json* t = a.get("type");
        const std::string ts = (t && t->kind == Json::Str) ? t->str : "";
        acc.type_count = ts == "SCALAR" ? 1 : ts == "VEC2" ? 2 : ts == "VEC3" ? 3 : ts == "VEC4" ? 4 : ts == "MAT2" ? 4 : ts == "MAT3" ? 9 : ts == "MAT4" ? 16 : 0;
        if (!acc.type_count || !acc.component_size()) fail(PT_ERR_GLTF_PARSE, "glTF: accessor with an unknown type or componentType");
        g.accessors.push_back(acc);
    }
    for (const Json& v : j_array(doc, "bufferViews")) {
        BufferView bv;
        bv.buffer = (uint32_t)j_u64(v.get("buffer"), 0);
        bv.byte_offset = j_u64(v.get("byteOffset"), 0);
        bv.byte_length = j_u64(v.get("byteLength"), 0);
        g.buffer_views.push_back(bv);
    }
    for (const Json& b : j_array(doc, "buffers")) {
        GBuffer gb;
        const Json* u = b.get("uri");
        if (u && u->kind == Json::Str) gb.uri = u->str;
        gb.byte_length = j_u64(b.get("byteLength"), 0);
        g.buffers.push_back(gb);
    }
    for (const Json& i : j_array(doc, "images")) {
        GImage gi;
        const Json* u = i.get("uri");
        if (u && u->kind == Json::Str) gi.uri = u->str;
        gi.buffer_view = j_index(i.get("bufferView"));
        g.images.push_back(gi);
    }
    for (const Json& t : j_array(doc, "textures")) g.textures.push_back(j_index(t.get("source")));
    for (const Json& m : j_array(doc, "materials")) {
        GMaterial gm;
        if (const Json* pbr = m.get("pbrMetallicRoughness")) {
            j_floats(pbr->get("baseColorFactor"), gm.base_color, 4);
            gm.metallic = j_f32(pbr->get("metallicFactor"), 1.0f);
            gm.roughness = j_f32(pbr->get("roughnessFactor"), 1.0f);
            gm.base_color_texture = j_texture(pbr->get("baseColorTexture"));
            gm.metallic_roughness_texture = j_texture(pbr->get("metallicRoughnessTexture"));
        }
        j_floats(m.get("emissiveFactor"), gm.emissive, 3);
        gm.emissive_texture = j_texture(m.get("emissiveTexture"));
        gm.normal_texture = j_texture(m.get("normalTexture"));
        g.materials.push_back(gm);
    }
    for (const Json& m : j_array(doc, "meshes")) {
        GMesh gm;
        for (const Json& p : j_array(m, "primitives")) {
            GPrimitive gp;
            gp.mode = (uint32_t)j_u64(p.get("mode"), 4);
            gp.indices = j_index(p.get("indices"));
            gp.material = j_index(p.get("material"));
            if (const Json* at = p.get("attributes")) {
                gp.position = j_index(at->get("POSITION"));
                gp.normal = j_index(at->get("NORMAL"));
                gp.tangent = j_index(at->get("TANGENT"));
                gp.texcoord_0 = j_index(at->get("TEXCOORD_0"));
            }
            gm.primitives.push_back(gp);
        }
        g.meshes.push_back(std::move(gm));
    }
    for (const Json& n : j_array(doc, "nodes")) {
        GNode gn;
        gn.mesh = j_index(n.get("mesh"));
        j_floats(n.get("translation"), gn.translation, 3);
        j_floats(n.get("rotation"), gn.rotation, 4);
        j_floats(n.get("scale"), gn.scale, 3);
        for (const Json& c : j_array(n, "children")) gn.children.push_back(j_index(&c));
        g.nodes.push_back(std::move(gn));
    }
    for (const Json& s : j_array(doc, "scenes")) {
        std::vector<uint32_t> roots;
        for (const Json& c : j_array(s, "nodes")) roots.push_back(j_index(&c));
        g.scenes.push_back(std::move(roots));
    }
    // index validation (zgltf would hand out-of-range indices straight to a trapping slice access)
    auto chk = [](uint32_t i, size_t n, const char* what) { if (i != kNone && i >= n) fail(PT_ERR_GLTF_PARSE, std::string("glTF: ") + what + " index out of range"); };
    for (const auto& a : g.accessors) chk(a.buffer_view, g.buffer_views.size(), "bufferView");
    for (const auto& v : g.buffer_views) chk(v.buffer, g.buffers.size(), "buffer");
    for (const auto& i : g.images) chk(i.buffer_view, g.buffer_views.size(), "image bufferView");
    for (uint32_t s : g.textures) { if (s == kNone) fail(PT_ERR_GLTF_PARSE, "glTF: texture without a source image"); chk(s, g.images.size(), "image"); }
    for (const auto& m : g.meshes)
        for (const auto& p : m.primitives)
            for (uint32_t a : {p.indices, p.position, p.normal, p.tangent, p.texcoord_0}) chk(a, g.accessors.size(), "accessor");
    for (const auto& n : g.nodes) {
        chk(n.mesh, g.meshes.size(), "mesh");
        for (uint32_t c : n.children) { if (c == kNone) fail(PT_ERR_GLTF_PARSE, "glTF: bad child index"); chk(c, g.nodes.size(), "node"); }
    }
    for (const auto& s : g.scenes) for (uint32_t r : s) { if (r == kNone) fail(PT_ERR_GLTF_PARSE, "glTF: bad root index"); chk(r, g.nodes.size(), "node"); }
    return g;
}

// ------------------------------------------------------------------ files
struct File {
    FILE* f = nullptr;
    std::string name;
    File() = default;
    File(const File&) = delete;
    File& operator=(const File&) = delete;
    ~File() { if (f) std::fclose(f); }
    void open(const std::string& path) {   // std.fs.cwd().openFile: relative paths resolve against the cwd
        name = path;
        f = std::fopen(path.c_str(), "rb");
        if (!f) fail(PT_ERR_IO, "cannot open " + path);
    }
    // seekTo + read of exactly n bytes, error.EndOfFile otherwise (Scene.zig:269-271)
    void pread_exact(uint64_t offset, void* dst, uint64_t n) const {
        if (fseeko(f, (off_t)offset, SEEK_SET) != 0) fail(PT_ERR_IO, "seek failed in " + name);
        if (n && std::fread(dst, 1, n, f) != n) fail(PT_ERR_IO, "EndOfFile: short read in " + name);
    }
    std::vector<uint8_t> read_all() const {
        if (fseeko(f, 0, SEEK_END) != 0) fail(PT_ERR_IO, "seek failed in " + name);
        const off_t n = ftello(f);
        std::vector<uint8_t> out((size_t)std::max<off_t>(n, 0));
        pread_exact(0, out.data(), out.size());
        return out;
    }
};
struct Buffer {        // Scene.zig:63-66
    uint64_t offset = 0;
    const File* file = nullptr;
};

std::string extension(const std::string& path) {   // std.fs.path.extension
    const size_t slash = path.find_last_of('/');
    const size_t base = slash == std::string::npos ? 0 : slash + 1;
    const size_t dot = path.find_last_of('.');
    if (dot == std::string::npos || dot <= base) return "";
    return path.substr(dot);
}

// ------------------------------------------------------------------ zalgebra restated (f32, no contraction)
// Mat4 as rows of M.  mul = plain product with k-ascending accumulation; translate / scale
// PRE-multiply (T*self, S*self) as zalgebra's Mat4.translate / Mat4.scale do (assumed
// from the published library, not verifiable here), so Scene.zig:516-526 yields M = S*((T*P)*R).
struct Mat4 { float m[4][4]; };
Mat4 identity() {
    Mat4 r{};
    for (int i = 0; i < 4; ++i) r.m[i][i] = 1.0f;
    return r;
}
Mat4 mul(const Mat4& a, const Mat4& b) {
    Mat4 r{};
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            float s = 0.0f;
            for (int k = 0; k < 4; ++k) { const float p = a.m[i][k] * b.m[k][j]; s = s + p; }
            r.m[i][j] = s;
        }
    return r;
}
Mat4 quat_to_mat4(float w, float x, float y, float z) {   // Quat.toMat4 (normalises first)
    const float n = std::sqrt(w * w + x * x + y * y + z * z);
    w = w / n; x = x / n; y = y / n; z = z / n;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
    Mat4 r = identity();
    r.m[0][0] = 1.0f - 2.0f * (yy + zz); r.m[0][1] = 2.0f * (xy - wz); r.m[0][2] = 2.0f * (xz + wy);
    r.m[1][0] = 2.0f * (xy + wz); r.m[1][1] = 1.0f - 2.0f * (xx + zz); r.m[1][2] = 2.0f * (yz - wx);
    r.m[2][0] = 2.0f * (xz - wy); r.m[2][1] = 2.0f * (yz + wx); r.m[2][2] = 1.0f - 2.0f * (xx + yy);
    return r;
}

// ------------------------------------------------------------------ the loaded scene
struct Scene {
    std::vector<uint8_t> triangle_data;
    std::vector<PtPrimitive> primitives;
    std::vector<PtMesh> meshes;
    std::vector<PtMaterial> materials;
    std::vector<std::vector<uint8_t>> texture_data;
    std::vector<PtTexture> textures;
    std::vector<PtInstance> instances;
};

uint64_t ceil_pow2(uint64_t v) {   // std.math.ceilPowerOfTwoAssert
    uint64_t p = 1;
    while (p < v) p <<= 1;
    return p;
}
uint64_t align_forward(uint64_t v, uint64_t a) { return a ? (v + a - 1) / a * a : v; }

// Scene.zig:160-352
void load_meshes(Scene& sc, const Gltf& g, const std::vector<Buffer>& buffers) {
    uint64_t sizes[5] = {0, 0, 0, 0, 0};
    uint64_t largest = 0;
    size_t n_prims = 0;
    for (const GMesh& m : g.meshes) {
        for (const GPrimitive& p : m.primitives) {
            if (p.mode != 4) fail(PT_ERR_GLTF_NOT_TRIANGLES, "GltfNotATriangleTopology");                 // :176
            if (p.indices == kNone) fail(PT_ERR_GLTF_NO_INDICES, "GltfNoIndices");                         // :178
            if (p.position == kNone) fail(PT_ERR_GLTF_NO_POSITIONS, "GltfNoPositions");                    // :179
            if (p.normal == kNone) fail(PT_ERR_GLTF_NO_NORMALS, "GltfNoNormals");                          // :180
            if (p.tangent == kNone) fail(PT_ERR_GLTF_NO_TANGENTS, "GltfNoTangents");                       // :181
            if (p.texcoord_0 == kNone) fail(PT_ERR_GLTF_NO_TEXCOORDS, "GltfNoTextureCoordinates");         // :182
            const uint32_t acc[5] = {p.indices, p.position, p.normal, p.tangent, p.texcoord_0};
            for (int k = 0; k < 5; ++k) {                                                                  // :192-198
                const uint64_t s = g.accessors[acc[k]].byte_size();
                sizes[k] += s;
                largest = std::max(largest, s);
            }
        }
        PtMesh pm;
        pm.start = (uint32_t)n_prims;
        n_prims += m.primitives.size();
        pm.end = (uint32_t)n_prims;
        sc.meshes.push_back(pm);                                                                           // :201-206
    }
    largest = ceil_pow2(largest);                                                                          // :209
    uint64_t offsets[5], total = 0;
    for (int k = 0; k < 5; ++k) {                                                                          // :211-217
        offsets[k] = total;
        sizes[k] = align_forward(sizes[k], largest);
        total += sizes[k];
    }
    sc.primitives.resize(n_prims);
    sc.triangle_data.assign(total, 0);
    uint64_t cur[5] = {offsets[0], offsets[1], offsets[2], offsets[3], offsets[4]};
    size_t pi = 0;
    for (const GMesh& m : g.meshes) {
        for (const GPrimitive& p : m.primitives) {
            const Accessor& indices = g.accessors[p.indices];
            const Accessor& positions = g.accessors[p.position];
            if (p.material == kNone) fail(PT_ERR_NO_MATERIAL, "NoMaterial");                               // :251
            if (indices.component_type != 5123 && indices.component_type != 5125)
                fail(PT_ERR_INVALID_INDEX_TYPE, "index componentType must be UNSIGNED_SHORT or UNSIGNED_INT");  // `unreachable` :252-256
            cur[0] = align_forward(cur[0], indices.component_size());                                      // :236-240
            PtPrimitive& out = sc.primitives[pi++];
            out.indices_offset = cur[0]; out.positions_offset = cur[1]; out.normals_offset = cur[2];
            out.tangents_offset = cur[3]; out.uvs_offset = cur[4];
            out.max_vertex = (uint32_t)(positions.count - 1);                                              // :248
            out.triangle_count = (uint32_t)(indices.count / 3);                                            // :249
            out.info = (p.material & PT_MATERIAL_INDEX_MASK) | (indices.component_type == 5125 ? PT_UINT32_INDICES_MASK : 0u);
            out._pad = 0;
            const uint32_t acc[5] = {p.indices, p.position, p.normal, p.tangent, p.texcoord_0};
            static const uint32_t want[5] = {0, 3, 3, 4, 2};
            static const char* what[5] = {"", "InvalidPositionType", "InvalidNormalType", "InvalidTangentType", "InvalidTexcoordType"};
            for (int k = 0; k < 5; ++k) {                                                                  // :260-331
                const Accessor& a = g.accessors[acc[k]];
                if (k > 0 && a.type_count != want[k]) fail(PT_ERR_INVALID_ATTRIBUTE_TYPE, what[k]);
                const uint64_t size = k == 0 ? a.count * a.component_size() : a.byte_size();
                if (a.buffer_view == kNone) fail(PT_ERR_NO_BUFFER_VIEW, "NoBufferView");
                const BufferView& bv = g.buffer_views[a.buffer_view];
                const Buffer& b = buffers[bv.buffer];
                if (cur[k] + size > sc.triangle_data.size()) fail(PT_ERR_GLTF_PARSE, "accessor larger than its section");
                b.file->pread_exact(b.offset + a.byte_offset + bv.byte_offset, sc.triangle_data.data() + cur[k], size);
                cur[k] += size;
            }
        }
    }
}

// Scene.zig:354-446.  One RGBA8 image per glTF texture, decoded on up to hardware_concurrency threads.
void load_textures(Scene& sc, const Gltf& g, const std::vector<Buffer>& buffers) {
    const size_t n = g.textures.size();
    sc.texture_data.resize(n);
    sc.textures.resize(n);
    if (n == 0) return;
    // compressed bytes are fetched serially (FILE* is not shared between threads), decoding is threaded
    std::vector<std::vector<uint8_t>> raw(n);
    for (size_t i = 0; i < n; ++i) {
        const GImage& img = g.images[g.textures[i]];
        if (!img.uri.empty()) {
            File f;
            f.open(img.uri);                                                                               // dir = cwd, :414
            raw[i] = f.read_all();
        } else {
            if (img.buffer_view == kNone) fail(PT_ERR_NO_BUFFER_VIEW, "image without uri or bufferView");
            const BufferView& bv = g.buffer_views[img.buffer_view];
            const Buffer& b = buffers[bv.buffer];
            raw[i].resize(bv.byte_length);
            b.file->pread_exact(b.offset + bv.byte_offset, raw[i].data(), bv.byte_length);                // :421-431
        }
    }
    const size_t nthreads = std::max<size_t>(1, std::min<size_t>(n, std::thread::hardware_concurrency()));
    std::vector<std::string> errors(nthreads);
    std::vector<std::thread> threads;
    for (size_t t = 0; t < nthreads; ++t) {
        threads.emplace_back([&, t]() {
            for (size_t i = t; i < n; i += nthreads) {
                uint32_t w = 0, h = 0;
                std::string err;
                if (!pt::decode_image_rgba8(raw[i].data(), raw[i].size(), sc.texture_data[i], w, h, err)) {
                    errors[t] = "texture " + std::to_string(i) + ": " + err;
                    return;
                }
                sc.textures[i].rgba = sc.texture_data[i].data();
                sc.textures[i].width = w;
                sc.textures[i].height = h;
                std::vector<uint8_t>().swap(raw[i]);
            }
        });
    }
    for (auto& th : threads) th.join();
    for (const auto& e : errors) if (!e.empty()) fail(PT_ERR_IMAGE_DECODE, e);
}

// Scene.zig:448-485: factors truncated to 8 bits (@intFromFloat(x * 255.0)), alpha lane 0.
uint32_t unorm8_trunc(float x) {
    const float v = x * 255.0f;
    if (!(v >= 0.0f) || v >= 256.0f) fail(PT_ERR_INVALID_ARGUMENT, "material factor outside [0, 1] (a safety-checked trap in the reference, Scene.zig:458-474)");
    return (uint32_t)v & 0xFFu;
}
void load_materials(Scene& sc, const Gltf& g) {
    for (const GMaterial& m : g.materials) {
        PtMaterial o;
        o.albedo_factor = unorm8_trunc(m.base_color[0]) | unorm8_trunc(m.base_color[1]) << 8 | unorm8_trunc(m.base_color[2]) << 16;
        o.metal_roughness_factor = unorm8_trunc(m.roughness) << 8 | unorm8_trunc(m.metallic) << 16;
        o.emissive_factor = unorm8_trunc(m.emissive[0]) | unorm8_trunc(m.emissive[1]) << 8 | unorm8_trunc(m.emissive[2]) << 16;
        o.albedo_texture_index = m.base_color_texture;
        o.metal_roughness_texture_index = m.metallic_roughness_texture;
        o.emissive_texture_index = m.emissive_texture;
        o.normal_texture_index = m.normal_texture;
        sc.materials.push_back(o);
    }
}

// Scene.zig:487-546
void load_scene(Scene& sc, const Gltf& g) {
    if (g.scenes.empty()) fail(PT_ERR_NO_SCENE, "NoScene");                                                // :495
    if (g.scenes.size() > 1) fail(PT_ERR_TOO_MANY_SCENES, "TooManyScenes");                                // :496
    struct Item { uint32_t index; Mat4 matrix; };
    std::vector<Item> stack;
    size_t visited = 0;
    for (uint32_t root : g.scenes[0]) {
        stack.push_back(Item{root, identity()});
        while (!stack.empty()) {
            const Item n = stack.back();
            stack.pop_back();
            if (++visited > (1u << 26)) fail(PT_ERR_GLTF_PARSE, "node graph has a cycle");
            const GNode& node = g.nodes[n.index];
            Mat4 matrix = n.matrix;
            if (node.mesh != kNone) {
                Mat4 t = identity();
                t.m[0][3] = node.translation[0]; t.m[1][3] = node.translation[1]; t.m[2][3] = node.translation[2];
                matrix = mul(t, n.matrix);                                                                 // :516 translate
                matrix = mul(matrix, quat_to_mat4(node.rotation[3], node.rotation[0], node.rotation[1], node.rotation[2]));  // :518-523
                Mat4 s = identity();
                s.m[0][0] = node.scale[0]; s.m[1][1] = node.scale[1]; s.m[2][2] = node.scale[2];
                matrix = mul(s, matrix);                                                                   // :525 scale
                PtInstance inst;
                inst.mesh_index = node.mesh;
                std::memcpy(inst.transform, matrix.m, sizeof(inst.transform));                             // transposed column-major = rows of M, :529
                sc.instances.push_back(inst);
            }
            for (uint32_t c : node.children) stack.push_back(Item{c, matrix});                             // :536-541
        }
    }
}

std::unique_ptr<Scene> load(const std::string& path) {
    File main_file;
    main_file.open(path);                                                                                  // :80
    const std::string ext = extension(path);
    std::vector<std::unique_ptr<File>> owned;
    std::vector<Buffer> buffers;
    Gltf g;
    if (ext == ".glb") {                                                                                   // :91-107
        uint32_t hdr[3];
        main_file.pread_exact(0, hdr, 12);
        if (hdr[0] != 0x46546C67u) fail(PT_ERR_GLTF_PARSE, "GLB: bad magic");
        if (hdr[1] != 2u) fail(PT_ERR_GLTF_PARSE, "GLB: unsupported container version");
        uint32_t ch[2];
        main_file.pread_exact(12, ch, 8);
        if (ch[1] != 0x4E4F534Au) fail(PT_ERR_GLTF_PARSE, "GLB: first chunk is not JSON");
        std::vector<char> js(ch[0]);
        main_file.pread_exact(20, js.data(), js.size());
        g = build_model(JsonParser(js.data(), js.size()).parse());
        uint64_t bin_offset = 20ull + ch[0];
        if (bin_offset + 8 <= hdr[2]) {
            uint32_t bc[2];
            main_file.pread_exact(bin_offset, bc, 8);
            if (bc[1] != 0x004E4942u) fail(PT_ERR_GLTF_PARSE, "GLB: second chunk is not BIN");
            bin_offset += 8;
        }
        if (g.buffers.size() > 1) fail(PT_ERR_GLTF_PARSE, "GLB: more than one buffer (the reference maps only the BIN chunk, Scene.zig:98-102)");
        buffers.assign(std::max<size_t>(1, g.buffers.size()), Buffer{bin_offset, &main_file});
    } else if (ext == ".gltf") {                                                                           // :108-134
        const std::vector<uint8_t> js = main_file.read_all();
        g = build_model(JsonParser(reinterpret_cast<const char*>(js.data()), js.size()).parse());
        for (const GBuffer& b : g.buffers) {
            if (b.uri.rfind("data:", 0) == 0) fail(PT_ERR_IO, "data: URIs are not supported (the reference opens the uri as a file, Scene.zig:125)");
            owned.emplace_back(new File());
            owned.back()->open(b.uri);                                                                     // relative to the cwd
            buffers.push_back(Buffer{0, owned.back().get()});
        }
    } else {
        fail(PT_ERR_INVALID_FILE_EXTENSION, "InvalidFileExtension: " + path);                              // :135
    }
    std::unique_ptr<Scene> sc(new Scene());
    load_meshes(*sc, g, buffers);
    load_textures(*sc, g, buffers);
    load_materials(*sc, g);
    load_scene(*sc, g);
    for (const PtPrimitive& p : sc->primitives)
        if ((p.info & PT_MATERIAL_INDEX_MASK) >= sc->materials.size()) fail(PT_ERR_GLTF_PARSE, "primitive material index out of range");
    return sc;
}

}  // namespace

struct PtScene { std::unique_ptr<Scene> s; };

extern "C" {

int pt_scene_load(const char* path, PtScene** out) {
    if (!path || !out) { pt::set_error("pt_scene_load: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    *out = nullptr;
    try {
        std::unique_ptr<Scene> s = load(path);
        *out = new PtScene{std::move(s)};
        return PT_OK;
    } catch (const LoadError& e) {
        pt::set_error(std::string("pt_scene_load(") + path + "): " + e.msg);
        return e.status;
    } catch (const std::bad_alloc&) {
        pt::set_error("pt_scene_load: out of host memory");
        return PT_ERR_OUT_OF_MEMORY;
    } catch (const std::exception& e) {
        pt::set_error(std::string("pt_scene_load: ") + e.what());
        return PT_ERR_GLTF_PARSE;
    }
}

int pt_scene_desc(const PtScene* scene, PtSceneDesc* out) {
    if (!scene || !scene->s || !out) { pt::set_error("pt_scene_desc: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    const Scene& s = *scene->s;
    out->triangle_data = s.triangle_data.data(); out->triangle_data_len = s.triangle_data.size();
    out->primitives = s.primitives.data(); out->num_primitives = s.primitives.size();
    out->meshes = s.meshes.data(); out->num_meshes = s.meshes.size();
    out->materials = s.materials.data(); out->num_materials = s.materials.size();
    out->textures = s.textures.data(); out->num_textures = s.textures.size();
    out->instances = s.instances.data(); out->num_instances = s.instances.size();
    return PT_OK;
}

void pt_scene_free(PtScene* scene) { delete scene; }

}  // extern "C"
