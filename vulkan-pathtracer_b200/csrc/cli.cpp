// b200pt — headless command line of the B200 trace-loop backend.  Same flags, defaults and exit
// behaviour as the reference's `vulkan-pathtracer` command (src/main.zig:29-84): --num-samples/-c (1),
// --num-bounces/-b (2), --resolution-x/-x (1920), --resolution-y/-y (1080),
// --render-resolution-divider/-d (1), --enable-validation/-v (accepted, no effect: no Vulkan), and
// --scene-path/-s; help exits 0, a parse error exits 2, a missing scene path prints the reference's
// message and returns normally.  The window/swapchain loop (main.zig:232-345) becomes `--frames N`
// iterations of Camera -> RaytracingPass.record with frame_count = 0..N-1, and the present becomes a
// PFM / EXR / raw BGRA8 file (--output).  `--gpus N` (additive) splits every frame into interleaved 32x8
// tiles over N devices of this process — scene replicated, one renderer per device — and combines the shares with one
// NCCL reduce to device 0 (pt_comm_init_all / pt_reduce_image / pt_reduce_accum): the written image is bit-identical to
// the one-GPU image.  Only the C ABI of include/b200pt.h is used.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/b200pt.h"

namespace {

struct Args {
    uint32_t num_samples = 1, num_bounces = 2, resolution_x = 1920, resolution_y = 1080, divider = 1;
    bool enable_validation = false;
    std::string scene_path;
    // additive (headless) options
    std::string output = "out.pfm";
    std::string camera_path;
    uint32_t frames = 1;
    bool accumulate = false, quiet = false;
    float position[3] = {0.0f, 0.0f, 0.0f};
    float yaw = 0.0f, pitch = 0.0f;
    int device = 0;
    uint32_t gpus = 1;
};

void usage(FILE* f) {
    std::fputs(
        "usage: b200pt --scene-path <file.gltf|file.glb> [options]\n"
        "  -c, --num-samples <u32>                 samples per pixel per frame (default 1)\n"
        "  -b, --num-bounces <u32>                 ray segments per sample (default 2)\n"
        "  -x, --resolution-x <u32>                (default 1920)\n"
        "  -y, --resolution-y <u32>                (default 1080)\n"
        "  -d, --render-resolution-divider <u32>   render at resolution / d, nearest up-scale (default 1)\n"
        "  -v, --enable-validation <bool>          accepted for compatibility; no effect\n"
        "  -s, --scene-path <path>\n"
        "headless additions:\n"
        "  -o, --output <file.pfm|.exr|.bgra>      image written after the last frame (default out.pfm)\n"
        "  -f, --frames <u32>                      frames to render, frame_count = 0..N-1 (default 1)\n"
        "      --accumulate                        write the mean over all frames instead of the last frame\n"
        "      --camera-position <x,y,z>  --yaw <deg>  --pitch <deg>   (defaults 0,0,0 / 0 / 0)\n"
        "      --camera-path <file>                replay recorded input through Camera.update, one line per frame:\n"
        "                                          <delta_time_s> <keys|-> <mouse_left 0|1> <cursor_dx> <cursor_dy>\n"
        "                                          keys joined by '+': w s a d space left_ctrl left_shift\n"
        "      --gpus <n>                          tile-partition every frame over devices device..device+n-1, NCCL reduce to the first\n"
        "      --device <ordinal>  --quiet\n",
        f);
}

bool parse_u32(const char* s, uint32_t& out) {
    if (!s || !*s) return false;
    char* e = nullptr;
    const unsigned long long v = std::strtoull(s, &e, 10);
    if (!e || *e || v > 0xFFFFFFFFull || s[0] == '-') return false;
    out = (uint32_t)v;
    return true;
}
bool parse_f32(const char* s, float& out) {
    if (!s || !*s) return false;
    char* e = nullptr;
    out = std::strtof(s, &e);
    return e && !*e;
}
bool parse_bool(const char* s, bool& out) {
    if (!std::strcmp(s, "true") || !std::strcmp(s, "1")) { out = true; return true; }
    if (!std::strcmp(s, "false") || !std::strcmp(s, "0")) { out = false; return true; }
    return false;
}

// returns 0 = ok, 1 = help, 2 = parse error
int parse(int argc, char** argv, Args& a) {
    for (int i = 1; i < argc; ++i) {
        std::string key = argv[i];
        const char* val = nullptr;
        std::string inline_val;
        if (key.rfind("--", 0) == 0) {
            const size_t eq = key.find('=');
            if (eq != std::string::npos) { inline_val = key.substr(eq + 1); key = key.substr(0, eq); val = inline_val.c_str(); }
        }
        auto is = [&](const char* longf, const char* shortf) { return key == longf || (shortf && key == shortf); };
        if (is("--help", "-h")) return 1;
        if (is("--accumulate", nullptr)) { a.accumulate = true; continue; }
        if (is("--quiet", nullptr)) { a.quiet = true; continue; }
        if (is("--enable-validation", "-v")) {
            // bool flag: an optional true/false may follow
            if (val) { if (!parse_bool(val, a.enable_validation)) return 2; }
            else if (i + 1 < argc && parse_bool(argv[i + 1], a.enable_validation)) ++i;
            else a.enable_validation = true;
            continue;
        }
        if (!val) {
            if (i + 1 >= argc) { std::fprintf(stderr, "b200pt: missing value for %s\n", key.c_str()); return 2; }
            val = argv[++i];
        }
        bool ok = true;
        if (is("--num-samples", "-c")) ok = parse_u32(val, a.num_samples);
        else if (is("--num-bounces", "-b")) ok = parse_u32(val, a.num_bounces);
        else if (is("--resolution-x", "-x")) ok = parse_u32(val, a.resolution_x);
        else if (is("--resolution-y", "-y")) ok = parse_u32(val, a.resolution_y);
        else if (is("--render-resolution-divider", "-d")) ok = parse_u32(val, a.divider);
        else if (is("--scene-path", "-s")) a.scene_path = val;
        else if (is("--output", "-o")) a.output = val;
        else if (is("--frames", "-f")) ok = parse_u32(val, a.frames);
        else if (is("--camera-path", nullptr)) a.camera_path = val;
        else if (is("--yaw", nullptr)) ok = parse_f32(val, a.yaw);
        else if (is("--pitch", nullptr)) ok = parse_f32(val, a.pitch);
        else if (is("--device", nullptr)) { uint32_t d = 0; ok = parse_u32(val, d); a.device = (int)d; }
        else if (is("--gpus", nullptr)) ok = parse_u32(val, a.gpus) && a.gpus >= 1 && a.gpus <= 64;
        else if (is("--camera-position", nullptr)) ok = std::sscanf(val, "%f,%f,%f", &a.position[0], &a.position[1], &a.position[2]) == 3;
        else { std::fprintf(stderr, "b200pt: unknown argument %s\n", key.c_str()); return 2; }
        if (!ok) { std::fprintf(stderr, "b200pt: invalid value '%s' for %s\n", val, key.c_str()); return 2; }
    }
    return 0;
}

// recorded Input + Timer values for Camera.update (Camera.zig:48-79), one PtInput per frame
bool read_camera_path(const std::string& path, std::vector<PtInput>& out) {
    FILE* f = std::fopen(path.c_str(), "r");
    if (!f) { std::fprintf(stderr, "b200pt: cannot open camera path %s\n", path.c_str()); return false; }
    char line[512];
    int lineno = 0;
    while (std::fgets(line, sizeof(line), f)) {
        ++lineno;
        char* p = line;
        while (*p == ' ' || *p == '\t') ++p;
        if (*p == '#' || *p == '\n' || *p == '\r' || !*p) continue;
        char keys[256];
        PtInput in;
        std::memset(&in, 0, sizeof(in));
        unsigned mouse = 0;
        if (std::sscanf(p, "%f %255s %u %d %d", &in.delta_time, keys, &mouse, &in.cursor_delta_x, &in.cursor_delta_y) != 5) {
            std::fprintf(stderr, "b200pt: %s:%d: expected <dt> <keys|-> <mouse 0|1> <dx> <dy>\n", path.c_str(), lineno);
            std::fclose(f); return false;
        }
        in.mouse_left = mouse ? 1u : 0u;
        for (char* tok = std::strtok(keys, "+"); tok; tok = std::strtok(nullptr, "+")) {
            static const struct { const char* name; uint32_t bit; } names[] = {
                {"w", PT_KEY_W}, {"s", PT_KEY_S}, {"d", PT_KEY_D}, {"a", PT_KEY_A}, {"space", PT_KEY_SPACE},
                {"left_ctrl", PT_KEY_LEFT_CTRL}, {"left_shift", PT_KEY_LEFT_SHIFT}};
            if (!std::strcmp(tok, "-")) continue;
            bool known = false;
            for (const auto& n : names) if (!std::strcmp(tok, n.name)) { in.keys |= n.bit; known = true; }
            if (!known) { std::fprintf(stderr, "b200pt: %s:%d: unknown key '%s'\n", path.c_str(), lineno, tok); std::fclose(f); return false; }
        }
        out.push_back(in);
    }
    std::fclose(f);
    return true;
}

bool ends_with(const std::string& s, const char* suffix) {
    const size_t n = std::strlen(suffix);
    return s.size() >= n && !s.compare(s.size() - n, n, suffix);
}

int die(const char* what, int status) {
    std::fprintf(stderr, "b200pt: %s failed: %s (%s)\n", what, pt_last_error(), pt_status_name(status));
    return 1;
}

}  // namespace

int main(int argc, char** argv) {
    Args a;
    const int pr = parse(argc, argv, a);
    if (pr == 1) { usage(stdout); return 0; }                 // error.Help   -> exit(0)   main.zig:75
    if (pr == 2) { usage(stderr); return 2; }                 // error.Parser -> exit(2)   main.zig:76
    if (a.scene_path.empty()) {                               // main.zig:81-84
        std::fputs("Missing path to scene from arguments\n", stderr);
        return 0;
    }
    if (a.resolution_x == 0 || a.resolution_y == 0 || a.divider == 0 || a.frames == 0) {
        std::fputs("b200pt: resolution, divider and frames must be non-zero\n", stderr);
        return 2;
    }
    std::vector<PtInput> path;
    if (!a.camera_path.empty() && !read_camera_path(a.camera_path, path)) return 2;
    using clock = std::chrono::steady_clock;
    const auto t_load = clock::now();
    PtScene* scene = nullptr;
    int st = pt_scene_load(a.scene_path.c_str(), &scene);                         // Scene.load, RaytracingPass.zig:102
    if (st != PT_OK) return die("Scene.load", st);
    PtSceneDesc desc;
    st = pt_scene_desc(scene, &desc);
    if (st != PT_OK) return die("pt_scene_desc", st);
    const double load_s = std::chrono::duration<double>(clock::now() - t_load).count();

    PtRenderConfig cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.width = a.resolution_x; cfg.height = a.resolution_y; cfg.num_samples = a.num_samples; cfg.num_bounces = a.num_bounces;
    cfg.divider = a.divider; cfg.world = (int)a.gpus; cfg.partition = a.gpus > 1 ? PT_PARTITION_TILES : PT_PARTITION_NONE;
    const auto t_create = clock::now();
    std::vector<PtRenderer*> rs(a.gpus, nullptr);
    for (uint32_t g = 0; g < a.gpus && st == PT_OK; ++g) {                         // scene replicated per device
        cfg.device = a.device + (int)g; cfg.rank = (int)g;
        st = pt_create(&desc, &cfg, &rs[g]);                                      // RaytracingPass.init, main.zig:196-206
    }
    pt_scene_free(scene);                                                         // `defer scene.deinit()`, RaytracingPass.zig:103
    if (st != PT_OK) return die("RaytracingPass.init", st);
    if (a.gpus > 1 && (st = pt_comm_init_all(rs.data(), (int)a.gpus)) != PT_OK) return die("pt_comm_init_all", st);
    PtRenderer* r = rs[0];
    const double create_s = std::chrono::duration<double>(clock::now() - t_create).count();
    if (a.accumulate) for (PtRenderer* q : rs) pt_set_accumulate(q, 1);

    const float aspect = (float)a.resolution_x / (float)a.resolution_y;           // main.zig:222
    PtCameraState camera;
    st = pt_camera_state_new(3.14159265358979323846f * 0.25f, aspect, a.position, &camera);   // main.zig:220-224
    if (st != PT_OK) return die("Camera.new", st);
    uint64_t rays = 0;
    const auto t0 = clock::now();
    for (uint32_t frame = 0; frame < a.frames; ++frame) {                         // the present loop, main.zig:232-345
        PtCamera cam;
        if (!path.empty()) {
            // camera.update(input, timer) at the top of the loop body (main.zig:242); the trace holds its last line
            st = pt_camera_update(&camera, &path[frame < path.size() ? frame : path.size() - 1]);
            if (st == PT_OK) st = pt_camera_pack(&camera, frame, &cam);
        } else {
            st = pt_camera_new(3.14159265358979323846f * 0.25f, aspect, a.position, a.yaw, a.pitch, frame, &cam);
        }
        if (st != PT_OK) return die("Camera.update", st);
        for (PtRenderer* q : rs) {                                                // asynchronous: the devices run side by side
            st = pt_render(q, &cam, nullptr);                                     // RaytracingPass.record, main.zig:263-270
            if (st != PT_OK) return die("RaytracingPass.record", st);
        }
        if (!a.quiet) {
            PtStats s;
            for (PtRenderer* q : rs) if (pt_get_stats(q, &s) == PT_OK) rays += s.rays_traced;   // synchronises: the frame fence
        }
    }
    if (a.gpus > 1) {
        // the shares are disjoint and zero elsewhere: one reduce(sum) of the last frame (or of the accumulation buffers)
        st = pt_comm_group_start();
        for (PtRenderer* q : rs) if (st == PT_OK) st = a.accumulate ? pt_reduce_accum(q, 0, nullptr) : pt_reduce_image(q, 0, nullptr);
        if (st == PT_OK) st = pt_comm_group_end();
        if (st != PT_OK) return die("NCCL reduce", st);
    }
    for (PtRenderer* q : rs) if ((st = pt_sync(q)) != PT_OK) return die("pt_sync", st);
    const double render_s = std::chrono::duration<double>(clock::now() - t0).count();

    uint32_t w = 0, h = 0;
    pt_render_extent(r, &w, &h);
    int wr = PT_OK;
    if (ends_with(a.output, ".bgra")) {
        // the reference's own pixel format (b8g8r8a8_unorm storage image), raw, at the render extent
        std::vector<uint8_t> px((size_t)w * h * 4);
        st = pt_read_bgra8(r, px.data());
        if (st != PT_OK) return die("pt_read_bgra8", st);
        FILE* f = std::fopen(a.output.c_str(), "wb");
        if (!f || std::fwrite(px.data(), 1, px.size(), f) != px.size()) { std::fprintf(stderr, "b200pt: cannot write %s\n", a.output.c_str()); return 1; }
        std::fclose(f);
    } else {
        std::vector<float> img;
        uint32_t ow = w, oh = h;
        if (a.accumulate) {
            img.resize((size_t)w * h * 4);
            st = pt_read_accum_mean(r, img.data(), 0);
        } else if (a.divider > 1) {
            ow = a.resolution_x; oh = a.resolution_y;                              // the up-scaling blit, RaytracingPass.zig:599-635
            img.resize((size_t)ow * oh * 4);
            st = pt_read_rgba32f_upscaled(r, img.data());
        } else {
            img.resize((size_t)w * h * 4);
            st = pt_read_rgba32f(r, img.data());
        }
        if (st != PT_OK) return die("image read-back", st);
        wr = ends_with(a.output, ".exr") ? pt_write_exr(a.output.c_str(), img.data(), ow, oh) : pt_write_pfm(a.output.c_str(), img.data(), ow, oh);
        if (wr != PT_OK) return die("image write", wr);
    }
    if (!a.quiet) {
        PtStats s;
        std::memset(&s, 0, sizeof(s));
        pt_get_stats(r, &s);
        // Stats.zig:68-87: frames over the wall window and the mean frame time
        std::printf("scene: %llu triangles, %llu BVH8 nodes (%.1f MB), load %.2f s, init %.2f s (device build %.2f ms)\n",
                    (unsigned long long)s.num_triangles, (unsigned long long)s.bvh_nodes, s.bvh_bytes / 1.0e6, load_s, create_s, s.ms_build);
        if (a.gpus > 1) std::printf("gpus: %u (interleaved 32x8 tiles, NCCL reduce to device %d)\n", a.gpus, a.device);
        std::printf("frames: %u at %ux%u, %u spp, %u segments: %.3f ms/frame, %.1f fps, %.1f spp/s, %.1f Mrays/s -> %s\n", a.frames, w, h,
                    a.num_samples, a.num_bounces, 1e3 * render_s / a.frames, a.frames / render_s, a.num_samples * a.frames / render_s,
                    rays / render_s / 1e6, a.output.c_str());
    }
    for (PtRenderer* q : rs) pt_destroy(q);                                       // RaytracingPass.deinit, main.zig:206
    return 0;
}
