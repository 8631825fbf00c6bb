// image_decode.cpp — the texture decoders behind Scene.loadTexture (src/Scene.zig:396-446), which
// in the reference is stb_image (un-vendored, pinned f58f558c…): restated from the format
// specifications.  PNG (RFC 2083: all colour types, 1/2/4/8/16-bit, Adam7, PLTE/tRNS; inflate via
// zlib) decodes exactly in any conforming decoder; 16-bit samples keep their high byte and
// sub-byte greys are scaled to 0..255 as stb_image does.  JPEG (ITU T.81 baseline sequential
// Huffman, 8-bit, 1 or 3 components, any sampling factors, restart markers) uses a float IDCT and
// nearest chroma up-sampling, so texels may differ from stb_image by a few LSB — the synthetic
// scenes use PNG only for that reason (SURVEY.md §8c iv).
#include "image_decode.h"

#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstring>

namespace pt {
namespace {

uint32_t be32(const uint8_t* p) { return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3]; }

// ------------------------------------------------------------------ PNG
int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// un-filters one (sub-)image of w x h pixels in place; `src` holds h rows of (1 + stride) bytes
bool png_unfilter(uint8_t* src, size_t avail, uint32_t w, uint32_t h, uint32_t bits_per_pixel, std::vector<uint8_t>& out,
                  size_t& consumed) {
    const size_t stride = ((size_t)w * bits_per_pixel + 7) / 8;
    const size_t bpp = std::max<size_t>(1, bits_per_pixel / 8);
    consumed = (stride + 1) * h;
    if (consumed > avail) return false;
    out.assign(stride * h, 0);
    for (uint32_t y = 0; y < h; ++y) {
        const uint8_t* in = src + (stride + 1) * y;
        const uint8_t ft = *in++;
        uint8_t* cur = out.data() + stride * y;
        const uint8_t* up = y ? cur - stride : nullptr;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = up ? up[i] : 0, c = (up && i >= bpp) ? up[i - bpp] : 0;
            int v = in[i];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: return false;
            }
            cur[i] = (uint8_t)v;
        }
    }
    return true;
}

bool decode_png(const uint8_t* d, size_t n, std::vector<uint8_t>& rgba, uint32_t& W, uint32_t& H, std::string& err) {
    size_t p = 8;
    uint32_t depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, plte, trns;
    bool have_ihdr = false, done = false;
    while (!done && p + 12 <= n) {
        const uint32_t len = be32(d + p);
        const uint8_t* tag = d + p + 4;
        const uint8_t* body = d + p + 8;
        if ((size_t)len > n - p - 12) { err = "PNG: truncated chunk"; return false; }
        if (!std::memcmp(tag, "IHDR", 4)) {
            if (len < 13) { err = "PNG: bad IHDR"; return false; }
            W = be32(body); H = be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12];
            if (body[10] != 0 || body[11] != 0 || interlace > 1) { err = "PNG: unsupported compression/filter/interlace method"; return false; }
            have_ihdr = true;
        } else if (!std::memcmp(tag, "PLTE", 4)) plte.assign(body, body + len);
        else if (!std::memcmp(tag, "tRNS", 4)) trns.assign(body, body + len);
        else if (!std::memcmp(tag, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
        else if (!std::memcmp(tag, "IEND", 4)) done = true;
        p += 12 + (size_t)len;
    }
    if (!have_ihdr || W == 0 || H == 0 || W > (1u << 24) || H > (1u << 24)) { err = "PNG: missing or invalid IHDR"; return false; }
    static const int channels_of[7] = {1, 0, 3, 1, 2, 0, 4};
    if (ctype > 6 || !channels_of[ctype]) { err = "PNG: bad colour type"; return false; }
    const uint32_t ch = (uint32_t)channels_of[ctype];
    const bool depth_ok = (ctype == 0 && (depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)) ||
                          (ctype == 3 && (depth == 1 || depth == 2 || depth == 4 || depth == 8)) ||
                          ((ctype == 2 || ctype == 4 || ctype == 6) && (depth == 8 || depth == 16));
    if (!depth_ok) { err = "PNG: bad bit depth"; return false; }
    if (ctype == 3 && plte.size() < 3) { err = "PNG: palette image without PLTE"; return false; }
    const uint32_t bits_pp = ch * depth;
    // inflate
    size_t raw_size = 0;
    struct Pass { uint32_t x0, y0, dx, dy; };
    static const Pass adam7[7] = {{0, 0, 8, 8}, {4, 0, 8, 8}, {0, 4, 4, 8}, {2, 0, 4, 4}, {0, 2, 2, 4}, {1, 0, 2, 2}, {0, 1, 1, 2}};
    static const Pass whole[1] = {{0, 0, 1, 1}};
    const Pass* passes = interlace ? adam7 : whole;
    const int n_pass = interlace ? 7 : 1;
    for (int k = 0; k < n_pass; ++k) {
        const uint32_t pw = (W > passes[k].x0) ? (W - passes[k].x0 + passes[k].dx - 1) / passes[k].dx : 0;
        const uint32_t ph = (H > passes[k].y0) ? (H - passes[k].y0 + passes[k].dy - 1) / passes[k].dy : 0;
        if (pw && ph) raw_size += (((size_t)pw * bits_pp + 7) / 8 + 1) * ph;
    }
    std::vector<uint8_t> raw(raw_size);
    {
        z_stream zs;
        std::memset(&zs, 0, sizeof(zs));
        if (inflateInit(&zs) != Z_OK) { err = "PNG: inflateInit failed"; return false; }
        zs.next_in = idat.data(); zs.avail_in = (uInt)idat.size();
        size_t produced = 0;
        int rc = Z_OK;
        while (rc == Z_OK && produced < raw.size()) {
            const size_t chunk = std::min<size_t>(raw.size() - produced, 1u << 30);
            zs.next_out = raw.data() + produced; zs.avail_out = (uInt)chunk;
            rc = inflate(&zs, Z_NO_FLUSH);
            produced += chunk - zs.avail_out;
            if (zs.avail_in == 0 && rc == Z_OK && zs.avail_out != 0) break;
        }
        inflateEnd(&zs);
        if ((rc != Z_OK && rc != Z_STREAM_END) || produced < raw.size()) { err = "PNG: corrupt or short IDAT stream"; return false; }
    }
    rgba.assign((size_t)W * H * 4, 0);
    // colour key (tRNS for grey / RGB): compared on the stored sample values
    int key[3] = {-1, -1, -1};
    if (ctype == 0 && trns.size() >= 2) key[0] = trns[0] << 8 | trns[1];
    if (ctype == 2 && trns.size() >= 6) for (int k = 0; k < 3; ++k) key[k] = trns[2 * k] << 8 | trns[2 * k + 1];
    size_t off = 0;
    std::vector<uint8_t> img;
    for (int k = 0; k < n_pass; ++k) {
        const Pass& ps = passes[k];
        const uint32_t pw = (W > ps.x0) ? (W - ps.x0 + ps.dx - 1) / ps.dx : 0, ph = (H > ps.y0) ? (H - ps.y0 + ps.dy - 1) / ps.dy : 0;
        if (!pw || !ph) continue;
        size_t used = 0;
        if (!png_unfilter(raw.data() + off, raw.size() - off, pw, ph, bits_pp, img, used)) { err = "PNG: bad filter type"; return false; }
        off += used;
        const size_t stride = ((size_t)pw * bits_pp + 7) / 8;
        for (uint32_t y = 0; y < ph; ++y) {
            const uint8_t* row = img.data() + stride * y;
            for (uint32_t x = 0; x < pw; ++x) {
                int s[4] = {0, 0, 0, 0};   // stored sample values per channel
                if (depth == 8) for (uint32_t c = 0; c < ch; ++c) s[c] = row[(size_t)x * ch + c];
                else if (depth == 16) for (uint32_t c = 0; c < ch; ++c) s[c] = row[((size_t)x * ch + c) * 2] << 8 | row[((size_t)x * ch + c) * 2 + 1];
                else {
                    const size_t bit = (size_t)x * depth;
                    s[0] = (row[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1);
                }
                auto to8 = [&](int v) -> uint8_t {
                    if (depth == 16) return (uint8_t)(v >> 8);
                    if (depth == 8) return (uint8_t)v;
                    return (uint8_t)(v * (255 / ((1 << depth) - 1)));
                };
                uint8_t* o = rgba.data() + ((size_t)(ps.y0 + y * ps.dy) * W + (ps.x0 + x * ps.dx)) * 4;
                switch (ctype) {
                    case 0: o[0] = o[1] = o[2] = to8(s[0]); o[3] = (key[0] >= 0 && s[0] == key[0]) ? 0 : 255; break;
                    case 2: o[0] = to8(s[0]); o[1] = to8(s[1]); o[2] = to8(s[2]);
                            o[3] = (key[0] >= 0 && s[0] == key[0] && s[1] == key[1] && s[2] == key[2]) ? 0 : 255; break;
                    case 3: {
                        const size_t pi = (size_t)s[0];
                        if (pi * 3 + 2 < plte.size()) { o[0] = plte[pi * 3]; o[1] = plte[pi * 3 + 1]; o[2] = plte[pi * 3 + 2]; }
                        o[3] = pi < trns.size() ? trns[pi] : 255;
                        break;
                    }
                    case 4: o[0] = o[1] = o[2] = to8(s[0]); o[3] = to8(s[1]); break;
                    default: o[0] = to8(s[0]); o[1] = to8(s[1]); o[2] = to8(s[2]); o[3] = to8(s[3]); break;
                }
            }
        }
    }
    return true;
}

// ------------------------------------------------------------------ JPEG (baseline sequential)
struct Huff {
    uint8_t bits[17] = {0};
    uint8_t vals[256] = {0};
    int mincode[17], maxcode[18], valptr[17];
    bool present = false;
    void build() {
        int code = 0, k = 0;
        for (int l = 1; l <= 16; ++l) {
            valptr[l] = k; mincode[l] = code;
            code += bits[l]; k += bits[l];
            maxcode[l] = bits[l] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7fffffff;
        present = true;
    }
};

struct JpegBits {
    const uint8_t* d; size_t n, p;
    uint32_t acc = 0; int cnt = 0;
    bool hit_marker = false;
    int bit() {
        if (cnt == 0) {
            uint8_t b = 0;
            if (!hit_marker && p < n) {
                b = d[p];
                if (b == 0xFF) {
                    if (p + 1 < n && d[p + 1] == 0x00) p += 2;
                    else { hit_marker = true; b = 0; }   // leave the marker for the caller; feed zeros
                } else ++p;
            }
            acc = b; cnt = 8;
        }
        --cnt;
        return (acc >> cnt) & 1;
    }
    int receive(int s) { int v = 0; for (int i = 0; i < s; ++i) v = (v << 1) | bit(); return v; }
    static int extend(int v, int s) { return (s && v < (1 << (s - 1))) ? v - (1 << s) + 1 : v; }
    int decode(const Huff& h) {
        int code = 0;
        for (int l = 1; l <= 16; ++l) {
            code = (code << 1) | bit();
            if (h.maxcode[l] >= 0 && code <= h.maxcode[l] && code >= h.mincode[l]) return h.vals[h.valptr[l] + code - h.mincode[l]];
        }
        return -1;
    }
    void reset() { acc = 0; cnt = 0; hit_marker = false; }
};

const uint8_t kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                             35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

void idct8x8(const float* in, uint8_t* out, size_t out_stride) {
    static float c[8][8];
    static bool init = false;
    if (!init) {
        for (int x = 0; x < 8; ++x)
            for (int u = 0; u < 8; ++u) c[x][u] = (u == 0 ? std::sqrt(0.125f) : 0.5f) * std::cos((2 * x + 1) * u * 3.14159265358979323846f / 16.0f);
        init = true;
    }
    float tmp[64];
    for (int y = 0; y < 8; ++y)
        for (int x = 0; x < 8; ++x) {
            float s = 0.0f;
            for (int u = 0; u < 8; ++u) s += c[x][u] * in[y * 8 + u];
            tmp[y * 8 + x] = s;
        }
    for (int x = 0; x < 8; ++x)
        for (int y = 0; y < 8; ++y) {
            float s = 0.0f;
            for (int v = 0; v < 8; ++v) s += c[y][v] * tmp[v * 8 + x];
            const float r = std::floor(s + 128.5f);
            out[y * out_stride + x] = (uint8_t)std::min(255.0f, std::max(0.0f, r));
        }
}

bool decode_jpeg(const uint8_t* d, size_t n, std::vector<uint8_t>& rgba, uint32_t& W, uint32_t& H, std::string& err) {
    struct Comp { int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0, pred = 0; uint32_t bw = 0, bh = 0; std::vector<uint8_t> plane; };
    uint16_t qt[4][64] = {};
    Huff hdc[4], hac[4];
    Comp comp[3];
    int ncomp = 0, restart = 0;
    bool have_sof = false;
    size_t p = 2;
    while (p + 4 <= n) {
        if (d[p] != 0xFF) { ++p; continue; }
        const uint8_t m = d[p + 1];
        if (m == 0xFF) { ++p; continue; }
        p += 2;
        if (m == 0xD8 || (m >= 0xD0 && m <= 0xD7) || m == 0x01) continue;
        if (m == 0xD9) break;
        if (p + 2 > n) break;
        const size_t len = (size_t)d[p] << 8 | d[p + 1];
        if (len < 2 || p + len > n) { err = "JPEG: truncated segment"; return false; }
        const uint8_t* s = d + p + 2;
        const size_t sl = len - 2;
        if (m == 0xDB) {
            size_t q = 0;
            while (q < sl) {
                const int pq = s[q] >> 4, tq = s[q] & 15;
                ++q;
                if (tq > 3 || q + (pq ? 128 : 64) > sl) { err = "JPEG: bad DQT"; return false; }
                for (int i = 0; i < 64; ++i) { qt[tq][kZigzag[i]] = pq ? (uint16_t)(s[q] << 8 | s[q + 1]) : s[q]; q += pq ? 2 : 1; }
            }
        } else if (m == 0xC4) {
            size_t q = 0;
            while (q + 17 <= sl) {
                const int tc = s[q] >> 4, th = s[q] & 15;
                if (tc > 1 || th > 3) { err = "JPEG: bad DHT"; return false; }
                Huff& h = tc ? hac[th] : hdc[th];
                int total = 0;
                for (int l = 1; l <= 16; ++l) { h.bits[l] = s[q + l]; total += h.bits[l]; }
                q += 17;
                if (total > 256 || q + total > sl) { err = "JPEG: bad DHT"; return false; }
                std::memcpy(h.vals, s + q, total);
                q += total;
                h.build();
            }
        } else if (m == 0xC0 || m == 0xC1) {
            if (sl < 6 || s[0] != 8) { err = "JPEG: only 8-bit precision is supported"; return false; }
            H = (uint32_t)s[1] << 8 | s[2]; W = (uint32_t)s[3] << 8 | s[4]; ncomp = s[5];
            if ((ncomp != 1 && ncomp != 3) || sl < 6 + 3 * (size_t)ncomp || !W || !H) { err = "JPEG: unsupported component count or size"; return false; }
            for (int i = 0; i < ncomp; ++i) {
                comp[i].id = s[6 + 3 * i]; comp[i].h = s[7 + 3 * i] >> 4; comp[i].v = s[7 + 3 * i] & 15; comp[i].tq = s[8 + 3 * i] & 3;
                if (comp[i].h < 1 || comp[i].h > 4 || comp[i].v < 1 || comp[i].v > 4) { err = "JPEG: bad sampling factors"; return false; }
            }
            have_sof = true;
        } else if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)) {
            err = "JPEG: progressive / lossless / arithmetic coding is not supported (baseline only)";
            return false;
        } else if (m == 0xDD) {
            if (sl >= 2) restart = s[0] << 8 | s[1];
        } else if (m == 0xDA) {
            if (!have_sof) { err = "JPEG: SOS before SOF"; return false; }
            const int ns = s[0];
            if (ns != ncomp || sl < 1 + 2 * (size_t)ns + 3) { err = "JPEG: non-interleaved scans are not supported"; return false; }
            for (int i = 0; i < ns; ++i) {
                const int id = s[1 + 2 * i];
                for (int c = 0; c < ncomp; ++c) if (comp[c].id == id) { comp[c].td = s[2 + 2 * i] >> 4; comp[c].ta = s[2 + 2 * i] & 15; }
            }
            int hmax = 1, vmax = 1;
            for (int c = 0; c < ncomp; ++c) { hmax = std::max(hmax, comp[c].h); vmax = std::max(vmax, comp[c].v); }
            const uint32_t mcux = (W + 8 * hmax - 1) / (8 * hmax), mcuy = (H + 8 * vmax - 1) / (8 * vmax);
            for (int c = 0; c < ncomp; ++c) {
                if (comp[c].td > 3 || comp[c].ta > 3 || !hdc[comp[c].td].present || !hac[comp[c].ta].present) { err = "JPEG: missing Huffman table"; return false; }
                comp[c].bw = mcux * comp[c].h * 8; comp[c].bh = mcuy * comp[c].v * 8;
                comp[c].plane.assign((size_t)comp[c].bw * comp[c].bh, 0);
                comp[c].pred = 0;
            }
            JpegBits br{d, n, p + len};
            int until_restart = restart;
            float block[64];
            for (uint32_t my = 0; my < mcuy; ++my)
                for (uint32_t mx = 0; mx < mcux; ++mx) {
                    if (restart && until_restart == 0) {
                        // expect RSTn
                        br.reset();
                        while (br.p + 1 < n && !(d[br.p] == 0xFF && d[br.p + 1] >= 0xD0 && d[br.p + 1] <= 0xD7)) ++br.p;
                        br.p += 2;
                        for (int c = 0; c < ncomp; ++c) comp[c].pred = 0;
                        until_restart = restart;
                    }
                    for (int c = 0; c < ncomp; ++c)
                        for (int by = 0; by < comp[c].v; ++by)
                            for (int bx = 0; bx < comp[c].h; ++bx) {
                                std::memset(block, 0, sizeof(block));
                                const int t = br.decode(hdc[comp[c].td]);
                                if (t < 0 || t > 15) { err = "JPEG: bad DC code"; return false; }
                                comp[c].pred += JpegBits::extend(br.receive(t), t);
                                block[0] = (float)(comp[c].pred * qt[comp[c].tq][0]);
                                for (int k = 1; k < 64;) {
                                    const int rs = br.decode(hac[comp[c].ta]);
                                    if (rs < 0) { err = "JPEG: bad AC code"; return false; }
                                    const int r = rs >> 4, sz = rs & 15;
                                    if (sz == 0) { if (r == 15) { k += 16; continue; } break; }
                                    k += r;
                                    if (k > 63) { err = "JPEG: AC run past the block"; return false; }
                                    const int z = kZigzag[k];
                                    block[z] = (float)(JpegBits::extend(br.receive(sz), sz) * qt[comp[c].tq][z]);
                                    ++k;
                                }
                                uint8_t* dst = comp[c].plane.data() + ((size_t)(my * comp[c].v + by) * 8) * comp[c].bw + (size_t)(mx * comp[c].h + bx) * 8;
                                idct8x8(block, dst, comp[c].bw);
                            }
                    if (restart) --until_restart;
                }
            // colour conversion, nearest chroma up-sampling
            rgba.assign((size_t)W * H * 4, 255);
            for (uint32_t y = 0; y < H; ++y)
                for (uint32_t x = 0; x < W; ++x) {
                    float v[3] = {0, 128, 128};
                    for (int c = 0; c < ncomp; ++c) {
                        const uint32_t sx = x * comp[c].h / hmax, sy = y * comp[c].v / vmax;
                        v[c] = comp[c].plane[(size_t)sy * comp[c].bw + sx];
                    }
                    uint8_t* o = rgba.data() + ((size_t)y * W + x) * 4;
                    if (ncomp == 1) { o[0] = o[1] = o[2] = (uint8_t)v[0]; continue; }
                    const float cb = v[1] - 128.0f, cr = v[2] - 128.0f;
                    auto q8 = [](float f) { return (uint8_t)std::min(255.0f, std::max(0.0f, std::floor(f + 0.5f))); };
                    o[0] = q8(v[0] + 1.402f * cr);
                    o[1] = q8(v[0] - 0.344136f * cb - 0.714136f * cr);
                    o[2] = q8(v[0] + 1.772f * cb);
                }
            return true;
        }
        p += len;
    }
    err = "JPEG: no scan found";
    return false;
}

}  // namespace

bool decode_image_rgba8(const uint8_t* data, size_t size, std::vector<uint8_t>& rgba, uint32_t& width, uint32_t& height,
                        std::string& error) {
    static const uint8_t png_sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (size >= 8 && !std::memcmp(data, png_sig, 8)) return decode_png(data, size, rgba, width, height, error);
    if (size >= 3 && data[0] == 0xFF && data[1] == 0xD8 && data[2] == 0xFF) return decode_jpeg(data, size, rgba, width, height, error);
    error = "unknown image format (PNG and baseline JPEG are supported)";
    return false;
}

}  // namespace pt
