// pts_hdr.cpp — Radiance .hdr loader, the importance-sampling cache, a procedural sky and image writers.
// Reference: hdrloader.cpp:29-191 (three byte-identical copies in the reference tree) and
// ImportanceSampling_LowDiscrepancySequence/main.cpp:555-652 (calculateHdrCache, "M:").
#include "pts_scene.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace pts {
namespace {

struct Cursor {                                       // fgetc/feof over an in-memory copy of the file
    const unsigned char* p;
    size_t n, at;
    bool eof;
    int get() { if (at < n) return p[at++]; eof = true; return -1; }
    void unget() { if (at > 0) at--; }
};

typedef unsigned char RGBE[4];

// hdrloader.cpp:152-191 — flat or old-style run-length scanline
bool old_decrunch(RGBE* scan, int len, Cursor& c) {
    int rshift = 0;
    while (len > 0) {
        scan[0][0] = (unsigned char)c.get();
        scan[0][1] = (unsigned char)c.get();
        scan[0][2] = (unsigned char)c.get();
        scan[0][3] = (unsigned char)c.get();
        if (c.eof) return false;
        if (scan[0][0] == 1 && scan[0][1] == 1 && scan[0][2] == 1) {
            for (int i = scan[0][3] << rshift; i > 0 && len > 0; i--) {
                std::memcpy(&scan[0][0], &scan[-1][0], 4);
                scan++;
                len--;
            }
            rshift += 8;
        } else {
            scan++;
            len--;
            rshift = 0;
        }
    }
    return true;
}

// hdrloader.cpp:113-150 — new-style per-component RLE
bool decrunch(RGBE* scan, int len, Cursor& c) {
    if (len < 8 || len > 0x7fff) return old_decrunch(scan, len, c);
    int i = c.get();
    if (i != 2) { c.unget(); return old_decrunch(scan, len, c); }
    scan[0][1] = (unsigned char)c.get();
    scan[0][2] = (unsigned char)c.get();
    i = c.get();
    if (scan[0][1] != 2 || (scan[0][2] & 128)) {
        scan[0][0] = 2;
        scan[0][3] = (unsigned char)i;
        return old_decrunch(scan + 1, len - 1, c);
    }
    for (int comp = 0; comp < 4; comp++) {
        for (int j = 0; j < len;) {
            unsigned char code = (unsigned char)c.get();
            if (code > 128) {
                code &= 127;
                unsigned char val = (unsigned char)c.get();
                while (code-- && j < len) scan[j++][comp] = val;
            } else {
                while (code-- && j < len) scan[j++][comp] = (unsigned char)c.get();
            }
            if (c.eof) return false;
        }
    }
    return !c.eof;
}

// hdrloader.cpp:99-104: v = val/256.0f ; d = (float)pow(2, expo) ; v*d  (no special case for E = 0)
inline float convert_component(int expo, int val) {
    float v = (float)val / 256.0f;
    float d = (float)std::ldexp(1.0, expo);
    return v * d;
}

unsigned hash_u32(unsigned x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

}  // namespace
}  // namespace pts

using namespace pts;

extern "C" {

int pts_hdr_load(const char* path, int* width, int* height, float** data) {
    if (!path || !width || !height || !data) { set_error("pts_hdr_load: null argument"); return PTS_EINVAL; }
    FILE* f = std::fopen(path, "rb");
    if (!f) { set_error(std::string("cannot open ") + path); return PTS_EIO; }
    std::fseek(f, 0, SEEK_END);
    long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<unsigned char> buf((size_t)std::max(0L, sz));
    size_t got = buf.empty() ? 0 : std::fread(buf.data(), 1, buf.size(), f);
    std::fclose(f);
    if (got != buf.size() || got < 11 || std::memcmp(buf.data(), "#?RADIANCE", 10) != 0) {
        set_error("not a Radiance #?RADIANCE file");
        return PTS_EFORMAT;
    }
    Cursor c{buf.data(), buf.size(), 11, false};       // 10 magic bytes + fseek(1)   (hdrloader.cpp:38-44)
    int ch = 0, old;
    while (true) {                                     // header commands up to the empty line
        old = ch;
        ch = c.get();
        if (c.eof) { set_error("truncated header"); return PTS_EFORMAT; }
        if (ch == 0xa && old == 0xa) break;
    }
    char reso[200];
    int i = 0;
    while (true) {
        ch = c.get();
        if (c.eof || i >= 198) { set_error("bad resolution line"); return PTS_EFORMAT; }
        reso[i++] = (char)ch;
        if (ch == 0xa) break;
    }
    reso[i] = 0;
    int w = 0, h = 0;
    if (std::sscanf(reso, "-Y %d +X %d", &h, &w) != 2 || w <= 0 || h <= 0) {   // the reference's %ld-into-int, fixed
        set_error("unsupported resolution line (need -Y h +X w)");
        return PTS_EFORMAT;
    }
    float* cols = (float*)std::calloc((size_t)w * h * 3, sizeof(float));
    if (!cols) { set_error("out of memory"); return PTS_ENOMEM; }
    std::vector<unsigned char> scanbuf(((size_t)w + 1) * 4);
    RGBE* scan = reinterpret_cast<RGBE*>(scanbuf.data());
    float* out = cols;
    for (int y = h - 1; y >= 0; y--) {                 // rows in file order (hdrloader.cpp:84-89)
        if (!decrunch(scan, w, c)) { std::free(cols); set_error("truncated scanline data"); return PTS_EFORMAT; }
        for (int x = 0; x < w; x++) {
            int expo = scan[x][3] - 128;
            out[0] = convert_component(expo, scan[x][0]);
            out[1] = convert_component(expo, scan[x][1]);
            out[2] = convert_component(expo, scan[x][2]);
            out += 3;
        }
    }
    *width = w; *height = h; *data = cols;
    return PTS_OK;
}

// Flat (uncompressed) RGBE writer — enough for round-trip tests and for shipping synthetic maps.
int pts_hdr_write(const char* path, const float* data, int w, int h) {
    if (!path || !data || w <= 0 || h <= 0) return PTS_EINVAL;
    FILE* f = std::fopen(path, "wb");
    if (!f) { set_error(std::string("cannot write ") + path); return PTS_EIO; }
    std::fprintf(f, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n", h, w);
    std::vector<unsigned char> row((size_t)w * 4);
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            const float* p = data + 3 * ((size_t)y * w + x);
            float m = std::max(p[0], std::max(p[1], p[2]));
            unsigned char* o = &row[4 * (size_t)x];
            if (!(m > 1e-32f)) { o[0] = o[1] = o[2] = o[3] = 0; continue; }
            int e;
            float scale = std::frexp(m, &e) * 256.0f / m;
            o[0] = (unsigned char)(p[0] * scale); o[1] = (unsigned char)(p[1] * scale);
            o[2] = (unsigned char)(p[2] * scale); o[3] = (unsigned char)(e + 128);
        }
        // a flat scanline must not start with the new-RLE marker bytes (2,2,<128)
        if (w >= 8 && w <= 0x7fff && row[0] == 2 && row[1] == 2 && !(row[2] & 128)) row[0] = 3;
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
    return PTS_OK;
}

// calculateHdrCache, M:555-652.  fp32 running sums in the reference's order; the luminance is the one
// expression evaluated in double (0.2, 0.7, 0.1 are double literals, M:567).
int pts_hdr_cache(const float* HDR, int width, int height, float* cache) {
    if (!HDR || !cache || width <= 0 || height <= 0) { set_error("pts_hdr_cache: bad argument"); return PTS_EINVAL; }
    const size_t W = (size_t)width, H = (size_t)height;
    std::vector<float> pdf(W * H);
    float lumSum = 0.0f;
    for (size_t i = 0; i < H; i++)
        for (size_t j = 0; j < W; j++) {
            const float* p = HDR + 3 * (i * W + j);
            float lum = (float)(0.2 * (double)p[0] + 0.7 * (double)p[1] + 0.1 * (double)p[2]);
            pdf[i * W + j] = lum;
            lumSum += lum;
        }
    for (size_t k = 0; k < W * H; k++) pdf[k] /= lumSum;
    std::vector<float> pdf_x(W, 0.0f);                 // marginal over columns (M:579-583)
    for (size_t j = 0; j < W; j++) {
        float s = 0.0f;
        for (size_t i = 0; i < H; i++) s += pdf[i * W + j];
        pdf_x[j] = s;
    }
    std::vector<float> cdf_x = pdf_x;                  // M:586-588
    for (size_t j = 1; j < W; j++) cdf_x[j] += cdf_x[j - 1];
    std::vector<float> cdf_y(W * H);                   // stored column-major: cdf_y[j*H + i]  (M:591-609)
    for (size_t j = 0; j < W; j++) {
        float run = 0.0f;
        for (size_t i = 0; i < H; i++) {
            float cond = pdf[i * W + j] / pdf_x[j];
            run = (i == 0) ? cond : run + cond;
            cdf_y[j * H + i] = run;
        }
    }
    for (size_t j = 0; j < W; j++) {                   // M:621-636 (note the transposed roles of i and j)
        float xi_2 = (float)j / (float)width;
        for (size_t i = 0; i < H; i++) {
            float xi_1 = (float)i / (float)height;
            size_t x = std::lower_bound(cdf_x.begin(), cdf_x.end(), xi_1) - cdf_x.begin();
            if (x >= W) x = W - 1;                     // the reference would index past the end here (UB); never hit on real maps
            const float* col = &cdf_y[x * H];
            size_t y = std::lower_bound(col, col + H, xi_2) - col;
            float* o = cache + 3 * (i * W + j);
            o[0] = (float)x / (float)width;
            o[1] = (float)y / (float)height;
            o[2] = pdf[i * W + j];
        }
    }
    return PTS_OK;
}

// Procedural sky: horizon-brightened gradient, a small very bright sun, soft clouds from hashed value
// noise, dark ground.  Row 0 = zenith, matching the file order of a loaded .hdr.
int pts_hdr_synth(int width, int height, unsigned seed, float* out) {
    if (!out || width <= 0 || height <= 0) return PTS_EINVAL;
    const float PI = 3.14159265358979f;
    const float sun_phi = 0.9f, sun_theta = 0.75f;
    const float sx = std::cos(sun_theta) * std::cos(sun_phi), sy = std::sin(sun_theta), sz = std::cos(sun_theta) * std::sin(sun_phi);
    for (int i = 0; i < height; i++)
        for (int j = 0; j < width; j++) {
            float u = ((float)j + 0.5f) / (float)width, v = ((float)i + 0.5f) / (float)height;
            float phi = 2.0f * PI * (u - 0.5f), theta = PI * (0.5f - v);
            float dx = std::cos(theta) * std::cos(phi), dy = std::sin(theta), dz = std::cos(theta) * std::sin(phi);
            float r, g, b;
            if (dy >= 0.0f) {
                float t = 1.0f - dy;
                float t4 = t * t * t * t;
                r = 0.25f + 0.9f * t4; g = 0.45f + 0.8f * t4; b = 0.95f + 0.4f * t4;
                unsigned cx = (unsigned)(u * 64.0f), cy = (unsigned)(v * 32.0f);
                float fx = u * 64.0f - (float)cx, fy = v * 32.0f - (float)cy;
                auto val = [&](unsigned a, unsigned c2) { return (float)(hash_u32((a % 64u) * 7919u + c2 * 104729u + seed) >> 8) * (1.0f / 16777216.0f); };
                float n00 = val(cx, cy), n10 = val(cx + 1, cy), n01 = val(cx, cy + 1), n11 = val(cx + 1, cy + 1);
                float cloud = (n00 * (1 - fx) + n10 * fx) * (1 - fy) + (n01 * (1 - fx) + n11 * fx) * fy;
                cloud = std::max(0.0f, cloud - 0.55f) * 2.0f * dy;
                r += cloud; g += cloud; b += cloud;
                float c = dx * sx + dy * sy + dz * sz;
                if (c > 0.9995f) { r += 4000.0f; g += 3600.0f; b += 3000.0f; }
                else if (c > 0.0f) { float c2 = c * c, c8 = c2 * c2 * c2 * c2, c64 = c8 * c8 * c8 * c8 * c8 * c8 * c8 * c8; r += 6.0f * c64; g += 5.0f * c64; b += 4.0f * c64; }
            } else {
                float t = 1.0f + dy;
                r = 0.08f + 0.25f * t * t; g = 0.07f + 0.22f * t * t; b = 0.06f + 0.20f * t * t;
            }
            float* o = out + 3 * ((size_t)i * width + j);
            o[0] = r; o[1] = g; o[2] = b;
        }
    return PTS_OK;
}

// PFM keeps raw radiance (bottom row first, which is also the framebuffer's row order).
int pts_write_pfm(const char* path, const float* rgba, int w, int h) {
    if (!path || !rgba || w <= 0 || h <= 0) return PTS_EINVAL;
    FILE* f = std::fopen(path, "wb");
    if (!f) { set_error(std::string("cannot write ") + path); return PTS_EIO; }
    std::fprintf(f, "PF\n%d %d\n-1.0\n", w, h);
    std::vector<float> row((size_t)w * 3);
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) for (int c = 0; c < 3; c++) row[3 * (size_t)x + c] = rgba[4 * ((size_t)y * w + x) + c];
        std::fwrite(row.data(), sizeof(float), row.size(), f);
    }
    std::fclose(f);
    return PTS_OK;
}

// pass3.fsh:14-24: c * 1/(1 + lum/1.5), lum = 0.3r + 0.6g + 0.1b; written top row first as 8-bit PPM.
int pts_write_ppm_tonemapped(const char* path, const float* rgba, int w, int h) {
    if (!path || !rgba || w <= 0 || h <= 0) return PTS_EINVAL;
    FILE* f = std::fopen(path, "wb");
    if (!f) { set_error(std::string("cannot write ") + path); return PTS_EIO; }
    std::fprintf(f, "P6\n%d %d\n255\n", w, h);
    std::vector<unsigned char> row((size_t)w * 3);
    for (int y = h - 1; y >= 0; y--) {
        for (int x = 0; x < w; x++) {
            const float* p = rgba + 4 * ((size_t)y * w + x);
            float lum = 0.3f * p[0] + 0.6f * p[1] + 0.1f * p[2];
            float k = 1.0f / (1.0f + lum / 1.5f);
            for (int c = 0; c < 3; c++) {
                float vv = p[c] * k;
                vv = vv < 0.0f ? 0.0f : (vv > 1.0f ? 1.0f : vv);
                row[3 * (size_t)x + c] = (unsigned char)(vv * 255.0f + 0.5f);
            }
        }
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
    return PTS_OK;
}

// 8-bit PNG for the display stage (the reference's CPU tracer writes its image through svpng, Basic/main.cpp:169-190,
// a header that is absent from the tree).  Written from the PNG and zlib specifications: IHDR, ONE IDAT holding a zlib
// stream of stored (uncompressed) deflate blocks, IEND; every scanline starts with filter byte 0.
// `pixels` = height rows of width * channels bytes, top row first; channels 3 (RGB) or 4 (RGBA).
namespace {
struct PngOut {
    FILE* f;
    uint32_t crc;
    static uint32_t table_entry(uint32_t n) {
        for (int k = 0; k < 8; k++) n = (n & 1u) ? 0xedb88320u ^ (n >> 1) : n >> 1;
        return n;
    }
    void raw(const unsigned char* p, size_t n) {
        static uint32_t table[256];
        static bool ready = false;
        if (!ready) { for (uint32_t i = 0; i < 256; i++) table[i] = table_entry(i); ready = true; }
        for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xffu] ^ (crc >> 8);
        std::fwrite(p, 1, n, f);
    }
    void u8(unsigned v) { unsigned char b = (unsigned char)v; raw(&b, 1); }
    void u16le(unsigned v) { u8(v & 0xffu); u8((v >> 8) & 0xffu); }
    void u32be(uint32_t v) { u8(v >> 24); u8((v >> 16) & 0xffu); u8((v >> 8) & 0xffu); u8(v & 0xffu); }
    void begin(const char* type, uint32_t length) { u32be(length); crc = 0xffffffffu; raw((const unsigned char*)type, 4); }
    void end() { u32be(~crc); }
};
}  // namespace

int pts_write_png(const char* path, const unsigned char* pixels, int w, int h, int channels) {
    if (!path || !pixels || w <= 0 || h <= 0 || (channels != 3 && channels != 4)) return PTS_EINVAL;
    FILE* f = std::fopen(path, "wb");
    if (!f) { set_error(std::string("cannot write ") + path); return PTS_EIO; }
    PngOut o{f, 0};
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::fwrite(sig, 1, 8, f);
    o.begin("IHDR", 13);
    o.u32be((uint32_t)w); o.u32be((uint32_t)h);
    o.u8(8); o.u8(channels == 3 ? 2 : 6); o.u8(0); o.u8(0); o.u8(0);
    o.end();
    const size_t line = (size_t)w * channels + 1;                  // filter byte + pixels
    const size_t total = line * (size_t)h;
    const size_t nblocks = (total + 65534) / 65535;
    o.begin("IDAT", (uint32_t)(2 + nblocks * 5 + total + 4));
    o.u8(0x78); o.u8(0x01);                                        // zlib header: deflate, 32 K window, no preset dictionary
    uint32_t a = 1, b = 0;                                         // Adler-32 of the uncompressed bytes
    size_t in_block = 0, done = 0;
    auto put = [&](unsigned char v) {
        if (in_block == 0) {
            const size_t n = std::min<size_t>(65535, total - done);
            o.u8(done + n == total ? 1 : 0);                       // BFINAL, BTYPE = 00 (stored)
            o.u16le((unsigned)n); o.u16le((unsigned)(~n & 0xffffu));
            in_block = n;
        }
        o.u8(v);
        a = (a + v) % 65521u; b = (b + a) % 65521u;
        in_block--; done++;
    };
    for (int y = 0; y < h; y++) {
        put(0);
        const unsigned char* row = pixels + (size_t)y * w * channels;
        for (size_t i = 0; i < (size_t)w * channels; i++) put(row[i]);
    }
    o.u32be((b << 16) | a);
    o.end();
    o.begin("IEND", 0);
    o.end();
    const bool ok = !std::ferror(f);
    std::fclose(f);
    if (!ok) { set_error(std::string("write error on ") + path); return PTS_EIO; }
    return PTS_OK;
}

}  // extern "C"
