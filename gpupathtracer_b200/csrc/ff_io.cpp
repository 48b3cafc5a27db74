// ff_io.cpp -- headless image writers (PPM / PNG / EXR / PFM).  The OBJ / MTL reader lives in ff_obj.cpp.
//
// Writers: replace saveToPPM (utilities.h:842-856).  Row 0 is the top of the image (kernel orientation).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "ff_internal.h"

namespace ff {

void set_error(const std::string& msg);

// ---------------------------------------------------------------------------------------------- writers
static unsigned char to_srgb8(float v) {
    if (!(v > 0.f)) return 0;
    if (v >= 1.f) return 255;
    float s = v <= 0.0031308f ? 12.92f * v : 1.055f * std::pow(v, 1.0f / 2.4f) - 0.055f;
    int i = (int)(s * 255.0f + 0.5f);
    return (unsigned char)(i < 0 ? 0 : i > 255 ? 255 : i);
}
static uint32_t crc_table[256];
static void crc_init() {
    static bool done = false;
    if (done) return;
    for (uint32_t n = 0; n < 256; ++n) {
        uint32_t c = n;
        for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
        crc_table[n] = c;
    }
    done = true;
}
static uint32_t crc32(uint32_t crc, const unsigned char* buf, size_t len) {
    crc = ~crc;
    for (size_t i = 0; i < len; ++i) crc = crc_table[(crc ^ buf[i]) & 0xff] ^ (crc >> 8);
    return ~crc;
}
static void put_be32(std::vector<unsigned char>& v, uint32_t x) {
    v.push_back((unsigned char)(x >> 24)); v.push_back((unsigned char)(x >> 16)); v.push_back((unsigned char)(x >> 8)); v.push_back((unsigned char)x);
}
static void png_chunk(std::vector<unsigned char>& out, const char type[4], const std::vector<unsigned char>& data) {
    put_be32(out, (uint32_t)data.size());
    const size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    put_be32(out, crc32(0, out.data() + start, out.size() - start));
}
// 8-bit RGB PNG, zlib stream of stored (uncompressed) deflate blocks: no dependency on zlib.
static bool write_png_rgb8(const char* path, const unsigned char* rgb, int w, int h) {
    crc_init();
    std::vector<unsigned char> raw;
    raw.reserve((size_t)h * (3 * (size_t)w + 1));
    for (int y = 0; y < h; ++y) {
        raw.push_back(0);
        raw.insert(raw.end(), rgb + (size_t)y * w * 3, rgb + (size_t)(y + 1) * w * 3);
    }
    std::vector<unsigned char> z;
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;
    for (size_t off = 0; off < raw.size() || off == 0;) {
        const size_t n = raw.size() - off < 65535 ? raw.size() - off : 65535;
        const bool last = off + n >= raw.size();
        z.push_back(last ? 1 : 0);
        z.push_back((unsigned char)(n & 0xff)); z.push_back((unsigned char)(n >> 8));
        z.push_back((unsigned char)(~n & 0xff)); z.push_back((unsigned char)((~n >> 8) & 0xff));
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
        for (size_t i = 0; i < n; ++i) { a = (a + raw[off + i]) % 65521u; b = (b + a) % 65521u; }
        off += n;
        if (last) break;
    }
    put_be32(z, (b << 16) | a);
    std::vector<unsigned char> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<unsigned char> ihdr;
    put_be32(ihdr, (uint32_t)w); put_be32(ihdr, (uint32_t)h);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    png_chunk(out, "IHDR", ihdr);
    png_chunk(out, "IDAT", z);
    png_chunk(out, "IEND", std::vector<unsigned char>());
    FILE* f = fopen(path, "wb");
    if (!f) return false;
    const bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
    fclose(f);
    return ok;
}
static bool write_ppm_rgb8(const char* path, const unsigned char* rgb, int w, int h) {
    FILE* f = fopen(path, "wb");
    if (!f) return false;
    fprintf(f, "P6\n%d %d\n255\n", w, h);
    const size_t n = (size_t)w * h * 3;
    const bool ok = fwrite(rgb, 1, n, f) == n;
    fclose(f);
    return ok;
}
static bool write_pfm(const char* path, const float* rgba, int w, int h) {
    FILE* f = fopen(path, "wb");
    if (!f) return false;
    fprintf(f, "PF\n%d %d\n-1.0\n", w, h);  // little endian, bottom row first
    std::vector<float> row((size_t)w * 3);
    bool ok = true;
    for (int y = h - 1; y >= 0 && ok; --y) {
        for (int x = 0; x < w; ++x)
            for (int c = 0; c < 3; ++c) row[3 * (size_t)x + c] = rgba[4 * ((size_t)y * w + x) + c];
        ok = fwrite(row.data(), sizeof(float), row.size(), f) == row.size();
    }
    fclose(f);
    return ok;
}
// Minimal OpenEXR 2.0 scanline file: channels B, G, R (fp32), no compression, increasing Y.
static bool write_exr(const char* path, const float* rgba, int w, int h) {
    std::vector<unsigned char> o;
    auto put32 = [&](uint32_t x) { for (int k = 0; k < 4; ++k) o.push_back((unsigned char)(x >> (8 * k))); };
    auto put64 = [&](uint64_t x) { for (int k = 0; k < 8; ++k) o.push_back((unsigned char)(x >> (8 * k))); };
    auto puts0 = [&](const char* s) { o.insert(o.end(), s, s + strlen(s) + 1); };
    auto putf = [&](float f) { uint32_t u; memcpy(&u, &f, 4); put32(u); };
    put32(20000630u);
    put32(2u);
    puts0("channels"); puts0("chlist");
    put32(3 * 18 + 1);
    for (const char* ch : {"B", "G", "R"}) { puts0(ch); put32(2u); o.push_back(0); o.push_back(0); o.push_back(0); o.push_back(0); put32(1); put32(1); }
    o.push_back(0);
    puts0("compression"); puts0("compression"); put32(1); o.push_back(0);
    puts0("dataWindow"); puts0("box2i"); put32(16); put32(0); put32(0); put32((uint32_t)(w - 1)); put32((uint32_t)(h - 1));
    puts0("displayWindow"); puts0("box2i"); put32(16); put32(0); put32(0); put32((uint32_t)(w - 1)); put32((uint32_t)(h - 1));
    puts0("lineOrder"); puts0("lineOrder"); put32(1); o.push_back(0);
    puts0("pixelAspectRatio"); puts0("float"); put32(4); putf(1.0f);
    puts0("screenWindowCenter"); puts0("v2f"); put32(8); putf(0.f); putf(0.f);
    puts0("screenWindowWidth"); puts0("float"); put32(4); putf(1.0f);
    o.push_back(0);
    const size_t table = o.size();
    const size_t line_bytes = (size_t)w * 3 * 4;
    uint64_t off = table + 8 * (uint64_t)h;
    for (int y = 0; y < h; ++y) { put64(off); off += 8 + line_bytes; }
    for (int y = 0; y < h; ++y) {
        put32((uint32_t)y);
        put32((uint32_t)line_bytes);
        for (int c = 2; c >= 0; --c)
            for (int x = 0; x < w; ++x) putf(rgba[4 * ((size_t)y * w + x) + c]);
    }
    FILE* f = fopen(path, "wb");
    if (!f) return false;
    const bool ok = fwrite(o.data(), 1, o.size(), f) == o.size();
    fclose(f);
    return ok;
}
static bool has_ext(const char* path, const char* ext) {
    const size_t n = strlen(path), m = strlen(ext);
    if (n < m) return false;
    for (size_t i = 0; i < m; ++i) {
        char a = path[n - m + i], b = ext[i];
        if (a >= 'A' && a <= 'Z') a = (char)(a - 'A' + 'a');
        if (a != b) return false;
    }
    return true;
}
int write_image_float(const char* path, const float* rgba, int w, int h) {
    if (!path || !rgba || w <= 0 || h <= 0) { set_error("ff_write_image: bad argument"); return FF_ERR_INVALID; }
    bool ok;
    if (has_ext(path, ".exr")) ok = write_exr(path, rgba, w, h);
    else if (has_ext(path, ".pfm")) ok = write_pfm(path, rgba, w, h);
    else if (has_ext(path, ".png") || has_ext(path, ".ppm")) {
        std::vector<unsigned char> rgb((size_t)w * h * 3);
        for (size_t i = 0; i < (size_t)w * h; ++i)
            for (int c = 0; c < 3; ++c) rgb[3 * i + c] = to_srgb8(rgba[4 * i + c]);
        ok = has_ext(path, ".png") ? write_png_rgb8(path, rgb.data(), w, h) : write_ppm_rgb8(path, rgb.data(), w, h);
    } else { set_error(std::string("ff_write_image: unknown extension: ") + path); return FF_ERR_UNSUPPORTED; }
    if (!ok) { set_error(std::string("ff_write_image: cannot write ") + path); return FF_ERR_IO; }
    return FF_OK;
}
int write_image_u8(const char* path, const unsigned char* rgb, int w, int h) {
    if (!path || !rgb || w <= 0 || h <= 0) { set_error("ff_write_image_u8: bad argument"); return FF_ERR_INVALID; }
    bool ok;
    if (has_ext(path, ".png")) ok = write_png_rgb8(path, rgb, w, h);
    else if (has_ext(path, ".ppm")) ok = write_ppm_rgb8(path, rgb, w, h);
    else { set_error(std::string("ff_write_image_u8: unknown extension: ") + path); return FF_ERR_UNSUPPORTED; }
    if (!ok) { set_error(std::string("ff_write_image_u8: cannot write ") + path); return FF_ERR_IO; }
    return FF_OK;
}

}  // namespace ff
