// rt_image_io.cpp — host image output / input without stb (the reference links stb_image_write, color.cuh:9-12).
//   rt_write_ppm : P3 text, "r g b\n" per pixel, rows top to bottom  (output_image, gpu-version/main.cu:359-372)
//   rt_write_png : 8-bit RGB, zlib-compressed, filter 0               (write_image, gpu-version/color.cuh:15-35)
//   rt_read_png  : 8-bit grey / grey+alpha / RGB / RGBA / palette, non-interlaced (image textures)
#include <zlib.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "rt_ctx.h"

int rt_png_decode_file(const char *path, std::vector<uint8_t> &rgb, int &w, int &h, std::string &err);

static int rt_write_ppm_impl(const char *path, const unsigned char *rgb8, int width, int height) {
    if (!path || !rgb8 || width < 1 || height < 1) return RT_ERR_INVALID;
    FILE *f = fopen(path, "w");
    if (!f) return RT_ERR_IO;
    fprintf(f, "P3\n%d %d\n255\n", width, height);
    std::string buf;
    buf.reserve((size_t)width * 12 + 16);
    char tmp[16];
    for (int y = 0; y < height; ++y) {
        buf.clear();
        for (int x = 0; x < width; ++x) {
            const unsigned char *p = rgb8 + 3 * ((size_t)y * width + x);
            int n = snprintf(tmp, sizeof tmp, "%d %d %d\n", p[0], p[1], p[2]);
            buf.append(tmp, n);
        }
        if (fwrite(buf.data(), 1, buf.size(), f) != buf.size()) { fclose(f); return RT_ERR_IO; }
    }
    return fclose(f) == 0 ? RT_OK : RT_ERR_IO;
}
extern "C" int rt_write_ppm(const char *path, const unsigned char *rgb8, int width, int height) {
    return rt_guard(nullptr, "rt_write_ppm", [&] { return rt_write_ppm_impl(path, rgb8, width, height); });
}

static void put32(std::vector<uint8_t> &v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
static void chunk(std::vector<uint8_t> &out, const char *tag, const uint8_t *data, size_t n) {
    put32(out, (uint32_t)n);
    size_t start = out.size();
    out.insert(out.end(), tag, tag + 4);
    if (n) out.insert(out.end(), data, data + n);
    put32(out, (uint32_t)crc32(0, out.data() + start, (uInt)(n + 4)));
}

static int rt_write_png_impl(const char *path, const unsigned char *rgb8, int width, int height) {
    if (!path || !rgb8 || width < 1 || height < 1) return RT_ERR_INVALID;
    const size_t stride = (size_t)width * 3;
    std::vector<uint8_t> raw((stride + 1) * height);
    for (int y = 0; y < height; ++y) { raw[(stride + 1) * y] = 0; memcpy(&raw[(stride + 1) * y + 1], rgb8 + stride * y, stride); }
    uLongf clen = compressBound((uLong)raw.size());
    std::vector<uint8_t> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), (uLong)raw.size(), 1) != Z_OK) return RT_ERR_IO;
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<uint8_t> ihdr;
    put32(ihdr, (uint32_t)width); put32(ihdr, (uint32_t)height);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    chunk(out, "IHDR", ihdr.data(), ihdr.size());
    chunk(out, "IDAT", comp.data(), clen);
    chunk(out, "IEND", nullptr, 0);
    FILE *f = fopen(path, "wb");
    if (!f) return RT_ERR_IO;
    bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
    return (fclose(f) == 0 && ok) ? RT_OK : RT_ERR_IO;
}
extern "C" int rt_write_png(const char *path, const unsigned char *rgb8, int width, int height) {
    return rt_guard(nullptr, "rt_write_png", [&] { return rt_write_png_impl(path, rgb8, width, height); });
}

static inline int paeth(int a, int b, int c) {
    int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

int rt_png_decode_file(const char *path, std::vector<uint8_t> &rgb, int &w, int &h, std::string &err) {
    FILE *f = fopen(path, "rb");
    if (!f) { err = std::string("cannot open image ") + path; return RT_ERR_IO; }
    std::vector<uint8_t> file;
    uint8_t buf[65536]; size_t n;
    while ((n = fread(buf, 1, sizeof buf, f)) > 0) file.insert(file.end(), buf, buf + n);
    fclose(f);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (file.size() < 8 || memcmp(file.data(), sig, 8)) { err = std::string("not a PNG file: ") + path; return RT_ERR_PARSE; }
    size_t pos = 8;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, plte;
    w = h = 0;
    while (pos + 12 <= file.size()) {
        uint32_t len = (file[pos] << 24) | (file[pos + 1] << 16) | (file[pos + 2] << 8) | file[pos + 3];
        const char *tag = (const char *)&file[pos + 4];
        const uint8_t *data = &file[pos + 8];
        if (pos + 12 + len > file.size()) { err = "truncated PNG"; return RT_ERR_PARSE; }
        if (!memcmp(tag, "IHDR", 4) && len >= 13) {
            w = (data[0] << 24) | (data[1] << 16) | (data[2] << 8) | data[3];
            h = (data[4] << 24) | (data[5] << 16) | (data[6] << 8) | data[7];
            depth = data[8]; ctype = data[9]; interlace = data[12];
        } else if (!memcmp(tag, "PLTE", 4)) plte.assign(data, data + len);
        else if (!memcmp(tag, "IDAT", 4)) idat.insert(idat.end(), data, data + len);
        else if (!memcmp(tag, "IEND", 4)) break;
        pos += 12 + len;
    }
    if (w < 1 || h < 1 || depth != 8 || interlace != 0) { err = "unsupported PNG (need 8-bit, non-interlaced)"; return RT_ERR_PARSE; }
    int ch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!ch) { err = "unsupported PNG colour type"; return RT_ERR_PARSE; }
    const size_t stride = (size_t)w * ch;
    // the declared size must be plausible for the compressed data before anything is allocated for it (deflate expands at most
    // ~1032 : 1): a header that claims 2^31 x 2^31 pixels over a few bytes of IDAT is rejected here, not by bad_alloc
    if ((w > (1 << 20)) || (h > (1 << 20)) || (stride + 1) * (size_t)h > (idat.size() + 64) * 1100) { err = "PNG header declares more pixels than its data can hold"; return RT_ERR_PARSE; }
    std::vector<uint8_t> raw((stride + 1) * h);
    uLongf rlen = (uLongf)raw.size();
    if (uncompress(raw.data(), &rlen, idat.data(), (uLong)idat.size()) != Z_OK || rlen != raw.size()) { err = "PNG inflate failed"; return RT_ERR_PARSE; }
    std::vector<uint8_t> img(stride * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t *src = &raw[(stride + 1) * y + 1];
        const int ft = raw[(stride + 1) * y];
        uint8_t *dst = &img[stride * y];
        const uint8_t *up = y ? &img[stride * (y - 1)] : nullptr;
        for (size_t x = 0; x < stride; ++x) {
            int a = x >= (size_t)ch ? dst[x - ch] : 0, b = up ? up[x] : 0, c = (up && x >= (size_t)ch) ? up[x - ch] : 0;
            int v = src[x];
            switch (ft) {
            case 0: break; case 1: v += a; break; case 2: v += b; break; case 3: v += (a + b) >> 1; break;
            case 4: v += paeth(a, b, c); break;
            default: err = "bad PNG filter"; return RT_ERR_PARSE;
            }
            dst[x] = (uint8_t)v;
        }
    }
    rgb.resize((size_t)w * h * 3);
    for (size_t i = 0; i < (size_t)w * h; ++i) {
        const uint8_t *p = &img[i * ch];
        if (ctype == 2 || ctype == 6) { rgb[3 * i] = p[0]; rgb[3 * i + 1] = p[1]; rgb[3 * i + 2] = p[2]; }
        else if (ctype == 3) {
            size_t k = (size_t)p[0] * 3;
            if (k + 3 > plte.size()) { err = "PNG palette index out of range"; return RT_ERR_PARSE; }
            rgb[3 * i] = plte[k]; rgb[3 * i + 1] = plte[k + 1]; rgb[3 * i + 2] = plte[k + 2];
        } else { rgb[3 * i] = rgb[3 * i + 1] = rgb[3 * i + 2] = p[0]; }
    }
    return RT_OK;
}

static int rt_read_png_impl(const char *path, unsigned char **rgb8, int *width, int *height) {
    if (!path || !rgb8 || !width || !height) return RT_ERR_INVALID;
    std::vector<uint8_t> rgb; int w, h; std::string err;
    int rc = rt_png_decode_file(path, rgb, w, h, err);
    if (rc) return rc;
    unsigned char *out = (unsigned char *)malloc(rgb.size());
    if (!out) return RT_ERR_IO;
    memcpy(out, rgb.data(), rgb.size());
    *rgb8 = out; *width = w; *height = h;
    return RT_OK;
}
extern "C" int rt_read_png(const char *path, unsigned char **rgb8, int *width, int *height) {
    return rt_guard(nullptr, "rt_read_png", [&] { return rt_read_png_impl(path, rgb8, width, height); });
}
