// rt_image.cpp — output stage after the path: the reference only presents
// its three Rgba8Unorm textures to a window through main_fs (shader/src/lib.rs:368-391,
// src/renderer.rs:778-818) and has no file output (PLAN.md lists it as future work); headless,
// the composite / accumulation buffers read back through the C ABI are written as image files:
// binary PPM and uncompressed-deflate PNG for rgba8, PFM for the fp32 accumulation buffer.
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "../../include/rtb200_host.h"

namespace {
uint32_t crc32_update(uint32_t crc, const uint8_t* p, size_t n) {
    static uint32_t T[256];
    static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; i++) { uint32_t c = i; for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1; T[i] = c; }
        init = true;
    }
    for (size_t i = 0; i < n; i++) crc = T[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    return crc;
}
void be32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
void chunk(std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data) {
    be32(out, (uint32_t)data.size());
    size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    be32(out, crc32_update(0xFFFFFFFFu, out.data() + start, out.size() - start) ^ 0xFFFFFFFFu);
}
int write_all(const char* path, const void* p, size_t n) {
    FILE* f = fopen(path, "wb");
    if (!f) return -1;
    size_t w = fwrite(p, 1, n, f);
    return (fclose(f) == 0 && w == n) ? 0 : -1;
}
} // namespace

extern "C" {

int rt_write_ppm_rgba8(const char* path, const uint8_t* rgba, uint32_t width, uint32_t height) {
    if (!path || !rgba || !width || !height) return -1;
    char hdr[64];
    int hl = snprintf(hdr, sizeof hdr, "P6\n%u %u\n255\n", width, height);
    std::vector<uint8_t> buf((size_t)hl + (size_t)width * height * 3);
    memcpy(buf.data(), hdr, (size_t)hl);
    uint8_t* d = buf.data() + hl;
    for (size_t i = 0; i < (size_t)width * height; i++) { d[i * 3] = rgba[i * 4]; d[i * 3 + 1] = rgba[i * 4 + 1]; d[i * 3 + 2] = rgba[i * 4 + 2]; }
    return write_all(path, buf.data(), buf.size());
}

int rt_write_png_rgba8(const char* path, const uint8_t* rgba, uint32_t width, uint32_t height) {
    if (!path || !rgba || !width || !height) return -1;
    std::vector<uint8_t> raw; // filter byte 0 + scanline
    raw.reserve(((size_t)width * 4 + 1) * height);
    for (uint32_t y = 0; y < height; y++) { raw.push_back(0); raw.insert(raw.end(), rgba + (size_t)y * width * 4, rgba + (size_t)(y + 1) * width * 4); }
    std::vector<uint8_t> z;
    z.push_back(0x78); z.push_back(0x01); // zlib header, stored deflate blocks
    uint32_t a = 1, b = 0;
    for (size_t off = 0; off < raw.size() || off == 0;) {
        size_t n = raw.size() - off < 65535 ? raw.size() - off : 65535;
        z.push_back(off + n >= raw.size() ? 1 : 0);
        z.push_back(n & 0xFF); z.push_back(n >> 8); z.push_back(~n & 0xFF); z.push_back((~n >> 8) & 0xFF);
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
        for (size_t i = 0; i < n; i++) { a = (a + raw[off + i]) % 65521u; b = (b + a) % 65521u; }
        off += n;
        if (n == 0) break;
    }
    be32(z, (b << 16) | a);
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'}, ihdr;
    be32(ihdr, width); be32(ihdr, height);
    ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0); // 8-bit RGBA
    chunk(out, "IHDR", ihdr);
    chunk(out, "IDAT", z);
    chunk(out, "IEND", {});
    return write_all(path, out.data(), out.size());
}

/* accum: W*H float4 (rgb sum, sample count); writes rgb / count as a little-endian PFM (bottom row first) */
int rt_write_pfm_accum(const char* path, const float* accum, uint32_t width, uint32_t height) {
    if (!path || !accum || !width || !height) return -1;
    char hdr[64];
    int hl = snprintf(hdr, sizeof hdr, "PF\n%u %u\n-1.0\n", width, height);
    std::vector<uint8_t> buf((size_t)hl + (size_t)width * height * 12);
    memcpy(buf.data(), hdr, (size_t)hl);
    float* d = (float*)(buf.data() + hl);
    for (uint32_t y = 0; y < height; y++)
        for (uint32_t x = 0; x < width; x++) {
            const float* a = accum + ((size_t)(height - 1 - y) * width + x) * 4;
            float w = a[3] > 0.0f ? a[3] : 1.0f;
            float* o = d + ((size_t)y * width + x) * 3;
            o[0] = a[0] / w; o[1] = a[1] / w; o[2] = a[2] / w;
        }
    return write_all(path, buf.data(), buf.size());
}

} // extern "C"
