// image_io.cpp -- PNG writer for the film (host only).  Replaces the reference's stb_image_write call behind "Save PNG"
// (include/ui/ui_windows.h:205-208: stbi_flip_vertically_on_write(1) + stbi_write_png(path, w, h, 3, h_image, w*3)).
// The stream is standard PNG (8-bit RGB, filter 0 on every scanline) with the zlib payload in stored (uncompressed)
// deflate blocks: every decoder reads it, the file is w*h*3 bytes plus ~0.01 % framing, and writing it costs one pass.
#include "image_io.h"

#include <cstdio>
#include <cstring>

namespace ptx {

namespace {

uint32_t crc_table[256];
bool crc_ready = false;

void crc_init() {
    for (uint32_t n = 0; n < 256; n++) {
        uint32_t c = n;
        for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
        crc_table[n] = c;
    }
    crc_ready = true;
}

uint32_t crc32_update(uint32_t crc, const uint8_t* p, size_t n) {
    if (!crc_ready) crc_init();
    for (size_t i = 0; i < n; i++) crc = crc_table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    return crc;
}

void put_be32(std::vector<uint8_t>& out, uint32_t v) {
    out.push_back((uint8_t)(v >> 24)); out.push_back((uint8_t)(v >> 16)); out.push_back((uint8_t)(v >> 8)); out.push_back((uint8_t)v);
}

void put_chunk(std::vector<uint8_t>& out, const char type[4], const uint8_t* data, size_t n) {
    put_be32(out, (uint32_t)n);
    size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    if (n) out.insert(out.end(), data, data + n);
    uint32_t crc = crc32_update(0xFFFFFFFFu, out.data() + start, n + 4) ^ 0xFFFFFFFFu;
    put_be32(out, crc);
}

}  // namespace

void encode_png_rgb8(int w, int h, const uint8_t* rgb, bool flip_y, std::vector<uint8_t>& out) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    out.clear();
    out.insert(out.end(), sig, sig + 8);
    uint8_t ihdr[13];
    ihdr[0] = (uint8_t)(w >> 24); ihdr[1] = (uint8_t)(w >> 16); ihdr[2] = (uint8_t)(w >> 8); ihdr[3] = (uint8_t)w;
    ihdr[4] = (uint8_t)(h >> 24); ihdr[5] = (uint8_t)(h >> 16); ihdr[6] = (uint8_t)(h >> 8); ihdr[7] = (uint8_t)h;
    ihdr[8] = 8; ihdr[9] = 2; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;          // 8 bits, colour type 2 (RGB), deflate, adaptive, no interlace
    put_chunk(out, "IHDR", ihdr, 13);

    // raw scanlines: filter byte 0 + w*3 bytes
    const size_t row = (size_t)w * 3, raw_n = (row + 1) * (size_t)h;
    std::vector<uint8_t> raw(raw_n);
    for (int y = 0; y < h; y++) {
        const uint8_t* src = rgb + (size_t)(flip_y ? h - 1 - y : y) * row;
        raw[(size_t)y * (row + 1)] = 0;
        memcpy(raw.data() + (size_t)y * (row + 1) + 1, src, row);
    }
    // zlib container, stored blocks of <= 65535 bytes
    std::vector<uint8_t> z;
    z.reserve(raw_n + raw_n / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;                                                     // Adler-32 of the raw data
    size_t pos = 0;
    do {
        size_t n = raw_n - pos < 65535 ? raw_n - pos : 65535;
        z.push_back(pos + n == raw_n ? 1 : 0);                                 // BFINAL, BTYPE = 00
        z.push_back((uint8_t)(n & 0xFF)); z.push_back((uint8_t)(n >> 8));
        z.push_back((uint8_t)(~n & 0xFF)); z.push_back((uint8_t)((~n >> 8) & 0xFF));
        z.insert(z.end(), raw.begin() + (long)pos, raw.begin() + (long)(pos + n));
        for (size_t i = pos; i < pos + n;) {                                   // modulo deferred over runs of 5552 bytes
            size_t run = pos + n - i < 5552 ? pos + n - i : 5552;
            for (size_t k = 0; k < run; k++) { a += raw[i + k]; b += a; }
            a %= 65521u; b %= 65521u; i += run;
        }
        pos += n;
    } while (pos < raw_n);
    put_be32(z, (b << 16) | a);
    put_chunk(out, "IDAT", z.data(), z.size());
    put_chunk(out, "IEND", nullptr, 0);
}

bool write_png_rgb8(const char* path, int w, int h, const uint8_t* rgb, bool flip_y) {
    std::vector<uint8_t> png;
    encode_png_rgb8(w, h, rgb, flip_y, png);
    FILE* fp = fopen(path, "wb");
    if (!fp) return false;
    bool ok = fwrite(png.data(), 1, png.size(), fp) == png.size();
    return fclose(fp) == 0 && ok;
}

}  // namespace ptx
