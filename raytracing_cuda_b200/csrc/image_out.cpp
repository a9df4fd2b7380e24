/*
 * image_out.cpp -- headless replacement of the reference's display path.
 *
 * The reference maps a GL pixel-buffer object, renders into it and blits it to
 * the window (reference main.cpp:141-165, 176-213).  Here the framebuffer is a
 * plain device buffer; once copied to the host it is written as binary PPM or
 * as PNG (8-bit RGB, zlib deflate).  Pixels are the packed 0x00BBGGRR words of
 * reference kernel.cu:26-32 (bytes R,G,B,0 in memory), row 0 = top.
 */
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>
#include <zlib.h>

#include "../../include/rtb_api.h"

namespace {

void rgb_rows(const unsigned int* px, int w, int h, int filter_byte, std::vector<unsigned char>& out) {
    const size_t stride = (size_t)w * 3 + (filter_byte ? 1 : 0);
    out.resize(stride * h);
    for (int y = 0; y < h; y++) {
        unsigned char* row = out.data() + stride * y;
        if (filter_byte) *row++ = 0;
        for (int x = 0; x < w; x++) {
            unsigned int c = px[(size_t)y * w + x];
            row[3 * x + 0] = (unsigned char)(c & 255);
            row[3 * x + 1] = (unsigned char)((c >> 8) & 255);
            row[3 * x + 2] = (unsigned char)((c >> 16) & 255);
        }
    }
}

void put_be32(unsigned char* p, uint32_t v) {
    p[0] = (unsigned char)(v >> 24); p[1] = (unsigned char)(v >> 16);
    p[2] = (unsigned char)(v >> 8);  p[3] = (unsigned char)v;
}

bool write_chunk(FILE* f, const char* type, const unsigned char* data, uint32_t len) {
    unsigned char hdr[8];
    put_be32(hdr, len);
    memcpy(hdr + 4, type, 4);
    uint32_t crc = (uint32_t)crc32(0L, hdr + 4, 4);
    if (len) crc = (uint32_t)crc32(crc, data, len);
    unsigned char tail[4];
    put_be32(tail, crc);
    return fwrite(hdr, 1, 8, f) == 8 && (len == 0 || fwrite(data, 1, len, f) == len) &&
           fwrite(tail, 1, 4, f) == 4;
}

}  // namespace

extern "C" {

int rtb_write_ppm(const char* path, const unsigned int* pixels, int w, int h) {
    if (!path || !pixels || w <= 0 || h <= 0) return RT_ERR_ARG;
    FILE* f = fopen(path, "wb");
    if (!f) return RT_ERR_ARG;
    std::vector<unsigned char> rows;
    rgb_rows(pixels, w, h, 0, rows);
    fprintf(f, "P6\n%d %d\n255\n", w, h);
    bool ok = fwrite(rows.data(), 1, rows.size(), f) == rows.size();
    fclose(f);
    return ok ? RT_OK : RT_ERR_ARG;
}

int rtb_write_png(const char* path, const unsigned int* pixels, int w, int h) {
    if (!path || !pixels || w <= 0 || h <= 0) return RT_ERR_ARG;
    std::vector<unsigned char> rows;
    rgb_rows(pixels, w, h, 1, rows);
    uLongf clen = compressBound((uLong)rows.size());
    std::vector<unsigned char> comp(clen);
    if (compress2(comp.data(), &clen, rows.data(), (uLong)rows.size(), 6) != Z_OK) return RT_ERR_ARG;

    FILE* f = fopen(path, "wb");
    if (!f) return RT_ERR_ARG;
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    unsigned char ihdr[13];
    put_be32(ihdr, (uint32_t)w);
    put_be32(ihdr + 4, (uint32_t)h);
    ihdr[8] = 8;    /* bit depth */
    ihdr[9] = 2;    /* colour type RGB */
    ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;
    bool ok = fwrite(sig, 1, 8, f) == 8 && write_chunk(f, "IHDR", ihdr, 13) &&
              write_chunk(f, "IDAT", comp.data(), (uint32_t)clen) && write_chunk(f, "IEND", nullptr, 0);
    fclose(f);
    return ok ? RT_OK : RT_ERR_ARG;
}

}  // extern "C"
