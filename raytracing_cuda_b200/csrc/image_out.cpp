/*
 * image_out.cpp -- headless replacement of the reference's display path.
 *
 * The reference maps a GL pixel-buffer object, renders into it and blits it to
 * the window (reference main.cpp:141-165, 176-213).  Here the framebuffer is a
 * plain device buffer; once copied to the host it is written as binary PPM or
 * as PNG (8-bit RGB, zlib deflate).  Pixels are the packed 0x00BBGGRR words of
 * reference kernel.cu:26-32 (bytes R,G,B,0 in memory), row 0 = top.
 *
 * rtb_read_png is the input side: the reference decodes its four sky maps with
 * stbi_load(path, &w, &h, &n, 4) (reference scene.cpp:626-632), i.e. to RGBA8 rows
 * top-down whatever the file holds; here a small decoder on zlib for what PNG
 * files of that kind are (8-bit grey / grey+alpha / RGB / RGBA / palette,
 * non-interlaced) does the same.
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

uint32_t get_be32(const unsigned char* p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3];
}

/* PNG row filters (RFC 2083 section 6): undo filter `type` on `row` in place; `prev` is the finished row above
 * (zeros for the first row), bpp the bytes per complete pixel */
bool unfilter_row(int type, unsigned char* row, const unsigned char* prev, size_t n, size_t bpp) {
    switch (type) {
    case 0: return true;
    case 1: for (size_t i = bpp; i < n; i++) row[i] = (unsigned char)(row[i] + row[i - bpp]); return true;
    case 2: for (size_t i = 0; i < n; i++) row[i] = (unsigned char)(row[i] + prev[i]); return true;
    case 3:
        for (size_t i = 0; i < n; i++) {
            const int left = i >= bpp ? row[i - bpp] : 0;
            row[i] = (unsigned char)(row[i] + ((left + prev[i]) >> 1));
        }
        return true;
    case 4:
        for (size_t i = 0; i < n; i++) {
            const int a = i >= bpp ? row[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
            const int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
            row[i] = (unsigned char)(row[i] + ((pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c)));
        }
        return true;
    default: return false;
    }
}

}  // namespace

extern "C" {

/* Decode a PNG file to RGBA8 rows, top row first (what stbi_load(..., 4) returns, reference scene.cpp:626-632).
 * out = NULL: only *w, *h are set (size query).  capacity = bytes available at out (>= 4 w h).
 * 0 on success; RT_ERR_ARG: unreadable / not a PNG / capacity too small; RT_ERR_UNSUPPORTED: 16-bit, sub-byte or
 * interlaced files. */
int rtb_read_png(const char* path, int* w, int* h, unsigned char* out, size_t capacity) {
    if (!path || !w || !h) return RT_ERR_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) return RT_ERR_ARG;
    std::vector<unsigned char> file;
    {
        unsigned char buf[1 << 16];
        size_t n;
        while ((n = fread(buf, 1, sizeof(buf), f)) > 0) file.insert(file.end(), buf, buf + n);
        fclose(f);
    }
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (file.size() < 8 + 25 || memcmp(file.data(), sig, 8) != 0) return RT_ERR_ARG;
    uint32_t iw = 0, ih = 0;
    int depth = 0, ctype = -1, interlace = 0;
    std::vector<unsigned char> idat, plte, trns;
    for (size_t at = 8; at + 12 <= file.size();) {
        const uint32_t len = get_be32(&file[at]);
        const unsigned char* type = &file[at + 4];
        const unsigned char* data = &file[at + 8];
        if (at + 12 + (size_t)len > file.size()) return RT_ERR_ARG;
        if (!memcmp(type, "IHDR", 4) && len >= 13) {
            iw = get_be32(data); ih = get_be32(data + 4);
            depth = data[8]; ctype = data[9]; interlace = data[12];
        } else if (!memcmp(type, "PLTE", 4)) plte.assign(data, data + len);
        else if (!memcmp(type, "tRNS", 4)) trns.assign(data, data + len);
        else if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), data, data + len);
        else if (!memcmp(type, "IEND", 4)) break;
        at += 12 + (size_t)len;
    }
    if (iw == 0 || ih == 0 || iw > (1u << 20) || ih > (1u << 20)) return RT_ERR_ARG;
    *w = (int)iw;
    *h = (int)ih;
    if (!out) return RT_OK;
    if (depth != 8 || interlace != 0) return RT_ERR_UNSUPPORTED;
    const size_t ch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (ch == 0 || (ctype == 3 && plte.size() < 3)) return RT_ERR_UNSUPPORTED;
    if (capacity < (size_t)iw * ih * 4) return RT_ERR_ARG;
    const size_t stride = (size_t)iw * ch;
    std::vector<unsigned char> raw((stride + 1) * ih);
    uLongf rawlen = (uLongf)raw.size();
    if (uncompress(raw.data(), &rawlen, idat.data(), (uLong)idat.size()) != Z_OK || rawlen != raw.size()) return RT_ERR_ARG;
    std::vector<unsigned char> zero(stride, 0);
    for (uint32_t y = 0; y < ih; y++) {
        unsigned char* row = &raw[(stride + 1) * y + 1];
        const unsigned char* prev = y ? &raw[(stride + 1) * (y - 1) + 1] : zero.data();
        if (!unfilter_row(raw[(stride + 1) * y], row, prev, stride, ch)) return RT_ERR_ARG;
        unsigned char* o = out + (size_t)y * iw * 4;
        for (uint32_t x = 0; x < iw; x++, o += 4) {
            const unsigned char* p = row + (size_t)x * ch;
            switch (ctype) {
            case 0: o[0] = o[1] = o[2] = p[0]; o[3] = 255; break;
            case 2: o[0] = p[0]; o[1] = p[1]; o[2] = p[2]; o[3] = 255; break;
            case 3: {
                const size_t k = (size_t)p[0] * 3 + 2 < plte.size() ? p[0] : 0;
                o[0] = plte[3 * k]; o[1] = plte[3 * k + 1]; o[2] = plte[3 * k + 2];
                o[3] = k < trns.size() ? trns[k] : 255;
                break;
            }
            case 4: o[0] = o[1] = o[2] = p[0]; o[3] = p[1]; break;
            default: o[0] = p[0]; o[1] = p[1]; o[2] = p[2]; o[3] = p[3]; break;
            }
        }
    }
    return RT_OK;
}

/* frame `index` of a sequence dump: <prefix><index, six digits>.png (format 1) or .ppm (format 0) */
int rtb_write_frame(const char* prefix, int index, int format, const unsigned int* pixels, int w, int h) {
    if (!prefix || index < 0 || (format != 0 && format != 1)) return RT_ERR_ARG;
    char path[4096];
    const int n = snprintf(path, sizeof(path), "%s%06d.%s", prefix, index, format ? "png" : "ppm");
    if (n <= 0 || (size_t)n >= sizeof(path)) return RT_ERR_ARG;
    return format ? rtb_write_png(path, pixels, w, h) : rtb_write_ppm(path, pixels, w, h);
}

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
