// BMP / PNG -> 8-bit gray, reproducing what cv::imread(path, IMREAD_GRAYSCALE) returns for the
// frame files of a PICO run (RawParser.cpp:39, ZipParser.cpp:222).  Written against the file
// format specifications; zlib provides inflate.  Colour-to-gray weights:
//   BMP (OpenCV's own decoder): (B*1868 + G*9617 + R*4899 + 8192) >> 14
//   PNG (libpng rgb_to_gray set up by OpenCV with 0.299/0.587): (R*9797 + G*19234 + B*3737 + 16384) >> 15
// Both are the identity on gray-valued pixels, which is what detector frames hold.
#include "ImageCodec.hpp"

#include <stdio.h>
#include <string.h>
#include <zlib.h>

namespace abubcodec {

namespace {

inline uint32_t rd32le(const uint8_t* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
inline uint16_t rd16le(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }
inline uint32_t rd32be(const uint8_t* p) { return ((uint32_t)p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3]; }

inline uint8_t bgr2gray(int b, int g, int r) { return (uint8_t)((b * 1868 + g * 9617 + r * 4899 + (1 << 13)) >> 14); }
inline uint8_t png_rgb2gray(int r, int g, int b) {
    if (r == g && g == b) return (uint8_t)r;
    return (uint8_t)((r * 9797 + g * 19234 + b * 3737 + 16384) >> 15);
}

struct Sink {                 // where decoded rows go: an Image8 or caller memory
    uint8_t* px = nullptr;
    int rows = 0, cols = 0;
};

// ---------------------------------------------------------------- BMP
struct BmpInfo {
    int width = 0, height = 0, bpp = 0;
    bool top_down = false;
    uint32_t data_off = 0, pal_off = 0, pal_n = 0, pal_stride = 4;
};

bool bmp_header(const uint8_t* b, size_t n, BmpInfo& h) {
    if (n < 26 || b[0] != 'B' || b[1] != 'M') return false;
    h.data_off = rd32le(b + 10);
    const uint32_t dib = rd32le(b + 14);
    if (dib == 12) {                                   // BITMAPCOREHEADER
        h.width = rd16le(b + 18); h.height = rd16le(b + 20); h.bpp = rd16le(b + 24);
        h.pal_stride = 3;
    } else {
        if (n < 54 || dib < 40) return false;
        h.width = (int32_t)rd32le(b + 18);
        int32_t hh = (int32_t)rd32le(b + 22);
        h.top_down = hh < 0;
        h.height = hh < 0 ? -hh : hh;
        h.bpp = rd16le(b + 28);
        const uint32_t comp = rd32le(b + 30);
        if (!(comp == 0 || (comp == 3 && h.bpp == 32))) return false;       // RLE etc. not produced by the DAQ
        h.pal_n = rd32le(b + 46);
    }
    if (h.width <= 0 || h.height <= 0) return false;
    if (!(h.bpp == 1 || h.bpp == 4 || h.bpp == 8 || h.bpp == 24 || h.bpp == 32)) return false;
    h.pal_off = 14 + dib;
    if (h.bpp <= 8 && h.pal_n == 0) h.pal_n = 1u << h.bpp;
    if (h.bpp <= 8 && h.pal_n > 256) return false;
    return true;
}

bool bmp_decode(const uint8_t* b, size_t n, const BmpInfo& h, Sink& s) {
    const size_t stride = (((size_t)h.width * h.bpp + 31) / 32) * 4;
    if ((size_t)h.data_off + stride * h.height > n) return false;
    uint8_t gray_pal[256] = {0};
    if (h.bpp <= 8) {
        if ((size_t)h.pal_off + (size_t)h.pal_n * h.pal_stride > n) return false;
        for (uint32_t i = 0; i < h.pal_n; i++) {
            const uint8_t* e = b + h.pal_off + i * h.pal_stride;
            gray_pal[i] = bgr2gray(e[0], e[1], e[2]);
        }
    }
    bool identity = h.bpp == 8;                         // the usual greyscale palette: rows are copied as they are
    for (int i = 0; identity && i < 256; i++) identity = i < (int)h.pal_n ? gray_pal[i] == i : false;
    for (int y = 0; y < h.height; y++) {
        const uint8_t* src = b + h.data_off + stride * (h.top_down ? y : h.height - 1 - y);
        uint8_t* dst = s.px + (size_t)y * s.cols;
        if (identity) { memcpy(dst, src, (size_t)h.width); continue; }
        switch (h.bpp) {
            case 8:  for (int x = 0; x < h.width; x++) dst[x] = gray_pal[src[x]]; break;
            case 4:  for (int x = 0; x < h.width; x++) dst[x] = gray_pal[(src[x >> 1] >> ((x & 1) ? 0 : 4)) & 15]; break;
            case 1:  for (int x = 0; x < h.width; x++) dst[x] = gray_pal[(src[x >> 3] >> (7 - (x & 7))) & 1]; break;
            case 24: for (int x = 0; x < h.width; x++) dst[x] = bgr2gray(src[3 * x], src[3 * x + 1], src[3 * x + 2]); break;
            case 32: for (int x = 0; x < h.width; x++) dst[x] = bgr2gray(src[4 * x], src[4 * x + 1], src[4 * x + 2]); break;
        }
    }
    return true;
}

// ---------------------------------------------------------------- PNG
const uint8_t kPngSig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};

struct PngInfo {
    int width = 0, height = 0, depth = 0, ctype = 0, interlace = 0;
    int channels() const { return ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4; }
};

// libpng refuses a file whose critical chunk fails its CRC (cv2.imread then returns an empty Mat); same here, which also
// makes the compressed stream trustworthy for the device decoder (K9 does not compute the zlib Adler-32).
inline bool png_chunk_crc_ok(const uint8_t* chunk, uint32_t len) {
    return (uint32_t)crc32(crc32(0L, Z_NULL, 0), chunk + 4, len + 4) == rd32be(chunk + 8 + len);
}

bool png_header(const uint8_t* b, size_t n, PngInfo& h) {
    if (n < 33 || memcmp(b, kPngSig, 8) != 0) return false;
    if (rd32be(b + 8) != 13 || memcmp(b + 12, "IHDR", 4) != 0) return false;
    h.width = (int)rd32be(b + 16); h.height = (int)rd32be(b + 20);
    h.depth = b[24]; h.ctype = b[25]; h.interlace = b[28];
    if (h.width <= 0 || h.height <= 0 || b[26] != 0 || b[27] != 0) return false;
    if (!(h.ctype == 0 || h.ctype == 2 || h.ctype == 3 || h.ctype == 4 || h.ctype == 6)) return false;
    if (!(h.depth == 1 || h.depth == 2 || h.depth == 4 || h.depth == 8 || h.depth == 16)) return false;
    if ((h.ctype == 2 || h.ctype == 4 || h.ctype == 6) && h.depth < 8) return false;
    if (h.ctype == 3 && h.depth == 16) return false;
    return true;
}

inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

bool png_decode(const uint8_t* b, size_t n, const PngInfo& h, Sink& s) {
    if (h.interlace != 0) return false;                 // Adam7 is not produced by the DAQ
    // gather IDAT + PLTE
    std::vector<uint8_t> z;
    uint8_t pal[256][3];
    memset(pal, 0, sizeof(pal));
    size_t pos = 8;
    bool end = false;
    while (!end && pos + 12 <= n) {
        const uint32_t len = rd32be(b + pos);
        const uint8_t* type = b + pos + 4;
        if (pos + 12 + (size_t)len > n) return false;
        const uint8_t* d = b + pos + 8;
        if (!memcmp(type, "IDAT", 4)) { if (!png_chunk_crc_ok(b + pos, len)) return false; z.insert(z.end(), d, d + len); }
        else if (!memcmp(type, "PLTE", 4)) { for (uint32_t i = 0; i < len / 3 && i < 256; i++) memcpy(pal[i], d + 3 * i, 3); }
        else if (!memcmp(type, "IEND", 4)) end = true;
        pos += 12 + (size_t)len;
    }
    if (z.empty()) return false;
    const int ch = h.channels();
    const size_t bits_pp = (size_t)ch * h.depth;
    const size_t bpp = bits_pp >= 8 ? bits_pp / 8 : 1;                    // filter distance
    const size_t rowbytes = ((size_t)h.width * bits_pp + 7) / 8;
    std::vector<uint8_t> raw((rowbytes + 1) * (size_t)h.height);
    uLongf outlen = (uLongf)raw.size();
    if (uncompress(raw.data(), &outlen, z.data(), (uLong)z.size()) != Z_OK || outlen != raw.size()) return false;
    std::vector<uint8_t> prev(rowbytes, 0);
    for (int y = 0; y < h.height; y++) {
        uint8_t* line = raw.data() + (rowbytes + 1) * (size_t)y;
        const int ft = line[0];
        uint8_t* cur = line + 1;
        switch (ft) {
            case 0: break;
            case 1: for (size_t i = bpp; i < rowbytes; i++) cur[i] = (uint8_t)(cur[i] + cur[i - bpp]); break;
            case 2: for (size_t i = 0; i < rowbytes; i++) cur[i] = (uint8_t)(cur[i] + prev[i]); break;
            case 3:
                for (size_t i = 0; i < rowbytes; i++)
                    cur[i] = (uint8_t)(cur[i] + (((i >= bpp ? cur[i - bpp] : 0) + prev[i]) >> 1));
                break;
            case 4:
                for (size_t i = 0; i < rowbytes; i++)
                    cur[i] = (uint8_t)(cur[i] + paeth(i >= bpp ? cur[i - bpp] : 0, prev[i], i >= bpp ? prev[i - bpp] : 0));
                break;
            default: return false;
        }
        memcpy(prev.data(), cur, rowbytes);
        uint8_t* dst = s.px + (size_t)y * s.cols;
        const int step = h.depth == 16 ? 2 : 1;          // 16-bit samples: keep the high byte (png_set_strip_16)
        if (h.ctype == 0 || h.ctype == 4) {
            if (h.depth >= 8) {
                for (int x = 0; x < h.width; x++) dst[x] = cur[(size_t)x * ch * step];
            } else {
                const int mul = h.depth == 1 ? 255 : h.depth == 2 ? 85 : 17;
                const int per = 8 / h.depth, mask = (1 << h.depth) - 1;
                for (int x = 0; x < h.width; x++)
                    dst[x] = (uint8_t)(((cur[x / per] >> ((per - 1 - x % per) * h.depth)) & mask) * mul);
            }
        } else if (h.ctype == 3) {
            const int per = 8 / h.depth, mask = (1 << h.depth) - 1;
            for (int x = 0; x < h.width; x++) {
                const int idx = h.depth == 8 ? cur[x] : (cur[x / per] >> ((per - 1 - x % per) * h.depth)) & mask;
                dst[x] = png_rgb2gray(pal[idx][0], pal[idx][1], pal[idx][2]);
            }
        } else {
            for (int x = 0; x < h.width; x++) {
                const uint8_t* p = cur + (size_t)x * ch * step;
                dst[x] = png_rgb2gray(p[0], p[step], p[2 * step]);
            }
        }
    }
    return true;
}

bool decode_to(const uint8_t* buf, size_t n, Sink& s, Image8* own) {
    BmpInfo bh;
    PngInfo ph;
    int rows, cols;
    bool is_bmp = false;
    if (bmp_header(buf, n, bh)) { rows = bh.height; cols = bh.width; is_bmp = true; }
    else if (png_header(buf, n, ph)) { rows = ph.height; cols = ph.width; }
    else return false;
    if (own) { own->create(rows, cols); s.px = own->data.data(); s.rows = rows; s.cols = cols; }
    else if (rows != s.rows || cols != s.cols) return false;
    const bool ok = is_bmp ? bmp_decode(buf, n, bh, s) : png_decode(buf, n, ph, s);
    if (!ok && own) own->release();
    return ok;
}

}  // namespace

bool decode(const uint8_t* buf, size_t n, Image8& img) {
    Sink s;
    img.release();
    return decode_to(buf, n, s, &img);
}

bool decode_into(const uint8_t* buf, size_t n, uint8_t* dst, int rows, int cols) {
    Sink s;
    s.px = dst; s.rows = rows; s.cols = cols;
    return decode_to(buf, n, s, nullptr);
}

bool peek_size(const uint8_t* buf, size_t n, int* rows, int* cols) {
    BmpInfo bh;
    PngInfo ph;
    if (bmp_header(buf, n, bh)) { *rows = bh.height; *cols = bh.width; return true; }
    if (png_header(buf, n, ph)) { *rows = ph.height; *cols = ph.width; return true; }
    return false;
}

size_t png_gray8_stream(const uint8_t* b, size_t n, int rows, int cols, uint8_t* dst, size_t cap) {
    PngInfo h;
    if (!png_header(b, n, h)) return 0;
    if (h.ctype != 0 || h.depth != 8 || h.interlace != 0 || h.width != cols || h.height != rows) return 0;
    size_t pos = 8, zl = 0;
    while (pos + 12 <= n) {
        const uint32_t len = rd32be(b + pos);
        const uint8_t* type = b + pos + 4;
        if (pos + 12 + (size_t)len > n) return 0;
        if (!memcmp(type, "IDAT", 4)) {
            if (zl + len > cap || !png_chunk_crc_ok(b + pos, len)) return 0;
            memcpy(dst + zl, b + pos + 8, len);
            zl += len;
        } else if (!memcmp(type, "IEND", 4)) {
            break;
        }
        pos += 12 + (size_t)len;
    }
    return zl;
}

bool read_file(const std::string& path, std::vector<uint8_t>& out) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (sz < 0) { fclose(f); return false; }
    out.resize((size_t)sz);
    const size_t got = sz ? fread(out.data(), 1, (size_t)sz, f) : 0;
    fclose(f);
    return got == (size_t)sz;
}

bool write_bmp_gray(const std::string& path, const uint8_t* px, int rows, int cols) {
    const uint32_t stride = ((uint32_t)cols + 3) & ~3u, off = 14 + 40 + 1024, size = off + stride * rows;
    std::vector<uint8_t> b(size, 0);
    auto w32 = [&](size_t o, uint32_t v) { b[o] = v & 255; b[o + 1] = (v >> 8) & 255; b[o + 2] = (v >> 16) & 255; b[o + 3] = v >> 24; };
    b[0] = 'B'; b[1] = 'M'; w32(2, size); w32(10, off); w32(14, 40); w32(18, (uint32_t)cols); w32(22, (uint32_t)rows);
    b[26] = 1; b[28] = 8; w32(34, stride * rows); w32(46, 256);
    for (int i = 0; i < 256; i++) { b[54 + 4 * i] = b[55 + 4 * i] = b[56 + 4 * i] = (uint8_t)i; }
    for (int y = 0; y < rows; y++) memcpy(&b[off + stride * (size_t)(rows - 1 - y)], px + (size_t)y * cols, cols);
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) return false;
    const bool ok = fwrite(b.data(), 1, b.size(), f) == b.size();
    fclose(f);
    return ok;
}

}  // namespace abubcodec
