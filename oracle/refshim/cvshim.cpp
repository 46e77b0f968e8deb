/* oracle/refshim/cvshim.cpp -- TEST INFRASTRUCTURE: what stands behind the cv:: calls of the reference when it is
 * compiled into oracle/_ref/abub3hs_ref (see opencv2/opencv.hpp).  Pixel arithmetic = OpenCV 4.13.0's observable
 * behaviour as restated in oracle/abub_ref.c and checked against cv2 in tests/test_oracle_cpu.py. */
#include <float.h>
#include <zlib.h>

#include "opencv2/opencv.hpp"

extern "C" {
typedef struct {
    int32_t k, x, y, w, h, npts, first_px, reserved;
    int64_t sum_x, sum_y;
    double area, m00, m10, m01;
} ref_blob;
int ref_otsu(const uint32_t* h, size_t total);
int ref_find_contours(const uint8_t* bin, int rows, int cols, ref_blob* out, int cap, int32_t* pts_out, int pts_cap, int* npts_total);
void ref_measure_polygon(const int32_t* xy, int n, ref_blob* b);
}

namespace cv {

void shim_unsupported(const char* what) {
    fprintf(stderr, "oracle/refshim: cv::%s is not implemented (only dead or debug code of the reference should reach it)\n", what);
    abort();
}

static inline uchar sat_u8(int v) { return (uchar)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

static void same_shape(const Mat& a, const Mat& b, const char* what) {
    if (a.rows != b.rows || a.cols != b.cols || a.type() != b.type())
        throw Exception(std::string("(-209:Sizes of input arguments do not match) in function '") + what + "'");
}

/* contiguous copy of a (possibly ROI) CV_8U matrix */
static std::vector<uchar> packed(const Mat& m) {
    std::vector<uchar> v((size_t)m.rows * m.cols);
    for (int y = 0; y < m.rows; y++) memcpy(v.data() + (size_t)y * m.cols, m.ptr(y), (size_t)m.cols);
    return v;
}
static void unpack(const std::vector<uchar>& v, Mat& dst, int rows, int cols) {
    dst.create(rows, cols, CV_8U);
    for (int y = 0; y < rows; y++) memcpy(dst.ptr(y), v.data() + (size_t)y * cols, (size_t)cols);
}

Mat operator-(const Mat& a, const Mat& b) {
    same_shape(a, b, "arithm_op");
    Mat r(a.rows, a.cols, a.type());
    if (a.depth() == CV_8U) {
        for (int y = 0; y < a.rows; y++) {
            const uchar *pa = a.ptr(y), *pb = b.ptr(y);
            uchar* pr = r.ptr(y);
            for (int x = 0; x < a.cols; x++) pr[x] = sat_u8((int)pa[x] - (int)pb[x]);
        }
    } else {
        for (int y = 0; y < a.rows; y++)
            for (int x = 0; x < a.cols; x++) r.at<float>(y, x) = a.at<float>(y, x) - b.at<float>(y, x);
    }
    return r;
}
Mat operator+(const Mat& a, const Mat& b) {
    same_shape(a, b, "arithm_op");
    Mat r(a.rows, a.cols, a.type());
    if (a.depth() == CV_8U) {
        for (int y = 0; y < a.rows; y++)
            for (int x = 0; x < a.cols; x++) r.ptr(y)[x] = sat_u8((int)a.ptr(y)[x] + (int)b.ptr(y)[x]);
    } else {
        for (int y = 0; y < a.rows; y++)
            for (int x = 0; x < a.cols; x++) r.at<float>(y, x) = a.at<float>(y, x) + b.at<float>(y, x);
    }
    return r;
}
/* s * A: the reference only ever SUBTRACTS 6 * sigma from a CV_8U image.  OpenCV evaluates `X - 6*S` as one
 * addWeighted(X, 1, S, -6) = saturate(X - 6 S); evaluating 6*S first saturates it at 255, and saturate(X - min(6S, 255))
 * is the same value for every X in 0..255 (X - 255 <= 0 either way). */
Mat operator*(double s, const Mat& a) {
    Mat r(a.rows, a.cols, a.type());
    if (a.depth() == CV_8U) {
        for (int y = 0; y < a.rows; y++)
            for (int x = 0; x < a.cols; x++) r.ptr(y)[x] = sat_u8((int)lrint(a.ptr(y)[x] * s));
    } else {
        for (int y = 0; y < a.rows; y++)
            for (int x = 0; x < a.cols; x++) r.at<float>(y, x) = a.at<float>(y, x) * (float)s;
    }
    return r;
}
Mat operator*(const Mat& a, double s) { return s * a; }
/* A / s on a CV_32F histogram (Trainer.cpp:367, AnalyzerUnit.cpp:412): MatExpr scale alpha = 1/s applied by convertTo,
 * whose CV_32F -> CV_32F kernel multiplies by (float)alpha */
Mat operator/(const Mat& a, double s) {
    if (a.depth() != CV_32F) shim_unsupported("operator/ on a non-float Mat");
    const float alpha = (float)(1.0 / s);
    Mat r(a.rows, a.cols, a.type());
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) r.at<float>(y, x) = a.at<float>(y, x) * alpha;
    return r;
}
Mat& operator-=(Mat& a, const Mat& b) {
    Mat r = a - b;
    r.copyTo(a);                 /* in place: views of `a` see the result, as with cv::subtract(a, b, a) */
    return a;
}
Mat& operator+=(Mat& a, const Mat& b) {
    Mat r = a + b;
    r.copyTo(a);
    return a;
}
Mat& operator*=(Mat& a, double s) {
    Mat r = s * a;
    r.copyTo(a);
    return a;
}
void Mat::convertTo(Mat&, int, double, double) const { shim_unsupported("Mat::convertTo"); }

/* ---------------------------------------------------------------------------------------------------------------
 * imgcodecs: what cv::imread(path, IMREAD_GRAYSCALE) returns for the files of a run: 8-bit BMP (palette -> BGR ->
 * grey with OpenCV's fixed-point BT.601 weights, the identity for a grey palette) and 8-bit greyscale PNG.
 * Anything else -> empty Mat, as for an unreadable file.
 * --------------------------------------------------------------------------------------------------------------- */
static bool read_file(const std::string& path, std::vector<uchar>& out) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    const size_t got = n > 0 ? fread(out.data(), 1, (size_t)n, f) : 0;
    fclose(f);
    return n > 0 && got == (size_t)n;
}
static uint32_t le32(const uchar* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
static uint32_t be32(const uchar* p) { return ((uint32_t)p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3]; }

static Mat decode_bmp8(const std::vector<uchar>& b) {
    if (b.size() < 54 || b[0] != 'B' || b[1] != 'M') return Mat();
    const uint32_t off = le32(&b[10]), hdr = le32(&b[14]);
    const int w = (int)le32(&b[18]), hs = (int)le32(&b[22]);
    const int bpp = b[28] | (b[29] << 8);
    const uint32_t comp = le32(&b[30]);
    uint32_t ncol = le32(&b[46]);
    if (hdr < 40 || bpp != 8 || comp != 0 || w <= 0 || hs == 0) return Mat();
    if (ncol == 0) ncol = 256;
    const int h = hs < 0 ? -hs : hs;
    const size_t stride = ((size_t)w + 3) & ~(size_t)3;
    if (14 + hdr + 4 * (size_t)ncol > b.size() || off + stride * (size_t)h > b.size()) return Mat();
    uchar grey[256] = {0};
    for (uint32_t i = 0; i < ncol && i < 256; i++) {
        const uchar* q = &b[14 + hdr + 4 * i];                   /* B, G, R, 0 */
        grey[i] = (uchar)((q[0] * 1868 + q[1] * 9617 + q[2] * 4899 + 8192) >> 14);
    }
    Mat m(h, w, CV_8U);
    for (int y = 0; y < h; y++) {
        const uchar* src = &b[off + stride * (size_t)(hs < 0 ? y : h - 1 - y)];
        uchar* dst = m.ptr(y);
        for (int x = 0; x < w; x++) dst[x] = grey[src[x]];
    }
    return m;
}

static Mat decode_png_grey8(const std::vector<uchar>& b) {
    static const uchar sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n'};
    if (b.size() < 33 || memcmp(b.data(), sig, 8) != 0) return Mat();
    size_t pos = 8;
    int w = 0, h = 0;
    std::vector<uchar> z;
    while (pos + 12 <= b.size()) {
        const uint32_t n = be32(&b[pos]);
        if (pos + 12 + (size_t)n > b.size()) return Mat();
        const uchar* tag = &b[pos + 4];
        if (crc32(crc32(0, tag, 4), tag + 4, n) != be32(&b[pos + 8 + n])) return Mat();
        if (!memcmp(tag, "IHDR", 4)) {
            w = (int)be32(&b[pos + 8]);
            h = (int)be32(&b[pos + 12]);
            if (b[pos + 16] != 8 || b[pos + 17] != 0 || b[pos + 20] != 0) return Mat();    /* depth 8, grey, not interlaced */
        } else if (!memcmp(tag, "IDAT", 4)) {
            z.insert(z.end(), tag + 4, tag + 4 + n);
        } else if (!memcmp(tag, "IEND", 4)) {
            break;
        }
        pos += 12 + (size_t)n;
    }
    if (w <= 0 || h <= 0 || z.empty()) return Mat();
    std::vector<uchar> raw((size_t)h * (w + 1));
    uLongf len = (uLongf)raw.size();
    if (uncompress(raw.data(), &len, z.data(), (uLong)z.size()) != Z_OK || len != raw.size()) return Mat();
    Mat m(h, w, CV_8U);
    std::vector<uchar> zero((size_t)w, 0);
    for (int y = 0; y < h; y++) {
        const uchar* in = &raw[(size_t)y * (w + 1)];
        const uchar* up = y ? m.ptr(y - 1) : zero.data();
        uchar* out = m.ptr(y);
        const int ft = in[0];
        for (int x = 0; x < w; x++) {
            const int a = x ? out[x - 1] : 0, bb = up[x], c = x ? up[x - 1] : 0;
            int pred = 0;
            switch (ft) {
                case 0: pred = 0; break;
                case 1: pred = a; break;
                case 2: pred = bb; break;
                case 3: pred = (a + bb) >> 1; break;
                case 4: {
                    const int p = a + bb - c, pa = abs(p - a), pb = abs(p - bb), pc = abs(p - c);
                    pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? bb : c);
                    break;
                }
                default: return Mat();
            }
            out[x] = (uchar)(in[1 + x] + pred);
        }
    }
    return m;
}

Mat imread(const std::string& path, int flags) {
    if (flags != IMREAD_GRAYSCALE) shim_unsupported("imread with flags != IMREAD_GRAYSCALE");
    std::vector<uchar> b;
    if (!read_file(path, b)) return Mat();
    Mat m = decode_bmp8(b);
    if (m.empty()) m = decode_png_grey8(b);
    return m;
}
bool imwrite(const std::string&, const Mat&) { return true; }

}  // namespace cv

#include "cvshim_imgproc.inc"
