// Border following, TC89_L1 approximation and polygon measurement shared by the K7 kernels.
// Restates what cv::findContours(RETR_EXTERNAL, CHAIN_APPROX_TC89_L1) + boundingRect /
// contourArea / moments produce (OpenCV 4.13.0 behaviour) for the
// calls at L3Localizer.cpp:264,278,401-403,793,808,818-821.
//
// Everything is __host__ __device__ so that the exact device code path can also be fuzzed
// on the CPU box by tests/host_contour_driver.cu (test-only; the product calls these from
// kernels only).
#pragma once
#include <stdint.h>

#ifndef __CUDACC__
#define __host__
#define __device__
#endif

namespace abub_contour {

struct Pix {            // foreground predicate over a TOZERO'd image: value > thr
    const uint8_t* tz;
    int thr, rows, cols;
    __host__ __device__ bool fg(int x, int y) const {
        return (unsigned)x < (unsigned)cols && (unsigned)y < (unsigned)rows && tz[(size_t)y * cols + x] > thr;
    }
};

__host__ __device__ inline int cdx(int s) { return (s == 0 || s == 1 || s == 7) ? 1 : ((s >= 3 && s <= 5) ? -1 : 0); }
__host__ __device__ inline int cdy(int s) { return (s >= 1 && s <= 3) ? -1 : ((s >= 5 && s <= 7) ? 1 : 0); }

struct ChainView {      // one contour's region of the chain arena, L1 = len + 1 elements
    short2* P;
    int32_t* K;
    uint8_t* code;      // bits 0-2 chain code, bit 3: left crack examined, bit 4: right crack examined
    uint8_t* S;
    uint8_t* removed;
    __host__ __device__ static size_t bytes(int len) { return (size_t)12 * (len + 1); }
    __host__ __device__ ChainView(uint8_t* base, int len) {
        const size_t L1 = (size_t)len + 1;
        P = reinterpret_cast<short2*>(base);
        K = reinterpret_cast<int32_t*>(base + 4 * L1);
        code = base + 8 * L1;
        S = code + L1;
        removed = S + L1;
    }
};

struct TraceResult {
    int len;            // number of chain codes (0 = isolated pixel)
    int accepted;       // 0: met a pixel earlier than the start in raster order
    int minx, maxx, miny, maxy;
};

// The same predicate with a window of the binarised image held in fast memory (shared memory on the device): pixels inside
// the window are answered from it, the rest from the image itself.  win[(y - wy0) * ww + (x - wx0)] != 0 <=> foreground.
struct WinPix {
    Pix g;
    const uint8_t* win;
    int wx0, wy0, ww, wh;
    __host__ __device__ bool fg(int x, int y) const {
        const unsigned dx = (unsigned)(x - wx0), dy = (unsigned)(y - wy0);
        if (dx < (unsigned)ww && dy < (unsigned)wh) return win[dy * (unsigned)ww + dx] != 0;
        return g.fg(x, y);
    }
};

// Suzuki outer-border following from (px,py) whose W, NW, N, NE neighbours are background.
// store == nullptr: measure only.  max_len guards against runaway loops.
template <class PIX>
__host__ __device__ inline TraceResult trace_border(const PIX& im, int px, int py, ChainView* store, int max_len) {
    TraceResult r;
    r.len = 0; r.accepted = 1; r.minx = r.maxx = px; r.miny = r.maxy = py;
    int s = 4;
    bool found = false;
    for (int it = 0; it < 7; it++) {          // s = 3,2,1,0,7,6,5 (clockwise from NW)
        s = (s - 1) & 7;
        if (im.fg(px + cdx(s), py + cdy(s))) { found = true; break; }
    }
    if (!found) return r;                      // isolated pixel: single-point contour
    const int i1x = px + cdx(s), i1y = py + cdy(s);
    int i3x = px, i3y = py;
    for (;;) {
        int i4x, i4y;
        unsigned flags = 0;
        for (;;) {
            s = (s + 1) & 7;
            i4x = i3x + cdx(s); i4y = i3y + cdy(s);
            if (im.fg(i4x, i4y)) break;
            if (s == 4) flags |= 8;            // left neighbour examined and background
            if (s == 0) flags |= 16;           // right neighbour examined and background
        }
        if (store) {
            store->P[r.len] = make_short2((short)i3x, (short)i3y);
            store->code[r.len] = (uint8_t)(s | flags);
        }
        r.len++;
        if (i4y < py || (i4y == py && i4x < px)) { r.accepted = 0; return r; }
        if (i4x < r.minx) r.minx = i4x;
        if (i4x > r.maxx) r.maxx = i4x;
        if (i4y < r.miny) r.miny = i4y;
        if (i4y > r.maxy) r.maxy = i4y;
        if (i4x == px && i4y == py && i3x == i1x && i3y == i1y) break;
        if (r.len >= max_len) { r.accepted = 0; return r; }
        i3x = i4x; i3y = i4y;
        s = (s + 4) & 7;
    }
    return r;
}

__host__ __device__ inline int next_alive(const uint8_t* removed, int i, int L) {
    i++;
    while (i < L && removed[i]) i++;
    return i;
}

// TC89_L1 (Appendix A.4-A.9) in place on a stored chain of `len` >= 1 elements.
// Returns L (len, or len+1 when element 0 was appended as a copy); alive = !removed[i], i < L.
__host__ __device__ inline int approx_tc89_l1(ChainView& c, int len) {
    const int T[15] = {1, 2, 3, 4, 3, 2, 1, 0, 1, 2, 3, 4, 3, 2, 1};
    int L = len;
    short2* P = c.P;
    for (int i = 0; i < len; i++) {                                   // pass 0
        const int prev = c.code[(i + len - 1) % len] & 7;
        const int sv = T[(int)(c.code[i] & 7) - prev + 7];
        c.S[i] = (uint8_t)sv;
        c.removed[i] = (sv == 0);
        c.K[i] = 0;
    }
    for (int i = 0; i < len; i++) {                                   // pass 1: support region
        if (c.removed[i]) continue;
        int l = 0, d_num = 0, k;
        for (k = 1;; k++) {
            int i1 = i - k; i1 %= len; if (i1 < 0) i1 += len;
            const int i2 = (i + k) % len;
            const int dx = P[i2].x - P[i1].x, dy = P[i2].y - P[i1].y;
            const int lk = dx * dx + dy * dy;
            const int dk = (P[i].x - P[i1].x) * dy - (P[i].y - P[i1].y) * dx;
            const float d = (float)((double)d_num * (double)lk - (double)dk * (double)l);
            if (k > 1 && (l >= lk || (d_num > 0 && d <= 0) || (d_num < 0 && d >= 0))) break;
            d_num = dk;
            l = lk;
        }
        c.K[i] = k - 1;
    }
    for (int i = 0; i < len; i++) {                                   // pass 2: non-maximum suppression
        if (c.removed[i]) continue;
        const int k2 = c.K[i] >> 1, sv = c.S[i];
        bool kill = false;
        for (int j = 1; j <= k2 && !kill; j++) {
            int a = i - j; a %= len; if (a < 0) a += len;
            const int b = (i + j) % len;
            if (c.S[a] > sv || c.S[b] > sv) kill = true;
        }
        if (kill) { c.removed[i] = 1; c.S[i] = 0; }
    }
    for (int i = 0; i < len; i++) {                                   // pass 3
        if (c.removed[i] || c.K[i] != 1) continue;
        const int a = (i + len - 1) % len, b = (i + 1) % len;
        if (c.S[i] <= c.S[a] || c.S[i] <= c.S[b]) { c.removed[i] = 1; c.S[i] = 0; }
    }
    int start = 0;
    if (c.S[0] != 0 && c.S[len - 1] != 0) {                           // pass 4a: wrap-around run
        int i1 = 1, i2;
        while (i1 < len && c.S[i1] != 0) { c.S[i1 - 1] = 0; c.removed[i1 - 1] = 1; i1++; }
        if (i1 == len) {                                              // every point kept
            for (int i = 0; i < len; i++) c.removed[i] = 0;
            return len;
        }
        i1--;
        i2 = len - 2;
        while (i2 > 0 && c.S[i2] != 0) { c.S[i2 + 1] = 0; c.removed[i2 + 1] = 1; i2--; }
        i2++;
        if (i1 == 0 && i2 == len - 1) {
            i1 = next_alive(c.removed, 0, len);
            P[len] = P[0]; c.S[len] = c.S[0]; c.K[len] = c.K[0]; c.removed[len] = c.removed[0];
            L = len + 1;
        }
        start = i1;
    }
    int cur = start;                                                  // pass 4b: runs of consecutive indices
    if (cur < L && c.removed[cur]) cur = next_alive(c.removed, cur, L);
    int first = start, count = 1, prev = cur;
    while (cur < L) {
        const int nx = next_alive(c.removed, cur, L);
        if (nx >= L || nx - cur != 1) {
            if (count == 2) {
                const int a = prev, b = cur;
                if (c.S[a] > c.S[b] || (c.S[a] == c.S[b] && c.K[a] <= c.K[b])) c.removed[b] = 1; else c.removed[a] = 1;
            } else if (count > 2) {
                const int lo = next_alive(c.removed, first, L);
                for (int q = lo + 1; q < cur; q++) c.removed[q] = 1;
            }
            first = cur;
            count = 1;
        } else {
            count++;
        }
        prev = cur;
        cur = nx;
    }
    return L;
}

struct PolyMeasure {
    int x, y, w, h, npts;
    long long sum_x, sum_y;
    double area, m00, m10, m01;
};

// boundingRect / contourArea / moments over the alive vertices (Appendix A.10).  All
// accumulations are exact integers in double (|coords| < 2^15), so FMA contraction cannot
// change them; the two scalings are single IEEE multiplications.
__host__ __device__ inline PolyMeasure measure_alive(const ChainView& c, int L) {
    PolyMeasure m;
    int last = -1, n = 0;
    for (int i = 0; i < L; i++) if (!c.removed[i]) { last = i; n++; }
    int minx = 1 << 30, maxx = -(1 << 30), miny = 1 << 30, maxy = -(1 << 30);
    long long sx = 0, sy = 0;
    double a00 = 0, a10 = 0, a01 = 0;
    double xp = c.P[last].x, yp = c.P[last].y;
    for (int i = 0; i < L; i++) {
        if (c.removed[i]) continue;
        const int xi_ = c.P[i].x, yi_ = c.P[i].y;
        if (xi_ < minx) minx = xi_;
        if (xi_ > maxx) maxx = xi_;
        if (yi_ < miny) miny = yi_;
        if (yi_ > maxy) maxy = yi_;
        sx += xi_; sy += yi_;
        const double xi = xi_, yi = yi_;
        const double dxy = xp * yi - xi * yp;
        a00 += dxy;
        a10 += dxy * (xp + xi);
        a01 += dxy * (yp + yi);
        xp = xi; yp = yi;
    }
    m.x = minx; m.y = miny; m.w = maxx - minx + 1; m.h = maxy - miny + 1; m.npts = n;
    m.sum_x = sx; m.sum_y = sy;
    const double half = a00 * 0.5;
    m.area = half < 0 ? -half : half;
    const double aa = a00 < 0 ? -a00 : a00;
    if (aa > 1.1920928955078125e-07) {
        const double db1_2 = a00 > 0 ? 0.5 : -0.5, db1_6 = a00 > 0 ? 0.16666666666666666667 : -0.16666666666666666667;
        m.m00 = a00 * db1_2; m.m10 = a10 * db1_6; m.m01 = a01 * db1_6;
    } else {
        m.m00 = m.m10 = m.m01 = 0;
    }
    return m;
}

// Is pixel (qx,qy), which does not belong to the contour's component, inside its outer
// border (i.e. in one of its holes)?  Parity of the vertical boundary cracks of the filled
// outline that lie to the left of the pixel on its row; each crack was recorded exactly once
// while following the border (bit 3: crack on the left side of P[i], bit 4: right side).
__host__ __device__ inline bool inside_outer_border(const ChainView& c, int len, int qx, int qy) {
    int n = 0;
    for (int i = 0; i < len; i++) {
        if (c.P[i].y != qy) continue;
        const unsigned f = c.code[i];
        if ((f & 8) && c.P[i].x <= qx) n++;       // left crack at x - 1/2
        if ((f & 16) && c.P[i].x < qx) n++;       // right crack at x + 1/2
    }
    return (n & 1) != 0;
}

}  // namespace abub_contour
