/* abub_ref.c -- TEST INFRASTRUCTURE: cv2-free plain-C restatement of the integer/float
 * arithmetic of the AutoBub3hs hot path.  Never linked into the product; used by tests/
 * (through oracle/cref.py) as a second oracle and by bench.py as a timed CPU port.
 *
 * The reference delegates this arithmetic to OpenCV (unpinned, not under /root/reference):
 * call sites AnalyzerUnit.cpp:351-370 (subtract/scale, GaussianBlur, absdiff), :456
 * (calcHist), L3Localizer.cpp:252-264, 779-793 (threshold TOZERO / BINARY|OTSU, blur,
 * findContours RETR_EXTERNAL + CHAIN_APPROX_TC89_L1, boundingRect, contourArea, moments).
 * What is restated here is OpenCV 4.13.0's observable behaviour for those calls
 * (SURVEY.md Appendix A), and every function is checked against cv2 4.13.0 in
 * tests/test_oracle_cpu.py.  First-party arithmetic follows Trainer.cpp:144-216 and
 * AnalyzerUnit.cpp:435-532.
 *
 * PARITY PIN: "parity unpinned" w.r.t. reference-owned vectors (the reference has none);
 * pinned to cv2 4.13.0 outputs instead.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline int reflect101(int p, int n) {
    if (p < 0) p = -p;
    if (p >= n) p = 2 * (n - 1) - p;
    if (p < 0) p = 0;
    if (p >= n) p = n - 1;
    return p;
}

/* ---- a1: AnalyzerUnit::ProcessFrame (AnalyzerUnit.cpp:346-377) ---------------------- */
/* pos = sat(a-b-6s), neg = sat(b-a-6s); 5x5 Gaussian (sigma 0 -> [1 4 6 4 1]/16 fixed
 * point, (sum+128)>>8, BORDER_REFLECT_101); out = |P-N|. */
void ref_process_frame(const uint8_t* a, const uint8_t* b, const uint8_t* sigma, int rows, int cols,
                       uint8_t* out) {
    static const int k[5] = {1, 4, 6, 4, 1};
    size_t n = (size_t)rows * cols;
    uint8_t* pos = (uint8_t*)malloc(n);
    uint8_t* neg = (uint8_t*)malloc(n);
    uint16_t* hp = (uint16_t*)malloc(n * sizeof(uint16_t));
    uint16_t* hn = (uint16_t*)malloc(n * sizeof(uint16_t));
    for (size_t i = 0; i < n; i++) {
        int s = (int)a[i] - (int)b[i], t = 6 * (int)sigma[i];
        int p = s - t, q = -s - t;
        pos[i] = (uint8_t)(p > 0 ? p : 0);
        neg[i] = (uint8_t)(q > 0 ? q : 0);
    }
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++) {
            int sp = 0, sn = 0;
            for (int j = -2; j <= 2; j++) {
                int xx = reflect101(x + j, cols);
                sp += k[j + 2] * pos[(size_t)y * cols + xx];
                sn += k[j + 2] * neg[(size_t)y * cols + xx];
            }
            hp[(size_t)y * cols + x] = (uint16_t)sp;
            hn[(size_t)y * cols + x] = (uint16_t)sn;
        }
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++) {
            int sp = 0, sn = 0;
            for (int j = -2; j <= 2; j++) {
                int yy = reflect101(y + j, rows);
                sp += k[j + 2] * hp[(size_t)yy * cols + x];
                sn += k[j + 2] * hn[(size_t)yy * cols + x];
            }
            int P = (sp + 128) >> 8, N = (sn + 128) >> 8;
            out[(size_t)y * cols + x] = (uint8_t)(P > N ? P - N : N - P);
        }
    free(pos); free(neg); free(hp); free(hn);
}

/* ---- a2: calcHist 256 bins (AnalyzerUnit.cpp:456) ---------------------------------- */
void ref_hist256(const uint8_t* img, size_t n, uint32_t* hist) {
    memset(hist, 0, 256 * sizeof(uint32_t));
    for (size_t i = 0; i < n; i++) hist[img[i]]++;
}

/* ---- a9 front end: absdiff(frame,avg) - 6 sigma, blur 3x3 (L3Localizer.cpp:779-785) -- */
/* cv::blur 3x3 on CV_8U: integer box sum scaled by 1/9 and rounded to nearest
 * (no ties possible) = (sum + 4) / 9, BORDER_REFLECT_101. */
void ref_post_over(const uint8_t* frame, const uint8_t* avg, const uint8_t* sigma, int rows, int cols,
                   uint8_t* out) {
    size_t n = (size_t)rows * cols;
    uint8_t* d = (uint8_t*)malloc(n);
    for (size_t i = 0; i < n; i++) {
        int v = abs((int)frame[i] - (int)avg[i]) - 6 * (int)sigma[i];
        d[i] = (uint8_t)(v > 0 ? v : 0);
    }
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++) {
            int s = 0;
            for (int j = -1; j <= 1; j++)
                for (int i = -1; i <= 1; i++)
                    s += d[(size_t)reflect101(y + j, rows) * cols + reflect101(x + i, cols)];
            out[(size_t)y * cols + x] = (uint8_t)((s + 4) / 9);
        }
    free(d);
}

/* ---- THRESH_TOZERO + THRESH_BINARY|THRESH_OTSU (L3Localizer.cpp:252-254, 786-787) ---- */
/* Otsu as OpenCV's getThreshVal_Otsu_8u. */
int ref_otsu(const uint32_t* h, size_t total) {
    double scale = 1.0 / (double)total, mu = 0;
    for (int i = 0; i < 256; i++) mu += i * (double)h[i];
    mu *= scale;
    double mu1 = 0, q1 = 0, max_sigma = 0;
    int t = 0;
    for (int i = 0; i < 256; i++) {
        double p_i = h[i] * scale, q2, mu2, sg;
        mu1 *= q1;
        q1 += p_i;
        q2 = 1. - q1;
        if (fmin(q1, q2) < 1.1920928955078125e-07 || fmax(q1, q2) > 1. - 1.1920928955078125e-07) continue;
        mu1 = (mu1 + i * p_i) / q1;
        mu2 = (mu - q1 * mu1) / q2;
        sg = q1 * q2 * (mu1 - mu2) * (mu1 - mu2);
        if (sg > max_sigma) { max_sigma = sg; t = i; }
    }
    return t;
}

/* src > thres ? src : 0, then Otsu over the result, then src > t ? 255 : 0.  Returns t. */
int ref_tozero_otsu_binary(const uint8_t* src, size_t n, int tozero, uint8_t* out) {
    uint32_t h[256];
    memset(h, 0, sizeof(h));
    for (size_t i = 0; i < n; i++) {
        uint8_t v = src[i] > tozero ? src[i] : 0;
        out[i] = v;
        h[v]++;
    }
    int t = ref_otsu(h, n);
    for (size_t i = 0; i < n; i++) out[i] = out[i] > t ? 255 : 0;
    return t;
}

/* ---- findContours(RETR_EXTERNAL, CHAIN_APPROX_TC89_L1) + boundingRect/contourArea/moments
 * (L3Localizer.cpp:264,278,401-403 / :793,808,818-821) -- SURVEY.md Appendix A ------------ */
typedef struct {
    int32_t k, x, y, w, h, npts, first_px, reserved;
    int64_t sum_x, sum_y;
    double area, m00, m10, m01;
} ref_blob;                          /* same layout as abub_blob (include/abub_cuda.h) */

static const int DX[8] = {1, 1, 0, -1, -1, -1, 0, 1};
static const int DY[8] = {0, -1, -1, -1, 0, 1, 1, 1};

typedef struct { int x, y; } ipt;

/* Suzuki outer-border following from the component's first pixel p0 (Appendix A.3).
 * img is the padded 0/1 image with stride W.  Returns chain length; pts = positions before
 * each code.  Isolated pixel -> len 0 (single point). */
static int trace_border(const uint8_t* img, int W, int px, int py, uint8_t** codes_io, ipt** pts_io, int* cap_io) {
    int s = 4, s_end = 4;
    int cap = *cap_io, len = 0;
    uint8_t* codes = *codes_io;
    ipt* pts = *pts_io;
    do {
        s = (s - 1) & 7;
        if (img[(py + DY[s]) * W + (px + DX[s])]) break;
    } while (s != s_end);
    if (s == s_end) return 0;
    int i1x = px + DX[s], i1y = py + DY[s];
    int i3x = px, i3y = py;
    for (;;) {
        int i4x, i4y;
        for (;;) {
            s = (s + 1) & 7;
            i4x = i3x + DX[s]; i4y = i3y + DY[s];
            if (img[i4y * W + i4x]) break;
        }
        if (len == cap) {
            cap *= 2;
            codes = (uint8_t*)realloc(codes, cap);
            pts = (ipt*)realloc(pts, cap * sizeof(ipt));
        }
        codes[len] = (uint8_t)s;
        pts[len].x = i3x; pts[len].y = i3y;
        len++;
        if (i4x == px && i4y == py && i3x == i1x && i3y == i1y) break;
        i3x = i4x; i3y = i4y;
        s = (s + 4) & 7;
    }
    *codes_io = codes; *pts_io = pts; *cap_io = cap;
    return len;
}

static int next_alive(const uint8_t* removed, int i, int L) {
    i++;
    while (i < L && removed[i]) i++;
    return i;
}

/* TC89_L1 approximation (Appendix A.4-A.10).  In: chain of `len` points/codes.  Out: alive
 * points in index order into out_pts (capacity >= len + 1).  Returns the vertex count. */
static int approx_tc89_l1(const uint8_t* codes, const ipt* P0, int len, ipt* out_pts) {
    static const int T[15] = {1, 2, 3, 4, 3, 2, 1, 0, 1, 2, 3, 4, 3, 2, 1};
    int L = len;
    ipt* P = (ipt*)malloc((len + 1) * sizeof(ipt));
    int* S = (int*)malloc((len + 1) * sizeof(int));
    int* K = (int*)calloc(len + 1, sizeof(int));
    uint8_t* removed = (uint8_t*)calloc(len + 1, 1);
    memcpy(P, P0, len * sizeof(ipt));
    /* pass 0 */
    for (int i = 0; i < len; i++) {
        int prev = codes[(i + len - 1) % len];
        S[i] = T[(int)codes[i] - prev + 7];
        removed[i] = (S[i] == 0);
    }
    /* pass 1: support region */
    for (int i = 0; i < len; i++) {
        if (removed[i]) continue;
        int l = 0, d_num = 0, k;
        for (k = 1;; k++) {
            int i1 = ((i - k) % len + len) % len, i2 = (i + k) % len;
            int dx = P[i2].x - P[i1].x, dy = P[i2].y - P[i1].y;
            int lk = dx * dx + dy * dy;
            int dk = (P[i].x - P[i1].x) * dy - (P[i].y - P[i1].y) * dx;
            float d = (float)((double)d_num * lk - (double)dk * l);
            if (k > 1 && (l >= lk || (d_num > 0 && d <= 0) || (d_num < 0 && d >= 0))) break;
            d_num = dk;
            l = lk;
        }
        K[i] = k - 1;
    }
    /* pass 2: non-maximum suppression using live S */
    for (int i = 0; i < len; i++) {
        if (removed[i]) continue;
        int k2 = K[i] >> 1, s = S[i], kill = 0;
        for (int j = 1; j <= k2 && !kill; j++) {
            int a = ((i - j) % len + len) % len, b = (i + j) % len;
            if (S[a] > s || S[b] > s) kill = 1;
        }
        if (kill) { removed[i] = 1; S[i] = 0; }
    }
    /* pass 3 */
    for (int i = 0; i < len; i++) {
        if (removed[i] || K[i] != 1) continue;
        int a = (i + len - 1) % len, b = (i + 1) % len;
        if (S[i] <= S[a] || S[i] <= S[b]) { removed[i] = 1; S[i] = 0; }
    }
    /* pass 4a: wrap-around run */
    int start = 0, all = 0;
    if (S[0] != 0 && S[len - 1] != 0) {
        int i1 = 1, i2;
        while (i1 < len && S[i1] != 0) { S[i1 - 1] = 0; removed[i1 - 1] = 1; i1++; }
        if (i1 == len) {
            all = 1;
        } else {
            i1--;
            i2 = len - 2;
            while (i2 > 0 && S[i2] != 0) { S[i2 + 1] = 0; removed[i2 + 1] = 1; i2--; }
            i2++;
            if (i1 == 0 && i2 == len - 1) {
                i1 = next_alive(removed, 0, len);
                P[len] = P[0]; S[len] = S[0]; K[len] = K[0]; removed[len] = removed[0];
                L = len + 1;
            }
            start = i1;
        }
    }
    int n = 0;
    if (all) {
        for (int i = 0; i < len; i++) out_pts[n++] = P0[i];
    } else {
        /* pass 4b: runs of consecutive alive indices */
        int cur = start;
        if (cur < L && removed[cur]) cur = next_alive(removed, cur, L);
        int first = start, count = 1, prev = cur;
        while (cur < L) {
            int nx = next_alive(removed, cur, L);
            if (nx >= L || nx - cur != 1) {
                if (count == 2) {
                    int a = prev, b = cur;
                    if (S[a] > S[b] || (S[a] == S[b] && K[a] <= K[b])) removed[b] = 1; else removed[a] = 1;
                } else if (count > 2) {
                    int f = removed[first] ? next_alive(removed, first, L) : first;
                    /* remove every index strictly between next_alive(first) and cur */
                    int lo = next_alive(removed, first, L);
                    (void)f;
                    for (int q = lo + 1; q < cur; q++) removed[q] = 1;
                }
                first = cur;
                count = 1;
            } else {
                count++;
            }
            prev = cur;
            cur = nx;
        }
        for (int i = 0; i < L; i++)
            if (!removed[i]) out_pts[n++] = P[i];
    }
    free(P); free(S); free(K); free(removed);
    return n;
}

/* boundingRect, contourArea, moments of an integer polygon (Appendix A.10). */
static void measure_polygon(const ipt* v, int n, ref_blob* b) {
    int minx = v[0].x, maxx = v[0].x, miny = v[0].y, maxy = v[0].y;
    int64_t sx = 0, sy = 0;
    for (int i = 0; i < n; i++) {
        if (v[i].x < minx) minx = v[i].x;
        if (v[i].x > maxx) maxx = v[i].x;
        if (v[i].y < miny) miny = v[i].y;
        if (v[i].y > maxy) maxy = v[i].y;
        sx += v[i].x; sy += v[i].y;
    }
    b->x = minx; b->y = miny; b->w = maxx - minx + 1; b->h = maxy - miny + 1;
    b->npts = n; b->sum_x = sx; b->sum_y = sy;
    /* contourArea: a00 accumulated as xi_1*yi - xi*yi_1 in double, *0.5, fabs */
    double a00 = 0, a10 = 0, a01 = 0, area = 0;
    double xp = v[n - 1].x, yp = v[n - 1].y;
    for (int i = 0; i < n; i++) {
        double xi = v[i].x, yi = v[i].y;
        area += xp * yi - xi * yp;
        double dxy = xp * yi - xi * yp;
        a00 += dxy;
        a10 += dxy * (xp + xi);
        a01 += dxy * (yp + yi);
        xp = xi; yp = yi;
    }
    b->area = fabs(area * 0.5);
    if (fabs(a00) > 1.1920928955078125e-07) {
        double db1_2 = a00 > 0 ? 0.5 : -0.5, db1_6 = a00 > 0 ? 0.16666666666666666667 : -0.16666666666666666667;
        b->m00 = a00 * db1_2; b->m10 = a10 * db1_6; b->m01 = a01 * db1_6;
    } else {
        b->m00 = b->m10 = b->m01 = 0;
    }
}

/* the same measures for a polygon given as (x, y) int32 pairs: cv::boundingRect / contourArea / moments of a contour
 * (used by oracle/refshim, where the reference's own code calls them on the contours it holds) */
void ref_measure_polygon(const int32_t* xy, int n, ref_blob* b) {
    ipt* v = (ipt*)malloc((size_t)(n > 0 ? n : 1) * sizeof(ipt));
    for (int i = 0; i < n; i++) { v[i].x = xy[2 * i]; v[i].y = xy[2 * i + 1]; }
    if (n > 0) measure_polygon(v, n, b);
    free(v);
}

/* External contours of a binary image (non-zero = foreground) in findContours order.
 * pts_out (may be NULL): polygon vertices (x,y int32 pairs) of all blobs back to back,
 * capacity pts_cap pairs.  Returns the number of contours (<= cap written), or -1. */
int ref_find_contours(const uint8_t* bin, int rows, int cols, ref_blob* out, int cap, int32_t* pts_out,
                      int pts_cap, int* npts_total) {
    const int W = cols + 2, Hh = rows + 2;
    uint8_t* img = (uint8_t*)calloc((size_t)W * Hh, 1);
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++) img[(y + 1) * W + x + 1] = bin[(size_t)y * cols + x] ? 1 : 0;
    /* outer background: 4-connected flood from the padding ring -> 2 */
    int* stack = (int*)malloc((size_t)W * Hh * sizeof(int));
    int sp = 0;
    stack[sp++] = 0; img[0] = 2;
    while (sp) {
        int p = stack[--sp], x = p % W, y = p / W;
        if (x > 0 && img[p - 1] == 0) { img[p - 1] = 2; stack[sp++] = p - 1; }
        if (x < W - 1 && img[p + 1] == 0) { img[p + 1] = 2; stack[sp++] = p + 1; }
        if (y > 0 && img[p - W] == 0) { img[p - W] = 2; stack[sp++] = p - W; }
        if (y < Hh - 1 && img[p + W] == 0) { img[p + W] = 2; stack[sp++] = p + W; }
    }
    /* foreground components (8-connected), raster order of first pixel; externality */
    int ncomp = 0, ccap = 1024;
    int* firsts = (int*)malloc(ccap * sizeof(int));
    for (int p = 0; p < W * Hh; p++) {
        if (img[p] != 1) continue;
        int external = 0;
        sp = 0; stack[sp++] = p; img[p] = 3;
        while (sp) {
            int q = stack[--sp];
            if (img[q - 1] == 2 || img[q + 1] == 2 || img[q - W] == 2 || img[q + W] == 2) external = 1;
            for (int s = 0; s < 8; s++) {
                int r = q + DY[s] * W + DX[s];
                if (img[r] == 1) { img[r] = 3; stack[sp++] = r; }
            }
        }
        if (external) {
            if (ncomp == ccap) { ccap *= 2; firsts = (int*)realloc(firsts, ccap * sizeof(int)); }
            firsts[ncomp++] = p;
        }
    }
    /* binary view for tracing: foreground = 3 -> 1, everything else 0 */
    for (int p = 0; p < W * Hh; p++) img[p] = (img[p] == 3);
    int tcap = 256, total_pts = 0, nout = 0;
    uint8_t* codes = (uint8_t*)malloc(tcap);
    ipt* pts = (ipt*)malloc(tcap * sizeof(ipt));
    for (int c = ncomp - 1; c >= 0; c--) {           /* descending raster order */
        int px = firsts[c] % W, py = firsts[c] / W;
        int len = trace_border(img, W, px, py, &codes, &pts, &tcap);
        ipt* poly = (ipt*)malloc((len + 2) * sizeof(ipt));
        int n;
        if (len == 0) { poly[0].x = px; poly[0].y = py; n = 1; }
        else n = approx_tc89_l1(codes, pts, len, poly);
        for (int i = 0; i < n; i++) { poly[i].x -= 1; poly[i].y -= 1; }   /* remove padding offset */
        if (nout < cap) {
            ref_blob* b = &out[nout];
            memset(b, 0, sizeof(*b));
            measure_polygon(poly, n, b);
            b->first_px = (py - 1) * cols + (px - 1);
            if (pts_out) {
                for (int i = 0; i < n && total_pts + i < pts_cap; i++) {
                    pts_out[2 * (total_pts + i)] = poly[i].x;
                    pts_out[2 * (total_pts + i) + 1] = poly[i].y;
                }
            }
        }
        total_pts += n;
        nout++;
        free(poly);
    }
    if (npts_total) *npts_total = total_pts;
    free(codes); free(pts); free(firsts); free(stack); free(img);
    return nout;
}

/* ---- a6: Welford mean/sigma (Trainer.cpp:144-216), float32, frame index innermost ---- */
void ref_welford(const uint8_t* const* frames, int n, size_t npix, uint8_t* mean_out, uint8_t* sigma_out) {
    for (size_t p = 0; p < npix; p++) {
        volatile float mean = 0, m2 = 0;        /* volatile: no contraction / reassociation */
        for (int f = 0; f < n; f++) {
            float x = (float)frames[f][p];
            float delta = x - mean;
            mean = mean + delta / (float)(f + 1);
            float t = x - mean;
            float prod = delta * t;
            m2 = m2 + prod;
        }
        float var = m2 / (float)(size_t)(n - 1);
        float sg = sqrtf(var);
        int si = (sg != sg) ? 0 : (int)sg;      /* (int)NaN -> INT_MIN on x86 -> low byte 0 */
        sigma_out[p] = (uint8_t)si;
        mean_out[p] = (uint8_t)(float)mean;
    }
}
