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
