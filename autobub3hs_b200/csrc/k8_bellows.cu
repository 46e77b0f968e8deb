// K8 -- bellows veto: L3Localizer::TrackAFeature (L3Localizer.cpp:473-543),
// GetDiffROI (:462-470) and the block :292-390 of CalculateInitialBubbleParams.
//
// TrackAFeature is cv::matchTemplate(CV_TM_CCORR_NORMED) of a ~557x178 template over the whole 1050x1680 frame,
// min-max normalisation, first maximum, 3x3 value-weighted centroid.  Here:
//   k8_sq_h / k8_sq_v   exact window sums of squares  W(y,x) = sum I^2   (u32 sliding rows, u64 sliding columns)
//   k8_ncc_kernel       exact integer numerators      N(y,x) = sum I*T   with IDP.4A (4 byte products per instruction),
//                       and the quotient N / (sqrt(W) * |T|) in double with OpenCV's clamps -> float map
//   k8_minmax / k8_argmax   min, max, and the raster-first maximum of the NORMALISED map (normalisation can merge
//                       neighbours of the maximum into the same float, and minMaxLoc then takes the first)
// The 3x3 centroid is nine float operations on the host, in the reference's order (cv::Point2f arithmetic).
// OpenCV evaluates the numerator with a float32 DFT, so its map differs from this one in the last bits
// (about 1e-6); the position is compared with the oracle under a stated tolerance (tests/test_gpu_bellows.py), and
// everything after the two positions (paste, ROI ProcessFrame, subtraction, re-binarisation) is integer-exact.
//
// k8_ncc_kernel: a CTA of 128 threads owns a 16-row x 128-column tile of the map and walks down the image rows
// that tile needs (16 + th - 1).  The whole template sits in shared memory (th x pitch bytes, <= 200 KB); the
// current image row is staged as FOUR byte-shifted copies so that every dp4a operand is an aligned 32-bit word:
// output column x = 4m + c reads copy c at words m .. m + tw/4.  A thread computes 16 adjacent outputs of one map
// row (4 shifts x 4 words) and per 16 template pixels issues 9 LDS.128 for 64 dp4a.  u32 accumulators are folded
// into u64 every 128 rows (128 rows x 256 px x 255^2 < 2^32).
#include <math.h>

#include "abub_internal.cuh"

namespace {

constexpr int NCC_TX = 128;     // map columns per CTA
constexpr int NCC_TY = 16;      // map rows per CTA
constexpr int NCC_THREADS = 128;

struct BelArena {
    unsigned int* d_sqh = nullptr;          // [rows][rc] horizontal window sums of squares
    unsigned long long* d_w = nullptr;      // [rr][rc]   window sums of squares
    float* d_map = nullptr;                 // [rr][rc]   CCORR_NORMED
    unsigned long long* d_red = nullptr;    // [0]=min bits, [1]=max bits, [2]=argmax key
    uint8_t* d_over = nullptr;              // overTheSigma of the veto
    size_t cap_sqh = 0, cap_map = 0;
    int rr = 0, rc = 0;                     // size of the last map
};

__global__ void k8_sq_h_kernel(const uint8_t* __restrict__ img, int rows, int cols, int tw, int rc,
                               unsigned int* __restrict__ sqh) {
    // thread = (row, chunk of 64 map columns): first window summed directly, then slid
    const int r = blockIdx.y;
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 64;
    if (x0 >= rc) return;
    const uint8_t* row = img + (size_t)r * cols;
    unsigned int s = 0;
    for (int i = 0; i < tw; i++) { const unsigned int v = row[x0 + i]; s += v * v; }
    const int x1 = min(x0 + 64, rc);
    for (int x = x0;; x++) {
        sqh[(size_t)r * rc + x] = s;
        if (x + 1 >= x1) break;
        const unsigned int o = row[x], n = row[x + tw];
        s += n * n - o * o;
    }
}

__global__ void k8_sq_v_kernel(const unsigned int* __restrict__ sqh, int th, int rr, int rc, unsigned long long* __restrict__ w) {
    // thread = (map column, chunk of 32 map rows)
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y0 = blockIdx.y * 32;
    if (x >= rc || y0 >= rr) return;
    unsigned long long s = 0;
    for (int j = 0; j < th; j++) s += sqh[(size_t)(y0 + j) * rc + x];
    const int y1 = min(y0 + 32, rr);
    for (int y = y0;; y++) {
        w[(size_t)y * rc + x] = s;
        if (y + 1 >= y1) break;
        s += sqh[(size_t)(y + th) * rc + x];
        s -= sqh[(size_t)y * rc + x];
    }
}

// cv::matchTemplate's common_matchTemplate for CV_TM_CCORR_NORMED (one channel): t = sqrt(window sum of squares) *
// |template|; num/t below t, +-1 within 12.5 % above it, else 0.
__device__ __forceinline__ float ccorr_normed(unsigned long long num_i, unsigned long long w_i, double templ_norm) {
    double num = (double)num_i;
    const double t = sqrt((double)w_i) * templ_norm;
    if (fabs(num) < t) num /= t;
    else if (fabs(num) < t * 1.125) num = num > 0 ? 1 : -1;
    else num = 0;
    return (float)num;
}

__global__ void __launch_bounds__(NCC_THREADS, 2)
k8_ncc_kernel(const uint8_t* __restrict__ img, int rows, int cols, const uint8_t* __restrict__ templ, int th, int tp,
              const unsigned long long* __restrict__ w, double templ_norm, int rr, int rc, float* __restrict__ map) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t* s_t = smem;                                              // th * tp bytes
    const int cw = (NCC_TX + tp) / 4;                                 // words per shifted copy
    uint32_t* s_row = reinterpret_cast<uint32_t*>(smem + (size_t)th * tp);   // [2][4][cw]
    const int tid = threadIdx.x, a = tid >> 3, g = tid & 7;
    const int x0 = blockIdx.x * NCC_TX, y0 = blockIdx.y * NCC_TY;

    for (int i = tid; i < th * tp / 16; i += NCC_THREADS)
        reinterpret_cast<uint4*>(s_t)[i] = __ldg(reinterpret_cast<const uint4*>(templ) + i);

    auto load_row = [&](int r, uint32_t* dst) {
        const uint8_t* row = img + (size_t)r * cols;
        for (int idx = tid; idx < 4 * cw; idx += NCC_THREADS) {
            const int c = idx / cw, wd = idx - c * cw;
            const int bx = x0 + 4 * wd + c;
            uint32_t v = 0;
#pragma unroll
            for (int b = 0; b < 4; b++)
                if (bx + b < cols) v |= (uint32_t)__ldg(row + bx + b) << (8 * b);
            dst[idx] = v;
        }
    };

    uint32_t acc[16];
    unsigned long long tot[16];
#pragma unroll
    for (int i = 0; i < 16; i++) { acc[i] = 0; tot[i] = 0; }

    const int nrl = min(NCC_TY + th - 1, rows - y0);
    load_row(y0, s_row);
    __syncthreads();
    for (int rl = 0; rl < nrl; rl++) {
        const uint32_t* buf = s_row + (rl & 1) * 4 * cw;
        const int j = rl - a;
        if (j >= 0 && j < th) {
            const uint4* T = reinterpret_cast<const uint4*>(s_t + (size_t)j * tp);
            for (int q = 0; q < tp / 16; q++) {
                const uint4 t4 = T[q];
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const uint4* wp = reinterpret_cast<const uint4*>(buf + c * cw + 4 * g + 4 * q);
                    const uint4 w0 = wp[0], w1 = wp[1];
                    const uint32_t W[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                    for (int mm = 0; mm < 4; mm++) {
                        uint32_t s = acc[c * 4 + mm];
                        s = __dp4a(W[mm], t4.x, s);
                        s = __dp4a(W[mm + 1], t4.y, s);
                        s = __dp4a(W[mm + 2], t4.z, s);
                        s = __dp4a(W[mm + 3], t4.w, s);
                        acc[c * 4 + mm] = s;
                    }
                }
            }
        }
        if (rl + 1 < nrl) load_row(y0 + rl + 1, s_row + ((rl + 1) & 1) * 4 * cw);
        if ((rl & 127) == 127) {
#pragma unroll
            for (int i = 0; i < 16; i++) { tot[i] += acc[i]; acc[i] = 0; }
        }
        __syncthreads();
    }
    const int y = y0 + a;
    if (y >= rr) return;
#pragma unroll
    for (int c = 0; c < 4; c++)
#pragma unroll
        for (int mm = 0; mm < 4; mm++) {
            const int x = x0 + 16 * g + 4 * mm + c;
            if (x < rc) {
                const size_t o = (size_t)y * rc + x;
                map[o] = ccorr_normed(tot[c * 4 + mm] + acc[c * 4 + mm], w[o], templ_norm);
            }
        }
}

// map values lie in [0,1]: their bit patterns order like unsigned integers
__global__ void k8_minmax_kernel(const float* __restrict__ map, size_t n, unsigned long long* __restrict__ red) {
    unsigned int lo = 0xFFFFFFFFu, hi = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const unsigned int b = __float_as_uint(map[i]);
        lo = min(lo, b);
        hi = max(hi, b);
    }
    for (int o = 16; o; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xFFFFFFFFu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&red[0], (unsigned long long)lo);
        atomicMax(&red[1], (unsigned long long)hi);
    }
}

// cv::normalize(NORM_MINMAX, 0..1) coefficients for a CV_32F result: scale and shift are rounded to float and
// applied with one fused multiply-add (what cv2 4.13 does; checked against cv2.normalize)
__host__ __device__ inline void minmax_coeffs(float smin_f, float smax_f, float* a, float* b) {
    const double smin = (double)smin_f, smax = (double)smax_f;
    const double scale = (smax - smin > 2.220446049250313e-16) ? 1.0 / (smax - smin) : 0.0;
    *a = (float)scale;
    *b = 0.0f - (float)(smin * (double)*a);
}

// first maximum (raster order) of the normalised map: key = value bits << 32 | ~index
__global__ void k8_argmax_kernel(const float* __restrict__ map, size_t n, unsigned long long* __restrict__ red) {
    float a, b;
    minmax_coeffs(__uint_as_float((unsigned int)red[0]), __uint_as_float((unsigned int)red[1]), &a, &b);
    unsigned long long best = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        // the normalised minimum can come out as a tiny negative number (fused v*a + b): order-preserving bits
        const unsigned int nb = __float_as_uint(fmaf(map[i], a, b));
        const unsigned int ord = (nb & 0x80000000u) ? ~nb : (nb | 0x80000000u);
        const unsigned long long key = ((unsigned long long)ord << 32) | (0xFFFFFFFFu - (unsigned int)i);
        best = max(best, key);
    }
    for (int o = 16; o; o >>= 1) best = max(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
    if ((threadIdx.x & 31) == 0) atomicMax(&red[2], best);
}

__device__ __forceinline__ int reflect101(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * (n - 1) - p;
    return p;
}

// Block :347-363: both frames are zero except for the template pasted at (tx,ty) / (px,py); ProcessFrame
// (AnalyzerUnit.cpp:346-377) restricted to the ROI (blur reflects at the ROI's own edges), then
// overTheSigma -= diff (saturating).  One thread per ROI pixel; the 25 taps are evaluated directly.
__global__ void k8_veto_kernel(uint8_t* __restrict__ over, const uint8_t* __restrict__ sigma, int cols,
                               const uint8_t* __restrict__ templ, int th, int tw, int tp, int tx, int ty, int px, int py,
                               int rx, int ry, int rw, int rh) {
    const int u = blockIdx.x * blockDim.x + threadIdx.x, v = blockIdx.y * blockDim.y + threadIdx.y;
    if (u >= rw || v >= rh) return;
    const int kw[5] = {1, 4, 6, 4, 1};
    int sp = 0, sn = 0;
#pragma unroll
    for (int dj = -2; dj <= 2; dj++) {
        const int gy = ry + reflect101(v + dj, rh);
#pragma unroll
        for (int di = -2; di <= 2; di++) {
            const int gx = rx + reflect101(u + di, rw);
            int a = 0, b = 0;
            if (gx >= tx && gx < tx + tw && gy >= ty && gy < ty + th) a = templ[(size_t)(gy - ty) * tp + (gx - tx)];
            if (gx >= px && gx < px + tw && gy >= py && gy < py + th) b = templ[(size_t)(gy - py) * tp + (gx - px)];
            const int t = min(6 * (int)sigma[(size_t)gy * cols + gx], 255);
            const int wgt = kw[dj + 2] * kw[di + 2];
            sp += wgt * max(a - b - t, 0);
            sn += wgt * max(b - a - t, 0);
        }
    }
    const int d = abs(((sp + 128) >> 8) - ((sn + 128) >> 8));
    const size_t o = (size_t)(ry + v) * cols + (rx + u);
    over[o] = (uint8_t)max((int)over[o] - d, 0);
}

int grow_bytes(abub_ctx* ctx, void** p, size_t bytes) {
    if (*p) cudaFree(*p);
    *p = nullptr;
    ABUB_CHECK_CUDA(ctx, cudaMalloc(p, bytes));
    return ABUB_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------
static BelArena* bel_of(abub_ctx* ctx) {
    if (!ctx->bel) ctx->bel = new BelArena();
    return reinterpret_cast<BelArena*>(ctx->bel);
}

void abub_bel_free(abub_ctx* ctx) {
    BelArena* a = reinterpret_cast<BelArena*>(ctx->bel);
    if (!a) return;
    cudaFree(a->d_sqh); cudaFree(a->d_w); cudaFree(a->d_map); cudaFree(a->d_red); cudaFree(a->d_over);
    delete a;
    ctx->bel = nullptr;
}

static size_t ncc_smem_bytes(int th, int tp) { return (size_t)th * tp + (size_t)2 * 4 * ((NCC_TX + tp) / 4) * 4; }

extern "C" int abub_set_bellows_template(abub_ctx* ctx, int cam, const uint8_t* templ, int trows, int tcols) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams, "cam");
    cudaSetDevice(ctx->device);
    abub_cam_state& cs = ctx->cam[cam];
    cudaFree(cs.d_templ);
    cs.d_templ = nullptr;
    cs.templ_rows = cs.templ_cols = cs.templ_pitch = 0;
    if (!templ) return ABUB_OK;                                     // template not loadable (L3Localizer.cpp:297-301)
    ABUB_REQUIRE(ctx, trows >= 1 && tcols >= 1 && trows <= ctx->rows && tcols <= ctx->cols, "template larger than the frame");
    const int tp = (tcols + 15) & ~15;
    ABUB_REQUIRE(ctx, ncc_smem_bytes(trows, tp) <= 227 * 1024, "template does not fit shared memory (rows x pitch <= ~200 KB)");
    std::vector<uint8_t> padded((size_t)trows * tp, 0);
    double s2 = 0;
    for (int r = 0; r < trows; r++)
        for (int c = 0; c < tcols; c++) {
            const uint8_t v = templ[(size_t)r * tcols + c];
            padded[(size_t)r * tp + c] = v;
            s2 += (double)v * v;
        }
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_templ, padded.size()));
    ABUB_CHECK_CUDA(ctx, cudaMemcpy(cs.d_templ, padded.data(), padded.size(), cudaMemcpyHostToDevice));
    cs.templ_rows = trows;
    cs.templ_cols = tcols;
    cs.templ_pitch = tp;
    cs.templ_sum2 = s2;
    return ABUB_OK;
}

static int ncc_map(abub_ctx* ctx, BelArena* a, const abub_cam_state& cs, const uint8_t* d_img) {
    const int th = cs.templ_rows, tw = cs.templ_cols, tp = cs.templ_pitch;
    const int rr = ctx->rows - th + 1, rc = ctx->cols - tw + 1;
    int rc_;
    const size_t need_sqh = (size_t)ctx->rows * rc, need_map = (size_t)rr * rc;
    if (need_sqh > a->cap_sqh) {
        if ((rc_ = grow_bytes(ctx, (void**)&a->d_sqh, need_sqh * sizeof(unsigned int)))) return rc_;
        a->cap_sqh = need_sqh;
    }
    if (need_map > a->cap_map) {
        if ((rc_ = grow_bytes(ctx, (void**)&a->d_w, need_map * sizeof(unsigned long long)))) return rc_;
        if ((rc_ = grow_bytes(ctx, (void**)&a->d_map, need_map * sizeof(float)))) return rc_;
        a->cap_map = need_map;
    }
    if (!a->d_red) ABUB_CHECK_CUDA(ctx, cudaMalloc(&a->d_red, 4 * sizeof(unsigned long long)));
    cudaStream_t st = ctx->stream;
    ABUB_L(ctx), k8_sq_h_kernel<<<dim3(((rc + 63) / 64 + 63) / 64, ctx->rows), 64, 0, st>>>(d_img, ctx->rows, ctx->cols, tw, rc, a->d_sqh);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    ABUB_L(ctx), k8_sq_v_kernel<<<dim3((rc + 127) / 128, (rr + 31) / 32), 128, 0, st>>>(a->d_sqh, th, rr, rc, a->d_w);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    const size_t smem = ncc_smem_bytes(th, tp);
    ABUB_CHECK_CUDA(ctx, cudaFuncSetAttribute(k8_ncc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // templNorm of common_matchTemplate for CCORR_NORMED: sqrt(sum T^2)
    ABUB_L(ctx), k8_ncc_kernel<<<dim3((rc + NCC_TX - 1) / NCC_TX, (rr + NCC_TY - 1) / NCC_TY), NCC_THREADS, smem, st>>>(
        d_img, ctx->rows, ctx->cols, cs.d_templ, th, tp, a->d_w, sqrt(cs.templ_sum2), rr, rc, a->d_map);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    a->rr = rr;
    a->rc = rc;
    return ABUB_OK;
}

extern "C" int abub_bellows_track(abub_ctx* ctx, int slot, int cam, int frame, float* x, float* y, int* peak_xy) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && cam >= 0 && cam < ctx->ncams && x && y, "abub_bellows_track");
    ABUB_REQUIRE(ctx, frame >= 0 && frame < ctx->slots[slot].nframes[cam], "frame out of range");
    const abub_cam_state& cs = ctx->cam[cam];
    ABUB_REQUIRE(ctx, cs.d_templ != nullptr, "no bellows template set for this camera");
    cudaSetDevice(ctx->device);
    BelArena* a = bel_of(ctx);
    const uint8_t* d_img = ctx->slots[slot].d_frames + ((size_t)cam * ctx->max_frames + frame) * ctx->npix;
    int rc_ = ncc_map(ctx, a, cs, d_img);
    if (rc_) return rc_;
    cudaStream_t st = ctx->stream;
    const size_t n = (size_t)a->rr * a->rc;
    const unsigned long long init[4] = {0xFFFFFFFFull, 0, 0, 0};
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(a->d_red, init, sizeof(init), cudaMemcpyHostToDevice, st));
    ABUB_L(ctx), k8_minmax_kernel<<<148 * 4, 256, 0, st>>>(a->d_map, n, a->d_red);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    ABUB_L(ctx), k8_argmax_kernel<<<148 * 4, 256, 0, st>>>(a->d_map, n, a->d_red);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    unsigned long long red[4];
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(red, a->d_red, sizeof(red), cudaMemcpyDeviceToHost, st));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(st));
    const unsigned int idx = 0xFFFFFFFFu - (unsigned int)(red[2] & 0xFFFFFFFFull);
    const int my = (int)(idx / (unsigned)a->rc), mx = (int)(idx % (unsigned)a->rc);
    if (peak_xy) { peak_xy[0] = mx; peak_xy[1] = my; }
    if (!(mx >= 1 && mx < a->rc - 1 && my >= 1 && my < a->rr - 1)) {
        // L3Localizer.cpp:524 reads result.at<float>() around the peak without a bounds check: undefined there
        ctx->err = "bellows template: correlation peak on the edge of the map (undefined behaviour in the reference)";
        return ABUB_E_INVALID;
    }
    float nb[3][3];
    for (int j = 0; j < 3; j++)
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(nb[j], a->d_map + (size_t)(my - 1 + j) * a->rc + (mx - 1), 3 * sizeof(float),
                                             cudaMemcpyDeviceToHost, st));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(st));
    unsigned int lo_b = (unsigned int)red[0], hi_b = (unsigned int)red[1];
    float smin, smax, ca, cb;
    memcpy(&smin, &lo_b, 4);
    memcpy(&smax, &hi_b, 4);
    minmax_coeffs(smin, smax, &ca, &cb);
    // :522-533  i = x offset (outer loop), j = y offset; cv::Point2f * float and += in float
    volatile float sx = 0.0f, sy = 0.0f, total = 0.0f;
    for (int i = -1; i <= 1; i++)
        for (int j = -1; j <= 1; j++) {
            const float v = fmaf(nb[j + 1][i + 1], ca, cb);
            volatile float px_ = (float)(mx + i) * v, py_ = (float)(my + j) * v;
            sx = sx + px_;
            sy = sy + py_;
            total = total + v;
        }
    const float inv = (float)(1.0 / (double)total);                 // :531
    volatile float fx = sx * inv, fy = sy * inv;                    // :532
    *x = fx + 0.0f;                                                 // :538  + TemplateSearchZone.x (= 0)
    *y = fy + 0.0f;
    return ABUB_OK;
}

extern "C" int abub_bellows_map(abub_ctx* ctx, float* out, int cap, int* map_rows, int* map_cols) {
    if (!ctx) return ABUB_E_INVALID;
    BelArena* a = reinterpret_cast<BelArena*>(ctx->bel);
    ABUB_REQUIRE(ctx, a && a->rr > 0, "abub_bellows_map: no map yet (call abub_bellows_track)");
    if (map_rows) *map_rows = a->rr;
    if (map_cols) *map_cols = a->rc;
    if (!out) return ABUB_OK;
    ABUB_REQUIRE(ctx, (size_t)cap >= (size_t)a->rr * a->rc, "abub_bellows_map: capacity");
    cudaSetDevice(ctx->device);
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(out, a->d_map, (size_t)a->rr * a->rc * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ABUB_OK;
}

// Block :303-390 once the two TrackAFeature positions are known.
extern "C" int abub_localize_bellows(abub_ctx* ctx, int slot, int cam, int trig, int pre_trig, int loc_thres,
                                     const float* pos_trig, const float* pos_pre, abub_blob* out, int cap, int* n_out,
                                     int* status, int* roi_out) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && cam >= 0 && cam < ctx->ncams && pos_trig && pos_pre && n_out && status,
                 "abub_localize_bellows");
    const abub_cam_state& cs = ctx->cam[cam];
    ABUB_REQUIRE(ctx, cs.has_background, "background not installed");
    ABUB_REQUIRE(ctx, cs.d_templ != nullptr, "no bellows template set for this camera");
    const int nf = ctx->slots[slot].nframes[cam];
    ABUB_REQUIRE(ctx, trig >= 0 && trig < nf && pre_trig >= 0 && pre_trig < nf, "trig/pre_trig out of range");
    cudaSetDevice(ctx->device);
    *n_out = 0;
    *status = 0;
    const int th = cs.templ_rows, tw = cs.templ_cols;
    // :316-323  push the x positions one pixel apart (float ++ / --)
    volatile float tx = pos_trig[0], ty = pos_trig[1], px = pos_pre[0], py = pos_pre[1];
    if (tx >= px) { tx = tx + 1.0f; px = px - 1.0f; }
    else { tx = tx - 1.0f; px = px + 1.0f; }
    // :347, :349  cv::Rect(float, float, int, int): truncation; a Rect outside the frame throws (-> -6 in AnyCamAnalysis)
    const int itx = (int)tx, ity = (int)ty, ipx = (int)px, ipy = (int)py;
    auto inside = [&](int x, int y, int w, int h) { return 0 <= x && 0 <= w && x + w <= ctx->cols && 0 <= y && 0 <= h && y + h <= ctx->rows; };
    if (!inside(itx, ity, tw, th) || !inside(ipx, ipy, tw, th)) { *status = -6; return ABUB_OK; }
    // GetDiffROI (:462-470): float max/min, int truncation
    const float fmx = tx > px ? tx : px, fmy = ty > py ? ty : py;      // std::max(point1, point2) with point1 = trig
    const int start_x = (int)(tx < px ? px : tx), start_y = (int)(ty < py ? py : ty);
    (void)fmx; (void)fmy;
    volatile float dxf = (px < tx ? px : tx) + (float)tw;
    dxf = dxf - (float)start_x;
    volatile float dyf = (py < ty ? py : ty) + (float)th;
    dyf = dyf - (float)start_y;
    const int delta_x = (int)dxf, delta_y = (int)dyf;
    if (roi_out) { roi_out[0] = start_x; roi_out[1] = start_y; roi_out[2] = delta_x; roi_out[3] = delta_y; }
    // workingFrame(ROI) throws for a ROI outside the frame, GaussianBlur for an empty one
    if (!inside(start_x, start_y, delta_x, delta_y) || delta_x == 0 || delta_y == 0) { *status = -6; return ABUB_OK; }

    BelArena* a = bel_of(ctx);
    if (!a->d_over) ABUB_CHECK_CUDA(ctx, cudaMalloc(&a->d_over, ctx->npix));
    const uint8_t* base = ctx->slots[slot].d_frames + (size_t)cam * ctx->max_frames * ctx->npix;
    // overTheSigma of (trig, preTrig) (:232), not thresholded
    int rc_ = abub_k1_launch_single(ctx, base + (size_t)trig * ctx->npix, base + (size_t)pre_trig * ctx->npix, cs.d_k32, a->d_over,
                                    nullptr, -1);
    if (rc_) return rc_;
    const dim3 blk(32, 8), grd((delta_x + 31) / 32, (delta_y + 7) / 8);
    ABUB_L(ctx), k8_veto_kernel<<<grd, blk, 0, ctx->stream>>>(a->d_over, cs.d_sigma, ctx->cols, cs.d_templ, th, tw, cs.templ_pitch, itx, ity, ipx,
                                                ipy, start_x, start_y, delta_x, delta_y);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return abub_loc_run_prepared(ctx, a->d_over, loc_thres, out, cap, n_out);
}
