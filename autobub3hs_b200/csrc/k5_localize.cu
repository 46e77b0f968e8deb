// K5-K7 -- L3Localizer blob extraction on the device.
// Reference: L3Localizer::CalculateInitialBubbleParams (L3Localizer.cpp:215-290),
// CalculatePostTriggerFrameParams (:764-836), i.e. OpenCV threshold(TOZERO),
// threshold(BINARY|OTSU), blur 3x3, findContours(RETR_EXTERNAL, CHAIN_APPROX_TC89_L1),
// boundingRect, contourArea, moments.
//
// The binarised frames are almost empty (a few blobs), so nothing here labels the whole
// image.  Pipeline per chunk of binarisation jobs (one job = one frame of one analyzer):
//   K5a  genesis jobs : ProcessFrame + TOZERO(loc_thres) + histogram   (k1_processframe.cu)
//   K5b  post jobs    : absdiff(frame,avg) - 6 sigma, 3x3 box, TOZERO(3) + histogram
//   K5c  Otsu threshold per job from the histogram (fp64, sequential as OpenCV)
//   K6   raster-start candidates: foreground pixels with no foreground neighbour earlier in
//        raster order (W, NW, N, NE).  Every 8-connected component has exactly one
//        candidate that is the raster-first pixel of the border it starts (its own first
//        pixel); all other candidates meet an earlier pixel while following their border.
//   sort candidates by (job, descending raster index)  == findContours output order
//   K7a  trace pass 1: follow the border (Suzuki outer-border rule), reject on an earlier
//        pixel, measure the chain length
//   K7b  trace pass 2 into an exactly-sized arena + TC89_L1 approximation + polygon
//        boundingRect / area / moments
//   K7c  RETR_EXTERNAL: a component is dropped when its first pixel lies inside another
//        component's outer border (= in one of its holes), decided by crack parity on the
//        stored chain
//   compaction in order -> abub_blob records
#include <cstdlib>
#include <cub/cub.cuh>

#include "abub_internal.cuh"

namespace {

struct Cand {           // one traced border
    int job;
    int first_px;       // y*cols + x of the start pixel
    int len;            // chain length (0: isolated pixel)
    int accepted;       // start is the raster-first pixel of its border
    int minx, maxx, miny, maxy;   // bbox of the chain
    long long arena_off;
    int external;
    int pad;
};

struct LocArena {
    // job-level
    abub_loc_job* d_jobs = nullptr;  abub_loc_job* h_jobs = nullptr;  int jobs_cap = 0;
    uint8_t* d_tz = nullptr;         size_t tz_cap = 0;       // jobs_cap * npix
    size_t scratch_cap = 0;                                     // bytes of TOZERO'd frames one chunk may take (set at first use)
    uint8_t* d_segflags = nullptr;                              // jobs_cap * rows * nxw (see abub_loc_job::segflags)
    uint32_t* d_hist = nullptr;                                 // jobs_cap * 256
    int* d_thr = nullptr;                                       // jobs_cap
    // candidates
    unsigned long long* d_keys = nullptr; unsigned long long* d_keys_alt = nullptr; int cand_cap = 0;
    Cand* d_cand = nullptr;
    int* d_counters = nullptr;       // [0]=n candidates, [1]=n accepted, [2]=n external, [3]=overflow
    long long* d_scan = nullptr;     // arena offsets (exclusive scan)
    int* d_iscan = nullptr;
    void* d_cub_tmp = nullptr; size_t cub_tmp_bytes = 0;
    // chain arena: per element {x:int16,y:int16} + code/flags byte + S byte + K u16 + removed byte
    uint8_t* d_chain = nullptr; size_t chain_cap = 0;          // bytes
    abub_blob* d_blobs = nullptr; int blob_cap = 0;
    int* d_npf = nullptr;            // per job count of emitted blobs
    // parity tap of the last call
    int last_njobs = 0;
    // pinned read-back words (counters[8], then the chain total as two ints) and the event the host waits on: waiting for an
    // event instead of the stream lets work queued BEHIND the read-back (a deferred trigger search, ctx.cu) keep the GPU busy
    int* h_ctr = nullptr;
    cudaEvent_t ev = nullptr;
};


// ------------------------------------------------------------------------------------
// K5b: post-trigger frames
// ------------------------------------------------------------------------------------
__device__ __forceinline__ int refl(int p, int n) {
    if (p < 0) p = -p;
    if (p >= n) p = 2 * (n - 1) - p;
    return min(max(p, 0), n - 1);
}

constexpr int PT_W = 64, PT_H = 16;

__global__ void __launch_bounds__(256)
k5_post_kernel(const abub_loc_job* __restrict__ jobs, int rows, int cols) {
    __shared__ uint8_t d[(PT_H + 2)][PT_W + 2 + 2];
    __shared__ uint32_t sh_hist[256];
    const abub_loc_job job = jobs[blockIdx.z];
    if (job.genesis) return;
    const int x0 = blockIdx.x * PT_W, y0 = blockIdx.y * PT_H;
    sh_hist[threadIdx.x] = 0;
    for (int i = threadIdx.x; i < (PT_H + 2) * (PT_W + 2); i += 256) {
        const int ly = i / (PT_W + 2), lx = i % (PT_W + 2);
        const int y = refl(y0 + ly - 1, rows), x = refl(x0 + lx - 1, cols);
        const size_t p = (size_t)y * cols + x;
        const int v = abs((int)job.a[p] - (int)job.b[p]) - (int)job.t8[p];   // L3Localizer.cpp:779-782
        d[ly][lx] = (uint8_t)max(v, 0);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < PT_H * PT_W; i += 256) {
        const int ly = i / PT_W, lx = i % PT_W;
        const int y = y0 + ly, x = x0 + lx;
        if (y < rows && x < cols) {
            int s = 0;
#pragma unroll
            for (int j = 0; j < 3; j++) s += d[ly + j][lx] + d[ly + j][lx + 1] + d[ly + j][lx + 2];
            int v = (s + 4) / 9;                     // cv::blur 3x3 (:785)
            if (v <= job.tozero) v = 0;              // THRESH_TOZERO (:786)
            job.tz[(size_t)y * cols + x] = (uint8_t)v;
            if (v) atomicAdd(&sh_hist[v], 1u);
        }
    }
    __syncthreads();
    const uint32_t c = sh_hist[threadIdx.x];
    if (c && threadIdx.x) atomicAdd(&job.hist[threadIdx.x], c);
}

// Streaming form of K5b for pitches that allow 8-byte loads (cols % 8 == 0): one warp owns a strip of 240 output
// columns (lane l holds the 8 pixels x0-8+8l.., lanes 0/31 are the halo) and walks down PS_TH rows.
//   d   = relu(|a - avg| - t)                     VABSDIFF4, widened to two pixels per word (PRMT), then ONE
//                                                 VIADDMNMX.S16x2.RELU per word against the per-camera plane -t (s16)
//   3x3 = vertical sum of three rows in 16-bit lanes, horizontal sum after taking the halo words from the neighbour
//         lanes (SHFL) and one PRMT per word for the odd alignment
//   v   = (sum + 4) / 9  (cv::blur rounding), THRESH_TOZERO -- only evaluated when some sum in the lane can exceed
//         the threshold, which is rare: frames are almost empty after the 6-sigma cut
// After the 6-sigma cut against the TRAINED MEAN most 256-pixel rows hold nothing at all (unlike the trigger's frame
// pairs, K1): one vote per row tells whether the warp's row is all zero, and when this row and the two before it are,
// the 3x3 sums are zero without being formed -- the vertical / horizontal passes and the threshold test are skipped.
// The frame rows are the only operand that comes from DRAM; at ~1 us of latency a warp must keep kilobytes of them in
// flight, far more than a register prefetch holds: they go through a per-warp ring of PS_D rows in shared memory filled by
// cp.async (each lane copies, and later reads, its own 8 bytes: no barrier, only cp.async.wait_group).  avg and -t come
// from L2 and stay on a two-row register prefetch.
// 3 bytes read (avg and -t from L2) + at most 1 byte written per pixel.
constexpr int PS_TH = 64;
constexpr int PS_WARPS = 4;
constexpr int PS_D = 8;              // frame rows in flight per warp

__device__ __forceinline__ void ps_cp_async8(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void ps_cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void ps_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void ps_cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ uint2 ps_ld8(const uint8_t* p) { return __ldg(reinterpret_cast<const uint2*>(p)); }

struct PsSlot {             // one frame row of one warp in the ring: what lane l needs of the three planes
    uint2 a[32];            // frame[trig + k]
    uint2 m[32];            // TrainedAvgImage
    uint4 t[32];            // -min(6 sigma, 255), s16 per pixel
};

__global__ void __launch_bounds__(32 * PS_WARPS)
k5_post_stream_kernel(const abub_loc_job* __restrict__ jobs, int rows, int cols, int nxw, int nstrips) {
    __shared__ uint32_t sh_hist[256];
    __shared__ __align__(16) PsSlot sh_ring[PS_WARPS][PS_D];
    const abub_loc_job job = jobs[blockIdx.z];
    if (job.genesis) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sh_hist[i] = 0;
    __syncthreads();
    const int unit = blockIdx.x * PS_WARPS + warp;
    if (unit < nxw * nstrips) {
        const int xw = unit % nxw;
        const int x0 = xw * 240, y0 = (unit / nxw) * PS_TH;
        const int gx = x0 - 8 + lane * 8;
        const int gxl = min(max(gx, 0), cols - 8);
        const bool out_lane = lane >= 1 && lane <= 30 && gx < cols;
        const bool left_edge = (gx == 0), right_edge = (gx + 8 == cols);
        const int nout = min(PS_TH, rows - y0);
        const int nrow = nout + 2;                                      // image rows y0-1 .. y0+nout
        const uint32_t lim = 9u * (uint32_t)job.tozero + 5u;            // (sum+4)/9 <= tozero  <=>  sum < lim
        // Rows are consecutive in memory except for BORDER_REFLECT_101 at the image's top (row -1 -> 1) and bottom
        // (row `rows` -> rows-2): the three source pointers of the row being issued advance by one pitch, or step back.
        const int yfirst = (y0 == 0) ? 1 : y0 - 1;
        const uint8_t* pa = job.a + (size_t)yfirst * cols + gxl;
        const uint8_t* pm = job.b + (size_t)yfirst * cols + gxl;
        const int16_t* pt = job.nt16 + (size_t)yfirst * cols + gxl;
        int yi = yfirst;                                                // image row the pointers stand on
        int ri = 0;                                                     // rows issued so far
        PsSlot* const ring = sh_ring[warp];
        auto issue = [&]() {                                            // row ri -> ring slot ri % PS_D; one cp.async group per row
            if (ri < nrow) {
                PsSlot& sl = ring[ri & (PS_D - 1)];
                ps_cp_async8(&sl.a[lane], pa);
                ps_cp_async8(&sl.m[lane], pm);
                ps_cp_async16(&sl.t[lane], pt);
                // next image row: y0-1+ri+1, reflected
                const int ynext = y0 + ri;
                const int yn = (ynext >= rows) ? 2 * (rows - 1) - ynext : ynext;     // ynext >= 0 always (ri + 1 >= 1)
                const ptrdiff_t step = (ptrdiff_t)(yn - yi) * cols;
                pa += step; pm += step; pt += step;
                yi = yn;
            }
            ps_cp_commit();
            ri++;
        };
#pragma unroll
        for (int i = 0; i < PS_D; i++) issue();
        uint8_t* flagp = job.segflags + (size_t)y0 * nxw + xw;          // flag of output row y0 + rl - 2
        uint32_t zhist = 0;                                             // bit 0 / 1 / 2: this row / the one before / two before holds something (warp-uniform)
        uint32_t wa[4] = {0, 0, 0, 0}, wb[4] = {0, 0, 0, 0}, wc[4];    // three consecutive rows, widened (2 px per word); roles rotate

        // one image row: wn = this row, w1 = the row before, w2 = two rows before
        auto do_row = [&](const int rl, uint32_t (&wn)[4], const uint32_t (&w1)[4], const uint32_t (&w2)[4]) {
            ps_cp_wait<PS_D - 1>();                                     // the oldest group = row rl has landed
            const PsSlot& sl = ring[rl & (PS_D - 1)];
            const uint2 a = sl.a[lane], m = sl.m[lane];
            const uint4 t = sl.t[lane];
            issue();                                                    // the slot is free again
            const uint32_t d0 = __vabsdiffu4(a.x, m.x), d1 = __vabsdiffu4(a.y, m.y);   // L3Localizer.cpp:779
            wn[0] = __viaddmax_s16x2_relu(__byte_perm(d0, 0, 0x4140), t.x, t.x);        // :782  - 6 sigma, saturating
            wn[1] = __viaddmax_s16x2_relu(__byte_perm(d0, 0, 0x4342), t.y, t.y);
            wn[2] = __viaddmax_s16x2_relu(__byte_perm(d1, 0, 0x4140), t.z, t.z);
            wn[3] = __viaddmax_s16x2_relu(__byte_perm(d1, 0, 0x4342), t.w, t.w);
            const bool row_nz = __any_sync(0xffffffffu, (wn[0] | wn[1] | wn[2] | wn[3]) != 0u);
            zhist = (zhist << 1) | (row_nz ? 1u : 0u);
            if (rl < 2) return;
            const uint8_t prev = *flagp;                                // same byte for every lane
            uint2 o = make_uint2(0u, 0u);
            bool nz = false;
            if (zhist & 7u) {                                           // some pixel over the cut in these three rows
                uint32_t W[6];
#pragma unroll
                for (int k = 0; k < 4; k++) W[k + 1] = wn[k] + w1[k] + w2[k];
                W[0] = __shfl_up_sync(0xffffffffu, W[4], 1);
                W[5] = __shfl_down_sync(0xffffffffu, W[1], 1);
                if (left_edge) W[0] = __byte_perm(W[1], 0, 0x3232);     // column -1 -> column 1 (hi lane is what is read)
                if (right_edge) W[5] = __byte_perm(W[4], 0, 0x1010);    // column cols -> column cols-2 (lo lane is read)
                uint32_t H[4], mx = 0;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t sl_ = __byte_perm(W[k], W[k + 1], 0x5432);     // (x-1, x) pairs
                    const uint32_t sr_ = __byte_perm(W[k + 1], W[k + 2], 0x5432); // (x+1, x+2) pairs
                    H[k] = sl_ + W[k + 1] + sr_;
                    mx = __vmaxu2(mx, H[k]);
                }
                if (out_lane && ((mx & 0xFFFFu) >= lim || (mx >> 16) >= lim)) {
                    uint32_t ob[8];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        ob[2 * k] = (((H[k] & 0xFFFFu) + 4u) * 7282u) >> 16;       // (sum+4)/9, exact for sum <= 2295
                        ob[2 * k + 1] = (((H[k] >> 16) + 4u) * 7282u) >> 16;
                    }
#pragma unroll
                    for (int c = 0; c < 8; c++) {
                        if ((int)ob[c] <= job.tozero) ob[c] = 0;                   // THRESH_TOZERO (:786)
                        if (ob[c]) atomicAdd(&sh_hist[ob[c]], 1u);
                    }
                    o.x = ob[0] | (ob[1] << 8) | (ob[2] << 16) | (ob[3] << 24);
                    o.y = ob[4] | (ob[5] << 8) | (ob[6] << 16) | (ob[7] << 24);
                }
                nz = __any_sync(0xffffffffu, (o.x | o.y) != 0u);
            }
            // the segment is stored when it holds something, or when what the memory holds is not known to be zero
            if (nz || (prev & 2)) {
                if (out_lane) *reinterpret_cast<uint2*>(job.tz + (size_t)(y0 + rl - 2) * cols + gx) = o;
            }
            const uint8_t nf = nz ? 3 : 0;
            if (lane == 0 && nf != prev) *flagp = nf;
            flagp += nxw;
        };
        // three rows per trip: the roles (this row, the one before, two before) rotate without a register move
#pragma unroll 1
        for (int rl = 0; rl < nrow; rl += 3) {
            do_row(rl, wc, wb, wa);
            if (rl + 1 < nrow) do_row(rl + 1, wa, wc, wb);
            if (rl + 2 < nrow) do_row(rl + 2, wb, wa, wc);
        }
        ps_cp_wait<0>();
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        const uint32_t c = sh_hist[i];
        if (c && i) atomicAdd(&job.hist[i], c);
    }
}

// ------------------------------------------------------------------------------------
// K5c: Otsu (OpenCV getThreshVal_Otsu_8u), one thread per job; also fills bin 0
// ------------------------------------------------------------------------------------
// Prepared jobs (genesis == 2; the bellows veto, k8_bellows.cu): `a` already holds overTheSigma; THRESH_TOZERO at
// job.tozero (L3Localizer.cpp:365) + histogram of the non-zero bytes.  grid.y = job, other jobs return.
__global__ void __launch_bounds__(256) k5_prepared_kernel(const abub_loc_job* __restrict__ jobs, size_t npix) {
    const abub_loc_job job = jobs[blockIdx.y];
    if (job.genesis != 2) return;
    __shared__ uint32_t sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    const int t = job.tozero;
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += (size_t)gridDim.x * blockDim.x) {
        const uint8_t v = job.a[p];
        const uint8_t o = v > t ? v : 0;
        job.tz[p] = o;
        if (o) atomicAdd(&sh[o], 1u);
    }
    __syncthreads();
    if (threadIdx.x > 0 && sh[threadIdx.x]) atomicAdd(&job.hist[threadIdx.x], sh[threadIdx.x]);
}

__global__ void k5_otsu_kernel(const abub_loc_job* __restrict__ jobs, int* __restrict__ thr, int njobs, uint32_t npix) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= njobs) return;
    uint32_t* h = jobs[j].hist;
    uint32_t s = 0;
    for (int i = 1; i < 256; i++) s += h[i];
    h[0] = npix - s;
    const double scale = __ddiv_rn(1.0, (double)npix);
    double mu = 0;
    for (int i = 0; i < 256; i++) mu = __dadd_rn(mu, __dmul_rn((double)i, (double)h[i]));
    mu = __dmul_rn(mu, scale);
    double mu1 = 0, q1 = 0, max_sigma = 0;
    int t = 0;
    for (int i = 0; i < 256; i++) {
        const double p_i = __dmul_rn((double)h[i], scale);
        mu1 = __dmul_rn(mu1, q1);
        q1 = __dadd_rn(q1, p_i);
        const double q2 = __dsub_rn(1.0, q1);
        if (fmin(q1, q2) < 1.1920928955078125e-07 || fmax(q1, q2) > 1. - 1.1920928955078125e-07) continue;
        mu1 = __ddiv_rn(__dadd_rn(mu1, __dmul_rn((double)i, p_i)), q1);
        const double mu2 = __ddiv_rn(__dsub_rn(mu, __dmul_rn(q1, mu1)), q2);
        const double dm = __dsub_rn(mu1, mu2);
        const double sg = __dmul_rn(__dmul_rn(__dmul_rn(q1, q2), dm), dm);
        if (sg > max_sigma) { max_sigma = sg; t = i; }
    }
    thr[j] = t;
}

// ------------------------------------------------------------------------------------
// K6: raster-start candidates
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k6_candidates_kernel(const abub_loc_job* __restrict__ jobs, const int* __restrict__ thr, int rows, int cols,
                     unsigned long long* __restrict__ keys, int* __restrict__ counters, int cand_cap, int job0) {
    const int jl = blockIdx.y;
    const uint8_t* tz = jobs[jl].tz;
    const int t = thr[jl];
    const size_t npix = (size_t)rows * cols;
    if ((cols & 15) == 0) {
        // 16 pixels per load; the TOZERO'd frames are almost empty, so nearly every group is rejected by one OR
        const size_t ngroups = npix / 16;
        for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += (size_t)gridDim.x * blockDim.x) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(tz) + g);
            if ((v.x | v.y | v.z | v.w) == 0u) continue;
            const uint32_t wds[4] = {v.x, v.y, v.z, v.w};
            for (int c = 0; c < 16; c++) {
                if ((int)((wds[c >> 2] >> (8 * (c & 3))) & 0xFFu) <= t) continue;
                const size_t p = g * 16 + c;
                const int y = (int)(p / cols), x = (int)(p % cols);
                bool earlier = false;
                if (x > 0 && tz[p - 1] > t) earlier = true;
                if (y > 0) {
                    const uint8_t* up = tz + p - cols;
                    if (up[0] > t) earlier = true;
                    if (x > 0 && up[-1] > t) earlier = true;
                    if (x + 1 < cols && up[1] > t) earlier = true;
                }
                if (!earlier) {
                    const int slot = atomicAdd(&counters[0], 1);
                    if (slot < cand_cap)
                        keys[slot] = ((unsigned long long)(unsigned)(job0 + jl) << 32) | (unsigned)(0xFFFFFFFFu - (unsigned)p);
                    else
                        counters[3] = 1;
                }
            }
        }
        return;
    }
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += (size_t)gridDim.x * blockDim.x) {
        if (tz[p] <= t) continue;
        const int y = (int)(p / cols), x = (int)(p % cols);
        bool earlier = false;
        if (x > 0 && tz[p - 1] > t) earlier = true;
        if (y > 0) {
            const uint8_t* up = tz + p - cols;
            if (up[0] > t) earlier = true;
            if (x > 0 && up[-1] > t) earlier = true;
            if (x + 1 < cols && up[1] > t) earlier = true;
        }
        if (!earlier) {
            const int slot = atomicAdd(&counters[0], 1);
            if (slot < cand_cap)
                keys[slot] = ((unsigned long long)(unsigned)(job0 + jl) << 32) | (unsigned)(0xFFFFFFFFu - (unsigned)p);
            else
                counters[3] = 1;
        }
    }
}

// Jobs whose frame was written in full (genesis K1 store mode, prepared jobs, the tiled K5b fallback): every
// segment may hold anything.
__global__ void k5_flag_fill_kernel(const abub_loc_job* __restrict__ jobs, size_t nseg, int all_jobs) {
    const abub_loc_job job = jobs[blockIdx.y];
    if (!all_jobs && job.genesis == 0) return;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nseg; i += (size_t)gridDim.x * blockDim.x) job.segflags[i] = 3;
}

// K6 over the segment flags (cols % 16 == 0): one thread per 240-px row segment; only flagged segments are read.
__global__ void __launch_bounds__(256)
k6_candidates_seg_kernel(const abub_loc_job* __restrict__ jobs, const int* __restrict__ thr, int rows, int cols, int nxw,
                         unsigned long long* __restrict__ keys, int* __restrict__ counters, int cand_cap, int job0) {
    const int jl = blockIdx.y;
    const abub_loc_job job = jobs[jl];
    const uint8_t* tz = job.tz;
    const int t = thr[jl];
    const int nseg = rows * nxw;
    for (int sgi = blockIdx.x * blockDim.x + threadIdx.x; sgi < nseg; sgi += gridDim.x * blockDim.x) {
        if (!(job.segflags[sgi] & 1)) continue;
        const int y = sgi / nxw, xs = (sgi - y * nxw) * 240, xe = min(xs + 240, cols);
        for (int xg = xs; xg < xe; xg += 16) {
            const size_t pg = (size_t)y * cols + xg;
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(tz + pg));
            if ((v.x | v.y | v.z | v.w) == 0u) continue;
            const uint32_t wds[4] = {v.x, v.y, v.z, v.w};
            for (int c = 0; c < 16; c++) {
                if ((int)((wds[c >> 2] >> (8 * (c & 3))) & 0xFFu) <= t) continue;
                const size_t p = pg + c;
                const int x = xg + c;
                bool earlier = false;
                if (x > 0 && tz[p - 1] > t) earlier = true;
                if (y > 0) {
                    const uint8_t* up = tz + p - cols;
                    if (up[0] > t) earlier = true;
                    if (x > 0 && up[-1] > t) earlier = true;
                    if (x + 1 < cols && up[1] > t) earlier = true;
                }
                if (!earlier) {
                    const int slot = atomicAdd(&counters[0], 1);
                    if (slot < cand_cap)
                        keys[slot] = ((unsigned long long)(unsigned)(job0 + jl) << 32) | (unsigned)(0xFFFFFFFFu - (unsigned)p);
                    else
                        counters[3] = 1;
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------
// K7: border following / approximation / measurement, one thread per candidate
// ------------------------------------------------------------------------------------
}  // namespace
#include "contour_core.cuh"
namespace {
using namespace abub_contour;

__device__ __forceinline__ void decode_key(unsigned long long key, int job0, int* jl, int* px) {
    *jl = (int)(key >> 32) - job0;
    *px = (int)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull));
}

// One WARP per candidate.  Border following is a chain of dependent pixel reads (up to eight probes per border pixel): from
// global memory that is hundreds of cycles per probe and a kernel as long as its longest contour.  The warp therefore first
// copies a window of the binarised image into shared memory (all lanes, coalesced), then lane 0 follows the border with the
// window answering the probes (pixels outside it still come from the image).  Pass 2 also keeps the chain (points, codes and
// the TC89 work arrays) in shared memory while it is approximated and measured, and the warp copies it to the arena at the
// end (k7_external_kernel reads it for the hole test).
constexpr int K7_WARPS = 4;
constexpr int K7_W1W = 64, K7_W1H = 40;                  // pass 1 window: x in [px-32, px+32), y in [py-1, py+39)
constexpr int K7_W2_BYTES = 6144;                         // pass 2 window: the bounding box + 1 (clipped to this many bytes)
constexpr int K7_CHAIN_MAX = 255;                         // longest chain kept in shared memory (12 bytes per element)

__device__ __forceinline__ void k7_load_window(const Pix& g, uint8_t* win, int wx0, int wy0, int ww, int wh, int lane) {
    for (int idx = lane; idx < ww * wh; idx += 32) {
        const int y = wy0 + idx / ww, x = wx0 + idx % ww;
        win[idx] = g.fg(x, y) ? 1 : 0;
    }
    __syncwarp();
}

__global__ void __launch_bounds__(32 * K7_WARPS)
k7_trace1_kernel(const unsigned long long* __restrict__ keys, int ncand, const abub_loc_job* __restrict__ jobs,
                 const int* __restrict__ thr, int rows, int cols, int job0, Cand* __restrict__ cand,
                 long long* __restrict__ sizes) {
    __shared__ uint8_t sh_win[K7_WARPS][K7_W1W * K7_W1H];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * K7_WARPS + warp;
    if (i >= ncand) return;
    int jl, p;
    decode_key(keys[i], job0, &jl, &p);
    const int px = p % cols, py = p / cols;
    WinPix im{Pix{jobs[jl].tz, thr[jl], rows, cols}, sh_win[warp], px - K7_W1W / 2, py - 1, K7_W1W, K7_W1H};
    k7_load_window(im.g, sh_win[warp], im.wx0, im.wy0, im.ww, im.wh, lane);
    if (lane != 0) return;
    const TraceResult t = trace_border(im, px, py, nullptr, 4 * rows * cols);
    Cand c;
    c.job = jl; c.first_px = p; c.len = t.len; c.accepted = t.accepted;
    c.minx = t.minx; c.maxx = t.maxx; c.miny = t.miny; c.maxy = t.maxy;
    c.arena_off = 0; c.external = 1; c.pad = 0;
    cand[i] = c;
    sizes[i] = t.accepted ? (long long)((ChainView::bytes(t.len) + 15) & ~(size_t)15) : 0;
}

__global__ void __launch_bounds__(32 * K7_WARPS)
k7_trace2_kernel(int ncand, const abub_loc_job* __restrict__ jobs, const int* __restrict__ thr, int rows,
                 int cols, Cand* __restrict__ cand, const long long* __restrict__ offs,
                 uint8_t* __restrict__ arena, abub_blob* __restrict__ tmp) {
    __shared__ uint8_t sh_win[K7_WARPS][K7_W2_BYTES];
    __shared__ __align__(16) uint8_t sh_chain[K7_WARPS][12 * (K7_CHAIN_MAX + 1)];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * K7_WARPS + warp;
    if (i >= ncand) return;
    Cand c = cand[i];
    if (!c.accepted) return;                               // warp-uniform
    c.arena_off = offs[i];
    const int px = c.first_px % cols, py = c.first_px / cols;
    int ww = min(c.maxx - c.minx + 3, 128);
    int wh = min(c.maxy - c.miny + 3, K7_W2_BYTES / ww);
    WinPix im{Pix{jobs[c.job].tz, thr[c.job], rows, cols}, sh_win[warp], c.minx - 1, c.miny - 1, ww, wh};
    if (c.len > 0) k7_load_window(im.g, sh_win[warp], im.wx0, im.wy0, ww, wh, lane);
    const bool in_smem = c.len <= K7_CHAIN_MAX;
    uint8_t* const chain_mem = in_smem ? sh_chain[warp] : arena + c.arena_off;
    if (lane == 0) {
        cand[i].arena_off = c.arena_off;
        ChainView cv(chain_mem, c.len);
        int L;
        if (c.len == 0) {
            cv.P[0] = make_short2((short)px, (short)py);
            cv.removed[0] = 0;
            L = 1;
        } else {
            trace_border(im, px, py, &cv, 4 * rows * cols);
            L = approx_tc89_l1(cv, c.len);
        }
        const PolyMeasure m = measure_alive(cv, L);
        abub_blob b;
        b.k = jobs[c.job].k; b.x = m.x; b.y = m.y; b.w = m.w; b.h = m.h; b.npts = m.npts; b.first_px = c.first_px;
        b.reserved = jobs[c.job].item; b.sum_x = m.sum_x; b.sum_y = m.sum_y;
        b.area = m.area; b.m00 = m.m00; b.m10 = m.m10; b.m01 = m.m01;
        tmp[i] = b;
    }
    __syncwarp();
    if (in_smem) {                                         // the chain goes to the arena (same layout), all lanes
        const int nwords = (int)(ChainView::bytes(c.len) / 4);
        const uint32_t* src = reinterpret_cast<const uint32_t*>(sh_chain[warp]);
        uint32_t* dst = reinterpret_cast<uint32_t*>(arena + c.arena_off);
        for (int w = lane; w < nwords; w += 32) dst[w] = src[w];
    }
}

// RETR_EXTERNAL: drop components whose first pixel lies inside another component's outer border.
// Candidates are sorted by job, so the same-job range is found by walking both ways.
__global__ void k7_external_kernel(int ncand, int cols, Cand* __restrict__ cand, const uint8_t* __restrict__ arena,
                                   int* __restrict__ flags) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ncand) return;
    const Cand c = cand[i];
    int keep = c.accepted;
    if (keep) {
        const int qx = c.first_px % cols, qy = c.first_px / cols;
        for (int dir = -1; dir <= 1 && keep; dir += 2) {
            for (int j = i + dir; j >= 0 && j < ncand && keep; j += dir) {
                const Cand o = cand[j];
                if (o.job != c.job) break;
                if (!o.accepted || o.len == 0) continue;
                if (qx <= o.minx || qx >= o.maxx || qy <= o.miny || qy >= o.maxy) continue;
                const ChainView cv(const_cast<uint8_t*>(arena) + o.arena_off, o.len);
                if (inside_outer_border(cv, o.len, qx, qy)) keep = 0;
            }
        }
    }
    flags[i] = keep;
}

__global__ void k7_emit_kernel(int ncand, const Cand* __restrict__ cand, const int* __restrict__ flags,
                               const int* __restrict__ pos, const abub_blob* __restrict__ tmp, abub_blob* __restrict__ out,
                               int out_cap, int* __restrict__ npf, int* __restrict__ counters) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ncand || !flags[i]) return;
    const int o = pos[i];
    if (o < out_cap) out[o] = tmp[i];
    atomicAdd(&npf[cand[i].job], 1);
    atomicAdd(&counters[2], 1);
}

__global__ void k5_binary_kernel(const uint8_t* __restrict__ tz, const int* __restrict__ thr, int job, uint8_t* __restrict__ out,
                                 size_t npix) {
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < npix) out[p] = tz[(size_t)job * npix + p] > thr[job] ? 255 : 0;
}

template <typename T>
int grow(abub_ctx* ctx, T** p, size_t n_elems) {
    if (*p) cudaFree(*p);
    *p = nullptr;
    ABUB_CHECK_CUDA(ctx, cudaMalloc((void**)p, n_elems * sizeof(T)));
    return ABUB_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------
// host orchestration
// ------------------------------------------------------------------------------------
static LocArena* arena_of(abub_ctx* ctx) { return reinterpret_cast<LocArena*>(ctx->loc); }

const abub_loc_job* abub_loc_host_jobs(abub_ctx* ctx) { return arena_of(ctx)->h_jobs; }

int abub_loc_init(abub_ctx* ctx) {
    LocArena* a = new LocArena();
    ctx->loc = a;
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&a->d_counters, 8 * sizeof(int)));
    ABUB_CHECK_CUDA(ctx, cudaHostAlloc(&a->h_ctr, 16 * sizeof(int), cudaHostAllocDefault));
    ABUB_CHECK_CUDA(ctx, cudaEventCreateWithFlags(&a->ev, cudaEventDisableTiming));
    return ABUB_OK;
}

void abub_loc_free(abub_ctx* ctx) {
    LocArena* a = arena_of(ctx);
    if (!a) return;
    cudaFree(a->d_jobs); cudaFree(a->d_tz); cudaFree(a->d_segflags); cudaFree(a->d_hist); cudaFree(a->d_thr);
    cudaFree(a->d_keys); cudaFree(a->d_keys_alt); cudaFree(a->d_cand); cudaFree(a->d_counters);
    cudaFree(a->d_scan); cudaFree(a->d_iscan); cudaFree(a->d_cub_tmp); cudaFree(a->d_chain);
    cudaFree(a->d_blobs); cudaFree(a->d_npf);
    free(a->h_jobs);
    if (a->h_ctr) cudaFreeHost(a->h_ctr);
    if (a->ev) cudaEventDestroy(a->ev);
    delete a;
    ctx->loc = nullptr;
}

static int ensure_jobs(abub_ctx* ctx, LocArena* a, int njobs) {
    if (njobs <= a->jobs_cap) return ABUB_OK;
    // a regrow frees, reallocates and clears the whole scratch: leave an eighth of slack above a first request so that
    // batches whose job counts differ by a few frames do not trigger it again
    if (a->jobs_cap > 0) njobs += njobs / 8;
    int rc;
    if ((rc = grow(ctx, &a->d_jobs, (size_t)njobs))) return rc;
    if ((rc = grow(ctx, &a->d_tz, (size_t)njobs * ctx->npix))) return rc;
    const size_t nseg = (size_t)ctx->rows * ((ctx->cols + 239) / 240);
    if ((rc = grow(ctx, &a->d_segflags, (size_t)njobs * nseg))) return rc;
    // sparse TOZERO frames: a row segment is written only when it holds a non-zero pixel or may still hold one from an
    // earlier call, so the scratch starts out all zero with every flag clear
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(a->d_tz, 0, (size_t)njobs * ctx->npix, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(a->d_segflags, 0, (size_t)njobs * nseg, ctx->stream));
    if ((rc = grow(ctx, &a->d_hist, (size_t)njobs * 256))) return rc;
    if ((rc = grow(ctx, &a->d_thr, (size_t)njobs))) return rc;
    if ((rc = grow(ctx, &a->d_npf, (size_t)njobs))) return rc;
    free(a->h_jobs);
    a->h_jobs = (abub_loc_job*)malloc(sizeof(abub_loc_job) * njobs);
    a->jobs_cap = njobs;
    return ABUB_OK;
}

static int ensure_cands(abub_ctx* ctx, LocArena* a, int n) {
    if (n <= a->cand_cap) return ABUB_OK;
    int rc;
    if ((rc = grow(ctx, &a->d_keys, (size_t)n))) return rc;
    if ((rc = grow(ctx, &a->d_keys_alt, (size_t)n))) return rc;
    if ((rc = grow(ctx, &a->d_cand, (size_t)n))) return rc;
    if ((rc = grow(ctx, &a->d_scan, (size_t)n + 1))) return rc;
    if ((rc = grow(ctx, &a->d_iscan, (size_t)2 * n + 2))) return rc;
    if ((rc = grow(ctx, &a->d_blobs, (size_t)2 * n))) return rc;      // tmp records [0,n) + ordered output [n,2n)
    a->blob_cap = n;
    size_t t1 = 0, t2 = 0, t3 = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, t1, a->d_keys, a->d_keys_alt, 2 * n + 2);
    cub::DeviceScan::ExclusiveSum(nullptr, t2, a->d_scan, a->d_scan, 2 * n + 2);
    cub::DeviceScan::ExclusiveSum(nullptr, t3, a->d_iscan, a->d_iscan, 2 * n + 2);
    size_t t = t1 > t2 ? t1 : t2;
    if (t3 > t) t = t3;
    cudaFree(a->d_cub_tmp);
    a->d_cub_tmp = nullptr;
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&a->d_cub_tmp, t + 256));
    a->cub_tmp_bytes = t + 256;
    a->cand_cap = n;
    return ABUB_OK;
}

// one chunk of jobs already described in a->h_jobs[0..njobs)
static int run_chunk(abub_ctx* ctx, LocArena* a, int njobs, std::vector<abub_blob>& blobs_out, std::vector<int>& npf_out,
                     bool has_prepared = false) {
    cudaStream_t st = ctx->stream;
    const size_t npix = ctx->npix;
    int rc;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(a->d_jobs, a->h_jobs, sizeof(abub_loc_job) * njobs, cudaMemcpyHostToDevice, st));
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(a->d_hist, 0, (size_t)njobs * 256 * sizeof(uint32_t), st));
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(a->d_npf, 0, (size_t)njobs * sizeof(int), st));
    if ((rc = abub_k1_launch_genesis_jobs(ctx, a->d_jobs, njobs))) return rc;
    {
    abub_prof_scope prof(ctx, ABUB_PROF_K5B);
    static const bool no_tma = getenv("ABUB_K5B_STREAM") != nullptr;       // A/B switch
    if (!has_prepared && !no_tma && abub_k5b_tma_supported(ctx)) {
        if ((rc = abub_k5b_launch_tma(ctx, a->d_jobs, njobs))) return rc;
    } else if ((ctx->cols % 8) == 0 && ctx->cols >= 16) {
        const int nxw = (ctx->cols + 239) / 240, nstrips = (ctx->rows + PS_TH - 1) / PS_TH;
        dim3 grid((nxw * nstrips + PS_WARPS - 1) / PS_WARPS, 1, njobs);
        ABUB_L(ctx), k5_post_stream_kernel<<<grid, 32 * PS_WARPS, 0, st>>>(a->d_jobs, ctx->rows, ctx->cols, nxw, nstrips);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    } else {
        dim3 grid((ctx->cols + PT_W - 1) / PT_W, (ctx->rows + PT_H - 1) / PT_H, njobs);
        ABUB_L(ctx), k5_post_kernel<<<grid, 256, 0, st>>>(a->d_jobs, ctx->rows, ctx->cols);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    }
    if (has_prepared) {
        ABUB_L(ctx), k5_prepared_kernel<<<dim3(148 * 2, njobs), 256, 0, st>>>(a->d_jobs, npix);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    const int nxw_f = (ctx->cols + 239) / 240;
    const bool stream_k5b = (ctx->cols % 8) == 0 && ctx->cols >= 16;
    {
    abub_prof_scope prof(ctx, ABUB_PROF_K5C);
    // frames written in full have every segment flagged; the TMA genesis kernel keeps the flags itself
    static const bool k5a_legacy = getenv("ABUB_K5A_LEGACY") != nullptr;
    if (has_prepared || !stream_k5b || k5a_legacy || !abub_k5a_tma_supported(ctx)) {
        ABUB_L(ctx), k5_flag_fill_kernel<<<dim3(4, njobs), 256, 0, st>>>(a->d_jobs, (size_t)ctx->rows * nxw_f, stream_k5b ? 0 : 1);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    ABUB_L(ctx), k5_otsu_kernel<<<(njobs + 63) / 64, 64, 0, st>>>(a->d_jobs, a->d_thr, njobs, (uint32_t)npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    a->last_njobs = njobs;

    int counters[8];
    if (a->cand_cap == 0 && (rc = ensure_cands(ctx, a, 1 << 16))) return rc;
    for (;;) {
        ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(a->d_counters, 0, 8 * sizeof(int), st));
        {
        abub_prof_scope prof(ctx, ABUB_PROF_K6);
        if ((ctx->cols & 15) == 0 && !getenv("ABUB_K6_DENSE")) {
            const int nseg = ctx->rows * nxw_f;
            ABUB_L(ctx), k6_candidates_seg_kernel<<<dim3((nseg + 255) / 256, njobs), 256, 0, st>>>(a->d_jobs, a->d_thr, ctx->rows, ctx->cols, nxw_f,
                                                                                     a->d_keys, a->d_counters, a->cand_cap, 0);
        } else {
            const int gx = (int)((npix + 256 * 16 - 1) / (256 * 16));
            ABUB_L(ctx), k6_candidates_kernel<<<dim3(gx, njobs), 256, 0, st>>>(a->d_jobs, a->d_thr, ctx->rows, ctx->cols, a->d_keys,
                                                                 a->d_counters, a->cand_cap, 0);
        }
        }
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(a->h_ctr, a->d_counters, sizeof(counters), cudaMemcpyDeviceToHost, st));
        ABUB_CHECK_CUDA(ctx, cudaEventRecord(a->ev, st));
        // the front-end kernels of this chunk are queued: the first half of a deferred trigger search goes behind them
        if ((rc = abub_run_deferred(ctx, 0))) return rc;
        ABUB_CHECK_CUDA(ctx, cudaEventSynchronize(a->ev));
        memcpy(counters, a->h_ctr, sizeof(counters));
        if (counters[0] <= a->cand_cap) break;
        if ((rc = ensure_cands(ctx, a, counters[0] + counters[0] / 4 + 1024))) return rc;
    }
    const int ncand = counters[0];
    npf_out.assign(njobs, 0);
    if (ncand == 0) return ABUB_OK;
    size_t tmpb = a->cub_tmp_bytes;
    {
        abub_prof_scope prof(ctx, ABUB_PROF_SORT);
        cub::DeviceRadixSort::SortKeys(a->d_cub_tmp, tmpb, a->d_keys, a->d_keys_alt, ncand, 0, 64, st);
    }
    const int tb = 64, nb = (ncand + tb - 1) / tb;
    const int nbw = (ncand + K7_WARPS - 1) / K7_WARPS;          // the two tracing kernels run one warp per candidate
    {
    abub_prof_scope prof(ctx, ABUB_PROF_K7A);
    ABUB_L(ctx), k7_trace1_kernel<<<nbw, 32 * K7_WARPS, 0, st>>>(a->d_keys_alt, ncand, a->d_jobs, a->d_thr, ctx->rows, ctx->cols, 0, a->d_cand,
                                        a->d_scan);
    }
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    // arena offsets: exclusive scan over ncand+1 entries (the extra entry yields the total)
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(a->d_scan + ncand, 0, sizeof(long long), st));
    tmpb = a->cub_tmp_bytes;
    cub::DeviceScan::ExclusiveSum(a->d_cub_tmp, tmpb, a->d_scan, a->d_scan, ncand + 1, st);
    long long total = 0;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(a->h_ctr + 8, a->d_scan + ncand, sizeof(long long), cudaMemcpyDeviceToHost, st));
    ABUB_CHECK_CUDA(ctx, cudaEventRecord(a->ev, st));
    ABUB_CHECK_CUDA(ctx, cudaEventSynchronize(a->ev));
    memcpy(&total, a->h_ctr + 8, sizeof(long long));
    if ((size_t)total + 64 > a->chain_cap) {
        const size_t want = (size_t)total + (size_t)total / 2 + (1 << 20);
        if ((rc = grow(ctx, &a->d_chain, want))) return rc;
        a->chain_cap = want;
    }
    abub_blob* tmp = a->d_blobs;
    abub_blob* ordered = a->d_blobs + a->blob_cap;
    {
    abub_prof_scope prof(ctx, ABUB_PROF_K7B);
    ABUB_L(ctx), k7_trace2_kernel<<<nbw, 32 * K7_WARPS, 0, st>>>(ncand, a->d_jobs, a->d_thr, ctx->rows, ctx->cols, a->d_cand, a->d_scan, a->d_chain, tmp);
    }
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    int* flags = a->d_iscan;
    int* pos = a->d_iscan + ncand + 1;
    {
    abub_prof_scope prof7c(ctx, ABUB_PROF_K7C);
    ABUB_L(ctx), k7_external_kernel<<<nb, tb, 0, st>>>(ncand, ctx->cols, a->d_cand, a->d_chain, flags);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    tmpb = a->cub_tmp_bytes;
    cub::DeviceScan::ExclusiveSum(a->d_cub_tmp, tmpb, flags, pos, ncand, st);
    ABUB_L(ctx), k7_emit_kernel<<<nb, tb, 0, st>>>(ncand, a->d_cand, flags, pos, tmp, ordered, a->blob_cap, a->d_npf, a->d_counters);
    }
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(a->h_ctr, a->d_counters, sizeof(counters), cudaMemcpyDeviceToHost, st));
    ABUB_CHECK_CUDA(ctx, cudaEventRecord(a->ev, st));
    ABUB_CHECK_CUDA(ctx, cudaEventSynchronize(a->ev));
    memcpy(counters, a->h_ctr, sizeof(counters));
    const int nblob = counters[2];
    const size_t base = blobs_out.size();
    blobs_out.resize(base + nblob);
    // results leave on the stream BEFORE the rest of a deferred trigger search is queued behind them
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(npf_out.data(), a->d_npf, njobs * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (nblob)
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(blobs_out.data() + base, ordered, sizeof(abub_blob) * nblob, cudaMemcpyDeviceToHost, st));
    ABUB_CHECK_CUDA(ctx, cudaEventRecord(a->ev, st));
    ABUB_CHECK_CUDA(ctx, cudaEventSynchronize(a->ev));
    return ABUB_OK;
}

extern "C" int abub_localize_batch(abub_ctx* ctx, const abub_loc_item* items, int n_items, abub_blob* out, int cap,
                                   int* n_per_frame, int* n_total) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, items && n_items >= 1 && n_per_frame && (out || cap == 0), "abub_localize_batch");
    cudaSetDevice(ctx->device);
    LocArena* a = arena_of(ctx);
    const size_t npix = ctx->npix;
    for (int i = 0; i < n_items; i++) {
        const abub_loc_item& it = items[i];
        ABUB_REQUIRE(ctx, it.slot >= 0 && it.slot < ctx->max_slots && it.cam >= 0 && it.cam < ctx->ncams, "item slot/cam");
        ABUB_REQUIRE(ctx, ctx->cam[it.cam].has_background, "background not installed");
        const int nf = ctx->slots[it.slot].nframes[it.cam];
        ABUB_REQUIRE(ctx, it.trig >= 0 && it.trig < nf && it.pre_trig >= 0 && it.pre_trig < nf, "trig/pre_trig out of range");
        ABUB_REQUIRE(ctx, it.n_post >= 0 && it.n_post <= ABUB_MAX_TRACK && it.trig + it.n_post < nf, "n_post out of range");
    }
    // chunk so that the TOZERO'd scratch stays bounded: a quarter of the free HBM, between 4 and 24 GB
    // (ABUB_LOC_SCRATCH_MB overrides).  Every chunk costs three stream synchronisations and its own small launches
    // of K5c-K7 (a 125-event batch in one chunk instead of three: 6.8 -> 5.9 ms of localizer time), and 180 GB of HBM
    // leave room for a whole batch's frames; the arena only grows to what a chunk really needs.
    if (a->scratch_cap == 0) {                                         // asked once per context: cudaMemGetInfo is not cheap
        a->scratch_cap = (size_t)4 << 30;
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
            size_t want = (free_b + (size_t)a->jobs_cap * npix) / 4;   // what the arena already holds counts as free
            if (want > ((size_t)24 << 30)) want = (size_t)24 << 30;
            if (want > a->scratch_cap) a->scratch_cap = want;
        }
    }
    size_t scratch = a->scratch_cap;
    if (const char* e = getenv("ABUB_LOC_SCRATCH_MB")) { const long mb = atol(e); if (mb > 0) scratch = (size_t)mb << 20; }
    size_t max_jobs = scratch / npix;
    if (max_jobs < ABUB_MAX_TRACK + 1) max_jobs = ABUB_MAX_TRACK + 1;
    std::vector<abub_blob> blobs;
    std::vector<int> npf;
    int written = 0, rc;
    for (int i0 = 0; i0 < n_items;) {
        int i1 = i0, njobs = 0;
        while (i1 < n_items && njobs + items[i1].n_post + 1 <= (int)max_jobs) { njobs += items[i1].n_post + 1; i1++; }
        if ((rc = ensure_jobs(ctx, a, njobs))) return rc;
        int j = 0;
        for (int i = i0; i < i1; i++) {
            const abub_loc_item& it = items[i];
            const abub_cam_state& cs = ctx->cam[it.cam];
            const uint8_t* base = ctx->slots[it.slot].d_frames + (size_t)it.cam * ctx->max_frames * npix;
            for (int k = 0; k <= it.n_post; k++, j++) {
                abub_loc_job& jb = a->h_jobs[j];
                jb.genesis = (k == 0);
                jb.k = k;
                jb.item = i;
                jb.slot = it.slot;
                jb.cam = it.cam;
                jb.zframe = it.cam * ctx->max_frames + it.trig + k;
                jb.zb = it.cam * ctx->max_frames + it.pre_trig;
                jb.k32 = cs.d_k32;
                jb.t8 = cs.d_t8;
                jb.nt16 = cs.d_nt16;
                jb.tz = a->d_tz + (size_t)j * npix;
                jb.segflags = a->d_segflags + (size_t)j * ctx->rows * ((ctx->cols + 239) / 240);
                jb.hist = a->d_hist + (size_t)j * 256;
                if (k == 0) {
                    jb.a = base + (size_t)it.trig * npix;        // L3Localizer.cpp:929
                    jb.b = base + (size_t)it.pre_trig * npix;    // :932-934
                    jb.tozero = it.loc_thres;                    // :252
                } else {
                    jb.a = base + (size_t)(it.trig + k) * npix;  // :770
                    jb.b = cs.d_mean;                            // :779
                    jb.tozero = 3;                               // :786
                }
            }
        }
        blobs.clear();
        if ((rc = run_chunk(ctx, a, njobs, blobs, npf))) return rc;
        // last chunk: its results are on the host; what is left of a deferred trigger search is queued now, so that the
        // GPU works through the host-side tail of this call and the caller's next steps
        if (i1 == n_items && (rc = abub_run_deferred(ctx, 1))) return rc;
        j = 0;
        for (int i = i0; i < i1; i++)
            for (int k = 0; k <= ABUB_MAX_TRACK; k++)
                n_per_frame[i * (ABUB_MAX_TRACK + 1) + k] = (k <= items[i].n_post) ? npf[j++] : 0;
        for (const abub_blob& b : blobs) {
            if (written < cap) {
                out[written] = b;
                out[written].reserved = 0;
            }
            written++;
        }
        i0 = i1;
    }
    if (n_total) *n_total = written;
    if (written > cap) {
        ctx->err = "abub_localize_batch: output capacity too small (" + std::to_string(written) + " blobs)";
        return ABUB_E_CAPACITY;
    }
    return ABUB_OK;
}

// One prepared frame (d_over = overTheSigma already on the device): TOZERO at `tozero`, Otsu, every external contour.
// Used by the bellows veto (k8_bellows.cu); blobs come back with k = 0 in findContours order.
int abub_loc_run_prepared(abub_ctx* ctx, const uint8_t* d_over, int tozero, abub_blob* out, int cap, int* n_out) {
    LocArena* a = arena_of(ctx);
    int rc;
    if ((rc = ensure_jobs(ctx, a, 1))) return rc;
    abub_loc_job& jb = a->h_jobs[0];
    memset(&jb, 0, sizeof(jb));
    jb.a = d_over;
    jb.tz = a->d_tz;
    jb.segflags = a->d_segflags;
    jb.hist = a->d_hist;
    jb.tozero = tozero;
    jb.genesis = 2;
    std::vector<abub_blob> blobs;
    std::vector<int> npf;
    if ((rc = run_chunk(ctx, a, 1, blobs, npf, true))) return rc;
    int written = 0;
    for (const abub_blob& b : blobs) {
        if (written < cap) {
            out[written] = b;
            out[written].reserved = 0;
        }
        written++;
    }
    if (n_out) *n_out = written;
    if (written > cap) {
        ctx->err = "bellows veto: output capacity too small (" + std::to_string(written) + " blobs)";
        return ABUB_E_CAPACITY;
    }
    return ABUB_OK;
}

extern "C" int abub_localize(abub_ctx* ctx, int slot, int cam, int trig, int pre_trig, int loc_thres, int n_post,
                             abub_blob* out, int cap, int* n_per_frame) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, n_per_frame != nullptr, "n_per_frame");
    abub_loc_item it{slot, cam, trig, pre_trig, loc_thres, n_post};
    int npf[ABUB_MAX_TRACK + 1], total = 0;
    int rc = abub_localize_batch(ctx, &it, 1, out, cap, npf, &total);
    for (int k = 0; k <= ABUB_MAX_TRACK; k++) n_per_frame[k] = npf[k];
    return rc;
}

extern "C" int abub_localize_binary(abub_ctx* ctx, int k, uint8_t* out) {
    if (!ctx) return ABUB_E_INVALID;
    LocArena* a = arena_of(ctx);
    ABUB_REQUIRE(ctx, out && a && k >= 0 && k < a->last_njobs, "abub_localize_binary: no such frame in the last call");
    cudaSetDevice(ctx->device);
    ABUB_L(ctx), k5_binary_kernel<<<(unsigned)((ctx->npix + 255) / 256), 256, 0, ctx->stream>>>(a->d_tz, a->d_thr, k, ctx->d_tmp8[3], ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(out, ctx->d_tmp8[3], ctx->npix, cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ABUB_OK;
}
