// K2 -- the significance trigger state machine on the histogram tables produced by K1.
// Reference: AnalyzerUnit::calculateSignificanceFrame + CalcMean/CalcStdDev
// (AnalyzerUnit.cpp:435-504, 514-532) and AnalyzerUnit::FindTriggerFrame (:119-324).
//
// One warp per (slot, camera).  The reference's pix_counts history (vector<vector<int>>)
// is the H table itself: frame f contributes H[f][b] to bin b (bins the reference never
// reaches hold 0, which changes neither sum), and the history length is the number of
// stored frames, n = i when frame i has just been stored (frames 1..i).  Per frame the 254
// per-bin z-scores are evaluated in parallel; the clamped running sum over bins
// (`if (significance < 0) significance = 0`, :479) is order-dependent and is done by lane 0.
// All fp64 arithmetic uses the _rn intrinsics so that nothing is contracted into FMAs
// (the reference is built with -O0, CMakeLists.txt:35).
#include "abub_internal.cuh"

namespace {

constexpr int K2_WARPS = 4;

struct WarpScratch {
    double s1[256];     // sum of counts per bin over stored frames
    double s2[256];     // sum of (int)(count*count) per bin (int32 wrap, AnalyzerUnit.cpp:529)
    double term[256];   // (c-mean)/sd of the frame under evaluation
    uint32_t use[8];    // bit b: term b is added (`binEntry != mean || sigma > 0`, :478)
};

__device__ __forceinline__ double wrapsq(uint32_t c) { return (double)(int32_t)(c * c); }

// A search that stopped for want of frames (ABUB_STATUS_PENDING in the early-exit rounds of abub_trigger_early), kept so that
// the next round continues where it stopped instead of replaying frames 1..i: the history sums exactly as the replay would
// leave them (same additions in the same order), the frame to go on with, and whether that frame is already in the sums
// (the search stopped inside its look-ahead).  next_i == 0: nothing kept.
struct K2State {
    double s1[256];
    double s2[256];
    int next_i, added, scanned, loc_thres;
};

// Evaluates one histogram row against the current history.  Returns the significance on
// all lanes; first_over / max_adc as the reference tracks them (:483-485).
__device__ double significance(const uint32_t* __restrict__ Hrow, WarpScratch* ws, int n, int lane,
                               int* first_over, int* max_adc) {
    int maxbin = 0;
    uint32_t cus[8];
#pragma unroll
    for (int q = 0; q < 8; q++) {
        cus[q] = Hrow[q * 32 + lane];
        if (cus[q]) maxbin = q * 32 + lane;
    }
    maxbin = __reduce_max_sync(0xffffffffu, maxbin);
    // the running sum below stops at the last occupied bin of THIS frame (pix_rem reaches 0 there, :470/:493), so the
    // z-scores of higher bins are never read: skip their fp64 divisions and square roots (typically 7 of the 8 groups)
    const int qn = maxbin >> 5;
#pragma unroll
    for (int q = 0; q < 8; q++) {
        if (q > qn) break;
        const int b = q * 32 + lane;
        const uint32_t cu = cus[q];
        const double c = (double)(float)cu;                       // binEntry is float (:459)
        const double mean = __ddiv_rn(ws->s1[b], (double)n);      // CalcMean :514-521
        const double var = __dsub_rn(__ddiv_rn(ws->s2[b], (double)n), __dmul_rn(mean, mean));
        const double sd = __dsqrt_rn(var);                        // CalcStdDev :523-532
        const bool use = (c != mean) || (sd > 0.0);               // :478
        ws->term[b] = __ddiv_rn(__dsub_rn(c, mean), sd);
        const uint32_t m = __ballot_sync(0xffffffffu, use);
        if (lane == 0) ws->use[q] = m;
    }
    __syncwarp();
    double sig = 0.0;
    int fo = -1;
    if (lane == 0) {
        for (int b = 2; b <= maxbin; b++) {                       // bins 0,1 add nothing (:473)
            if ((ws->use[b >> 5] >> (b & 31)) & 1u) sig = __dadd_rn(sig, ws->term[b]);
            if (sig < 0.0) sig = 0.0;                             // :479
            if (sig > 3.5 && fo < 0) fo = b;                      // :483
        }
    }
    sig = __shfl_sync(0xffffffffu, sig, 0);
    fo = __shfl_sync(0xffffffffu, fo, 0);
    __syncwarp();
    *first_over = fo;
    *max_adc = maxbin;                                            // :485
    return sig;
}

// The two trigger variants the reference keeps next to the significance (commented alternatives at AnalyzerUnit.cpp:226-227,
// :290-291): AnalyzerUnit::calculateEntropyFrame (:386-423), the Shannon entropy of the 128-bin histogram of the
// ProcessFrame output -- bins of width 2, i.e. pairs of the K1 table -- with p = count * (float)(1/N) and
// E -= p*log2(p) accumulated in float in bin order (lane 0).  Returns the float entropy on all lanes.
__device__ float entropy128(const uint32_t* __restrict__ Hrow, WarpScratch* ws, float scale, int lane) {
    float* t = reinterpret_cast<float*>(ws->term);
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const int k = q * 32 + lane;
        const uint2 c2 = *reinterpret_cast<const uint2*>(Hrow + 2 * k);
        const float p = __fmul_rn((float)(c2.x + c2.y), scale);
        t[k] = p != 0.0f ? __fmul_rn(p, (float)log2((double)p)) : 0.0f;
    }
    __syncwarp();
    float e = 0.0f;
    if (lane == 0)
        for (int k = 0; k < 128; k++) e = __fsub_rn(e, t[k]);
    e = __shfl_sync(0xffffffffu, e, 0);
    __syncwarp();
    return e;
}

// AnalyzerUnit::calculateEntropySignificance (:425-433) on running sums: CalcMean / CalcStdDev add the stored
// entropies in push order, which is what a running sum does.
struct EntropyHistory {
    double s1 = 0.0, s2 = 0.0;
    int n = 0;
    __device__ void push(double e) { s1 = __dadd_rn(s1, e); s2 = __dadd_rn(s2, __dmul_rn(e, e)); n++; }
    __device__ double z(double e) const {
        const double mean = __ddiv_rn(s1, (double)n);
        const double sd = __dsqrt_rn(__dsub_rn(__ddiv_rn(s2, (double)n), __dmul_rn(mean, mean)));
        return __ddiv_rn(__dsub_rn(e, mean), sd);
    }
};

__global__ void __launch_bounds__(32 * K2_WARPS)
k2_trigger_scan_kernel(const uint32_t* __restrict__ hist, const int* __restrict__ nframes,
                       const int* __restrict__ start, const int* __restrict__ tss_cam,
                       int* __restrict__ result, double* __restrict__ sig_out, int slot_first, int nslots,
                       int ncams, int max_frames, const int* __restrict__ avail, int metric, float thr0, float inv_npix,
                       K2State* __restrict__ state, int resume) {
    __shared__ WarpScratch scratch[K2_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item = blockIdx.x * K2_WARPS + warp;
    if (item >= nslots * ncams) return;
    const int slot = slot_first + item / ncams, cam = item % ncams;
    const int sc = slot * ncams + cam;
    WarpScratch* ws = &scratch[warp];
    const uint32_t* H = hist + (size_t)sc * max_frames * 256;
    double* so = sig_out + (size_t)sc * max_frames;
    int* res = result + sc * 4;
    const int nf = nframes[sc];
    // streamed form (abub_trigger_stream): only the first `av` frames are resident; a decision that needs a
    // later frame is reported as ABUB_STATUS_PENDING and the search is repeated when more frames have arrived
    const int av = avail ? min(avail[sc], nf) : nf;
    if (avail && avail[sc] <= 0) return;                          // finished earlier (or nothing resident): leave its results
    const int tss = tss_cam[cam];
    // the significance searches of abub_trigger_early keep their state between rounds (`state`); `resume`: this is a later round
    K2State* const stp = (state && metric == ABUB_METRIC_SIGNIFICANCE) ? state + sc : nullptr;
    const bool resumed = stp && resume && stp->next_i > 0;
    if (!resumed)
        for (int f = lane; f < max_frames; f += 32) so[f] = 0.0;

    if (nf < 5) {                                                 // :122-126
        if (lane == 0) { res[0] = 0; res[1] = -9; res[2] = -1; res[3] = 0; }
        if (stp && lane == 0) stp->next_i = 0;
        return;
    }
    int s = start[sc];
    if (s < 1) s = 1;                                             // :163
    float thr = thr0;                                             // :140-142 (3.5 for the significance)
    if (tss < 6) thr = (float)((double)thr * (5 / 3.5));          // :185-188
    // history of frames 1..s-1 (empty when s == 1, :165-169)
    EntropyHistory eh;
    if (metric == ABUB_METRIC_ENTROPY_SIGNIFICANCE)
        for (int f = 1; f < s && f < nf; f++) eh.push((double)entropy128(H + (size_t)f * 256, ws, inv_npix, lane));
    int status = -3, trig = 0, loc_thres = -1, scanned = 0;      // :180
    int i_first = s, i_added = -1;                                // i_added: this frame is already in the sums
    if (resumed) {
#pragma unroll
        for (int q = 0; q < 8; q++) {
            const int b = q * 32 + lane;
            ws->s1[b] = stp->s1[b];
            ws->s2[b] = stp->s2[b];
        }
        i_first = stp->next_i;
        if (stp->added) i_added = i_first;
        scanned = stp->scanned;
        loc_thres = stp->loc_thres;
    } else {
#pragma unroll
        for (int q = 0; q < 8; q++) {
            const int b = q * 32 + lane;
            double a1 = 0.0, a2 = 0.0;
            for (int f = 1; f < s && f < nf; f++) {
                const uint32_t c = H[(size_t)f * 256 + b];
                a1 += (double)c;
                a2 += wrapsq(c);
            }
            ws->s1[b] = a1;
            ws->s2[b] = a2;
        }
    }
    __syncwarp();

    int keep_i = 0, keep_added = 0;                               // where a PENDING search goes on
    for (int i = i_first; i < nf; i++) {                          // :191
        if (i >= av) { status = ABUB_STATUS_PENDING; keep_i = i; keep_added = 0; break; }
        const uint32_t* Hi = H + (size_t)i * 256;
        double sg;
        if (metric == ABUB_METRIC_SIGNIFICANCE) {
            if (i != i_added) {
#pragma unroll
                for (int q = 0; q < 8; q++) {                     // store = true (:461)
                    const int b = q * 32 + lane;
                    const uint32_t c = Hi[b];
                    ws->s1[b] += (double)c;
                    ws->s2[b] += wrapsq(c);
                }
            }
            __syncwarp();
            int fo, madc;
            sg = significance(Hi, ws, i, lane, &fo, &madc);       // :228
            int lt = max(fo - 1, madc - 1);                       // :487-491
            if (lt < 2) lt = 2;
            if (lt > 3 || tss < 6) lt = 3;
            loc_thres = lt;
        } else {
            const double e = (double)entropy128(Hi, ws, inv_npix, lane);   // :226 / :227
            if (metric == ABUB_METRIC_ENTROPY_SIGNIFICANCE) { eh.push(e); sg = eh.z(e); }
            else sg = e;
            loc_thres = 3;                                        // never updated on these paths (:29)
        }
        if (i != i_added) scanned++;                              // a frame met again after a stop in its look-ahead counts once
        if (lane == 0) so[i] = sg;
        float single = (float)sg;
        if (single > thr && i >= 2) {                             // :254
            if (i != nf - 1) {                                    // :261
                double max_so_far = (double)single;              // :280
                for (int ii = 1; ii <= 2 && ii + i < nf; ii++) {  // :282
                    if (i + ii >= av) { status = ABUB_STATUS_PENDING; keep_i = i; keep_added = 1; break; }
                    int fo2, madc2;
                    double s2v;                                   // :290-292  (store = false)
                    if (metric == ABUB_METRIC_SIGNIFICANCE) {
                        s2v = significance(H + (size_t)(i + ii) * 256, ws, i, lane, &fo2, &madc2);
                    } else {
                        const double e2 = (double)entropy128(H + (size_t)(i + ii) * 256, ws, inv_npix, lane);
                        s2v = metric == ABUB_METRIC_ENTROPY_SIGNIFICANCE ? eh.z(e2) : e2;
                    }
                    single = (float)s2v;
                    const double lhs = __dadd_rn(__ddiv_rn((double)single, __dmul_rn(__ddiv_rn((double)thr, 3.5), 5.0)),
                                                 __ddiv_rn((double)single, max_so_far));
                    if (lhs <= 3.0) break;                        // :296
                    else if (ii == 2) { status = 0; trig = i; }   // :297-300
                    if ((double)single > max_so_far) max_so_far = (double)single;   // :302
                }
                if (status == 0 || status == ABUB_STATUS_PENDING) break;
            }
        }
    }
    if (lane == 0) { res[0] = trig; res[1] = status; res[2] = loc_thres; res[3] = scanned; }
    if (stp) {
        if (status == ABUB_STATUS_PENDING) {
            __syncwarp();
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const int b = q * 32 + lane;
                stp->s1[b] = ws->s1[b];
                stp->s2[b] = ws->s2[b];
            }
            if (lane == 0) { stp->next_i = keep_i; stp->added = keep_added; stp->scanned = scanned; stp->loc_thres = loc_thres; }
        } else if (lane == 0) {
            stp->next_i = 0;
        }
    }
}

// Resident early exit (abub_trigger_early): frames a camera may look at in the next round -- none once its search is
// decided, the next boundary while K2 reports that it needs more frames.  first round: result == nullptr.
__global__ void k2_next_avail_kernel(const int* __restrict__ result, const int* __restrict__ nframes, int* __restrict__ avail,
                                     int* __restrict__ start, int first, int n, int next) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int sc = first + i;
    if (result == nullptr) {
        start[sc] = 1;
        avail[sc] = max(1, min(next, nframes[sc]));       // >= 1: K2 must still report -9 for a camera without frames
    } else {
        avail[sc] = (result[sc * 4 + 1] == ABUB_STATUS_PENDING) ? min(next, nframes[sc]) : 0;
    }
}

}  // namespace

int abub_k2_next_avail(abub_ctx* ctx, int slot_first, int nslots, int next, bool first_round) {
    const int n = nslots * ctx->ncams;
    ABUB_L(ctx), k2_next_avail_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(first_round ? nullptr : ctx->d_trig, ctx->d_nframes, ctx->d_avail,
                                                                     ctx->d_start, slot_first * ctx->ncams, n, next);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

int abub_k2_launch(abub_ctx* ctx, int slot_first, int nslots) { return abub_k2_launch_avail(ctx, slot_first, nslots, nullptr); }

// keep_state: the rounds of abub_trigger_early -- a search that stops for want of frames leaves its history sums in
// ctx->d_k2_state; resume: this is a later round of the same searches (the first round passes false and starts afresh)
int abub_k2_launch_avail(abub_ctx* ctx, int slot_first, int nslots, const int* d_avail, bool keep_state, bool resume) {
    const int items = nslots * ctx->ncams;
    if (keep_state && !ctx->d_k2_state) {
        const size_t n = (size_t)ctx->max_slots * ctx->ncams;
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_k2_state, n * sizeof(K2State)));
        ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(ctx->d_k2_state, 0, n * sizeof(K2State), ctx->stream));
    }
    abub_prof_scope prof(ctx, ABUB_PROF_K2);
    ABUB_L(ctx), k2_trigger_scan_kernel<<<(items + K2_WARPS - 1) / K2_WARPS, 32 * K2_WARPS, 0, ctx->stream>>>(
        ctx->d_hist, ctx->d_nframes, ctx->d_start, ctx->d_tss, ctx->d_trig, ctx->d_sig, slot_first, nslots,
        ctx->ncams, ctx->max_frames, d_avail, ctx->trig_metric, ctx->trig_thr, (float)(1.0 / (double)ctx->npix),
        keep_state ? reinterpret_cast<K2State*>(ctx->d_k2_state) : nullptr, resume ? 1 : 0);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

extern "C" int abub_set_trigger_metric(abub_ctx* ctx, int metric, float threshold) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, metric >= ABUB_METRIC_SIGNIFICANCE && metric <= ABUB_METRIC_ENTROPY_SIGNIFICANCE, "trigger metric");
    static const float kDefault[3] = {3.5f, 8e-5f, 1.9f};        // AnalyzerUnit.cpp:142, :140, :141
    ctx->trig_metric = metric;
    ctx->trig_thr = threshold > 0.0f ? threshold : kDefault[metric];
    return ABUB_OK;
}
