// K5b (TMA-fed) -- L3Localizer::CalculatePostTriggerFrameParams front end (reference: L3Localizer.cpp:779-787):
//   overTheSigma = absdiff(frame[trig+k], TrainedAvgImage) - 6 * TrainedSigmaImage   (saturating)
//   blur 3x3 (BORDER_REFLECT_101, (sum + 4) / 9), THRESH_TOZERO at 3, 256-bin histogram for the Otsu step (K5c)
// for every post-trigger job of a chunk.  Same arithmetic as k5_post_stream_kernel (k5_localize.cu, the generic-pitch
// path); this file is the hot path for pitches TMA can address (cols % 16 == 0), built like the trigger kernel
// (k1_trigger_tma.cu): one warp (= one CTA) owns a strip of 240 output columns x 128 rows, lane l holds 8 adjacent pixels
// (lanes 0 and 31 are the halo), and rows arrive as TMA boxes of 272 pixels x 4 rows -- the frame (3-D u16 map of the
// slot), TrainedAvgImage (2-D u16 map) and -min(6 sigma, 255) as s16 (2-D u32 map) -- through a ring of stages completed
// on mbarriers.  No per-lane global addressing, no address arithmetic in the row loop.
//
// Rows: every stage is consumed whole (4 rows), so the row loop is unrolled without tests.  A strip first consumes a
// priming box that ENDS at the row above the strip (for the top strip that is rows -2 .. 1: TMA zero-fills the two rows
// outside the image and the sequence 0, 1 | 0, 1, 2 ... hands row 1 to output row 0 as BORDER_REFLECT_101 wants), then the
// boxes from its first row on; the bottom strip ends on the image's last row exactly (strip heights are multiples of 4,
// the top strip takes the remainder) and is followed by one box at rows-2 (reflection of the row below the image).
// An output row is formed when the row below it is consumed and stored only inside the strip.
//
// Pixels that survive the 3x3 mean and THRESH_TOZERO are rare: their histogram counts go straight to the job's table with
// global atomics (no shared table, which would cost a kilobyte of the shared memory that bounds the number of resident
// warps).  Skipping the 3x3 sums of rows whose three-row window is empty (one vote per row; ABUB_K5T_SKIP=1) was measured
// and does not pay: 70 % of the windows of a 256-pixel row hold a pixel over the cut.  The TOZERO'd frames stay sparse (segment flags, see
// k5_localize.cu): a row segment is written only when it holds something or may still hold something older.
#include <cuda.h>

#include <type_traits>

#include "abub_internal.cuh"

namespace {

#ifndef ABUB_K5T_S
#define ABUB_K5T_S 2
#endif
#ifndef ABUB_K5T_TH
#define ABUB_K5T_TH 128
#endif
#ifndef ABUB_K5T_MINCTA
#define ABUB_K5T_MINCTA 24
#endif
#ifndef ABUB_K5T_SKIP
#define ABUB_K5T_SKIP 0
#endif
#ifndef ABUB_K5T_FMA
#define ABUB_K5T_FMA 1      // 1: the adds of the 3x3 window go through IMAD (a * one + b with a run-time 1): the ALU pipe, which
#endif                      // carries PRMT / VIADDMNMX / VIMNMX / VABSDIFF4 / IADD3 / LOP3 and bounds this kernel, is relieved
constexpr int P_R = 4;                    // rows per TMA box
constexpr int P_S = ABUB_K5T_S;           // ring stages per warp
constexpr int P_OW = 240;                 // output columns per warp
constexpr int P_BW = 272;                 // box width in pixels: [x0-16, x0+256)
constexpr int P_TH = ABUB_K5T_TH;              // output rows per strip (a multiple of P_R)
constexpr int P_A_OFF = 0;                                        // frame rows        (P_R x 272 B)
constexpr int P_M_OFF = ((P_R * P_BW + 127) / 128) * 128;         // TrainedAvgImage rows
constexpr int P_T_OFF = 2 * P_M_OFF;                              // -t rows           (P_R x 544 B)
constexpr int P_TX_BYTES = P_R * P_BW * 4;
constexpr int P_STAGE_BYTES = ((P_T_OFF + P_R * P_BW * 2 + 127) / 128) * 128;
constexpr int P_MINCTA = ABUB_K5T_MINCTA;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int x, int y, int z, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int x, int y, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(bar) : "memory");
}

// grid: x = strip * nxg + xg, y = index into the list of post jobs ; block = 32
__global__ void __launch_bounds__(32, P_MINCTA)
k5_post_tma_kernel(const abub_loc_job* __restrict__ jobs, const int* __restrict__ post_list, const CUtensorMap* __restrict__ slot_maps,
                   const CUtensorMap* __restrict__ mean_maps, const CUtensorMap* __restrict__ nt_maps, int rows, int cols, int nxg,
                   int nstrips, int h0, uint32_t one) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const int lane = threadIdx.x;
    const abub_loc_job job = jobs[post_list[blockIdx.y]];
    const uint32_t ring = smem_u32(smem_raw);
    const uint32_t bars = ring + P_S * P_STAGE_BYTES;
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < P_S; s++) mbar_init(bars + s * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    const int strip = blockIdx.x / nxg, xg = blockIdx.x % nxg;
    // the top strip takes the remainder h0 = rows - P_TH * (nstrips - 1), so that every other strip is P_TH rows high and
    // the bottom one ends on the last image row
    const int y0 = strip == 0 ? 0 : h0 + (strip - 1) * P_TH;
    const int nout = strip == 0 ? h0 : P_TH;
    const bool bottom = (y0 + nout == rows);
    const int x0 = xg * P_OW;
    const int xe = (x0 - 16) / 2;
    const int gx = x0 - 8 + lane * 8;
    const bool out_lane = lane >= 1 && lane <= 30 && gx < cols;
    const bool left_edge = (gx == 0), right_edge = (gx + 8 == cols);
    const CUtensorMap* const smap = slot_maps + job.slot;
    const CUtensorMap* const mmap = mean_maps + job.cam;
    const CUtensorMap* const tmap = nt_maps + job.cam;
    const int za = job.zframe;
    const uint32_t lim = 9u * (uint32_t)job.tozero + 5u;                // (sum+4)/9 <= tozero  <=>  sum < lim
    const int tozero = job.tozero;
    const uint32_t nlim2 = ((0x10000u - (lim - 1u)) & 0xFFFFu) * 0x00010001u;           // -(lim - 1) in both s16 lanes
    // stages: 0 = priming box (rows y0-4 .. y0-1; the top strip: -2 .. 1), 1 .. nreg = boxes from y0 on (covering at least one
    // row below the strip), then for the bottom strip one box at rows-2.  ybox(s) = first image row of stage s.
    const int nreg = bottom ? nout / P_R : (nout + 1 + P_R - 1) / P_R;
    const int nst = 1 + nreg + (bottom ? 1 : 0);
    auto ybox = [&](int s) {
        if (s == 0) return (y0 == 0) ? -2 : y0 - P_R;
        if (s <= nreg) return y0 + (s - 1) * P_R;
        return rows - 2;
    };
    auto issue = [&](int s, int stage) {
        const uint32_t base = ring + stage * P_STAGE_BYTES, bar = bars + stage * 8;
        const int ys = ybox(s);
        mbar_expect_tx(bar, P_TX_BYTES);
        tma_load_3d(base + P_A_OFF, smap, xe, ys, za, bar);
        tma_load_2d(base + P_M_OFF, mmap, xe, ys, bar);
        tma_load_2d(base + P_T_OFF, tmap, xe, ys, bar);
    };
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < P_S; s++)
            if (s < nst) issue(s, s);
    }

    // state of the vertical 3-sum without register rotation: pr = the row before, pp = pr + the row before that
    uint32_t pr[4] = {0, 0, 0, 0}, pp[4] = {0, 0, 0, 0};
    uint32_t zhist = 0;                                                 // bit i: the row i rows back holds something (warp-uniform)
    const int nxw = (cols + P_OW - 1) / P_OW;
    uint8_t* const flags = job.segflags + xg;                           // [row][nxw]
    int stage = 0;
    uint32_t parity = 0;
    auto yo0_of = [&](int s) {
        // nominal output row of the first row of stage s: the row above it (stage 0 of the top strip: the box is rows -2 .. 1
        // and its last row plays row -1; the bottom extra box continues at rows-1)
        return (s == 0) ? ((y0 == 0) ? -5 : y0 - P_R - 1) : (s <= nreg ? y0 + (s - 1) * P_R - 1 : rows - 1);
    };
    // segment flags of the four output rows of a stage (lanes 0..3), fetched one stage ahead
    auto flag_load = [&](int s) -> uint32_t {
        const int yf = yo0_of(s) + (lane & 3);
        return (lane < 4 && s < nst && yf >= y0 && yf < y0 + nout) ? (uint32_t)flags[(size_t)yf * nxw] : 0u;
    };
    uint32_t fl_next = flag_load(0);
#pragma unroll 1
    for (int s = 0; s < nst; s++) {
        const int yo0 = yo0_of(s);
        const int yf = yo0 + (lane & 3);
        const bool frow = yf >= y0 && yf < y0 + nout;
        const uint32_t fl = fl_next;
        fl_next = flag_load(s + 1);
        mbar_wait(bars + stage * 8, parity);
        const uint8_t* const st = smem_raw + stage * P_STAGE_BYTES;
        uint32_t newfl = fl;
        // CHK = the stage may hold rows without an output (the priming box, the rows past the strip): interior stages
        // run the copy of the row body without the per-row test
        auto do_rows = [&](auto chk) {
        constexpr bool CHK = decltype(chk)::value;
#pragma unroll
        for (int r = 0; r < P_R; r++) {
            const uint2 a = *reinterpret_cast<const uint2*>(st + P_A_OFF + r * P_BW + 8 + lane * 8);
            const uint2 m = *reinterpret_cast<const uint2*>(st + P_M_OFF + r * P_BW + 8 + lane * 8);
            const uint4 t = *reinterpret_cast<const uint4*>(st + P_T_OFF + r * 2 * P_BW + 16 + lane * 16);
            const uint32_t d0 = __vabsdiffu4(a.x, m.x), d1 = __vabsdiffu4(a.y, m.y);          // L3Localizer.cpp:779
            uint32_t w[4];                                                                     // :782  - 6 sigma, saturating
            w[0] = __viaddmax_s16x2_relu(__byte_perm(d0, 0, 0x4140), t.x, t.x);
            w[1] = __viaddmax_s16x2_relu(__byte_perm(d0, 0, 0x4342), t.y, t.y);
            w[2] = __viaddmax_s16x2_relu(__byte_perm(d1, 0, 0x4140), t.z, t.z);
            w[3] = __viaddmax_s16x2_relu(__byte_perm(d1, 0, 0x4342), t.w, t.w);
            const bool row_nz = __any_sync(0xffffffffu, (w[0] | w[1] | w[2] | w[3]) != 0u);
            zhist = (zhist << 1) | (row_nz ? 1u : 0u);
            const int yo = yo0 + r;                                     // output row completed by this row
            const uint32_t prev = __shfl_sync(0xffffffffu, fl, r);      // its segment flag
            if (!CHK || (yo >= y0 && yo < y0 + nout)) {                 // warp-uniform
                uint2 o = make_uint2(0u, 0u);
                bool nz = false;
                if (!ABUB_K5T_SKIP || (zhist & 7u)) {                   // some pixel over the cut in these three rows
                    uint32_t W[6];
#pragma unroll
                    for (int k = 0; k < 4; k++) W[k + 1] = ABUB_K5T_FMA ? w[k] * one + pp[k] : w[k] + pp[k];
                    W[0] = __shfl_up_sync(0xffffffffu, W[4], 1);
                    W[5] = __shfl_down_sync(0xffffffffu, W[1], 1);
                    if (left_edge) W[0] = __byte_perm(W[1], 0, 0x3232);     // column -1 -> column 1
                    if (right_edge) W[5] = __byte_perm(W[4], 0, 0x1010);    // column cols -> column cols-2
                    uint32_t H[4], mx = 0;
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const uint32_t sl = __byte_perm(W[k], W[k + 1], 0x5432);      // (x-1, x) pairs
                        const uint32_t sr = __byte_perm(W[k + 1], W[k + 2], 0x5432);  // (x+1, x+2) pairs
                        H[k] = ABUB_K5T_FMA ? (sl * one + W[k + 1]) + sr : sl + W[k + 1] + sr;
                        mx = __vmaxu2(mx, H[k]);
                    }
                    // some 16-bit sum of the lane reaches lim: relu(mx - (lim - 1)) per s16 lane, one VIADDMNMX (sums <= 2295)
                    if (out_lane && __viaddmax_s16x2_relu(mx, nlim2, nlim2) != 0u) {
                        uint32_t ob[8];
#pragma unroll
                        for (int k = 0; k < 4; k++) {
                            ob[2 * k] = (((H[k] & 0xFFFFu) + 4u) * 7282u) >> 16;       // (sum+4)/9, exact for sum <= 2295
                            ob[2 * k + 1] = (((H[k] >> 16) + 4u) * 7282u) >> 16;
                        }
#pragma unroll
                        for (int c = 0; c < 8; c++) {
                            if ((int)ob[c] <= tozero) ob[c] = 0;                       // THRESH_TOZERO (:786)
                            if (ob[c]) atomicAdd(&job.hist[ob[c]], 1u);        // rare: straight to the job's table in L2
                        }
                        o.x = ob[0] | (ob[1] << 8) | (ob[2] << 16) | (ob[3] << 24);
                        o.y = ob[4] | (ob[5] << 8) | (ob[6] << 16) | (ob[7] << 24);
                    }
                    nz = __any_sync(0xffffffffu, (o.x | o.y) != 0u);
                }
                // the segment is stored when it holds something, or when what the memory holds is not known to be zero
                if (nz || (prev & 2u)) {
                    if (out_lane) *reinterpret_cast<uint2*>(job.tz + (size_t)yo * cols + gx) = o;
                }
                if (lane == r) newfl = nz ? 3u : 0u;
            }
#pragma unroll
            for (int k = 0; k < 4; k++) { pp[k] = pr[k] + w[k]; pr[k] = w[k]; }
        }
        };
        if (yo0 >= y0 && yo0 + P_R <= y0 + nout) do_rows(std::false_type{});
        else do_rows(std::true_type{});
        if (lane < 4 && frow && newfl != fl) flags[(size_t)yf * nxw] = (uint8_t)newfl;
        __syncwarp();                                                   // every lane is done with the stage
        if (lane == 0 && s + P_S < nst) issue(s + P_S, stage);
        if (++stage == P_S) { stage = 0; parity ^= 1u; }
    }
}

}  // namespace

bool abub_k5b_tma_supported(const abub_ctx* ctx) {
    return ctx->d_tmaps != nullptr && (ctx->cols % 16) == 0 && ctx->cols >= 16 && ctx->rows > P_TH + 2;
}

int abub_k5b_launch_tma(abub_ctx* ctx, const abub_loc_job* d_jobs, int njobs) {
    // the post-trigger jobs of the chunk (genesis == 0 in the host copy of the job table): their indices go up as a table
    const abub_loc_job* hj = abub_loc_host_jobs(ctx);
    int np = 0;
    for (int j = 0; j < njobs; j++) np += (hj[j].genesis == 0);
    if (np == 0) return ABUB_OK;
    if (np > ctx->post_cap) {
        if (ctx->post_ev) ABUB_CHECK_CUDA(ctx, cudaEventSynchronize(ctx->post_ev));
        cudaFree(ctx->d_post_list);
        if (ctx->h_post_list) cudaFreeHost(ctx->h_post_list);
        ctx->d_post_list = nullptr;
        ctx->h_post_list = nullptr;
        ctx->post_cap = 0;
        const int cap = np + np / 2 + 64;
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_post_list, (size_t)cap * sizeof(int)));
        ABUB_CHECK_CUDA(ctx, cudaHostAlloc(&ctx->h_post_list, (size_t)cap * sizeof(int), cudaHostAllocDefault));
        ctx->post_cap = cap;
    }
    if (!ctx->post_ev) ABUB_CHECK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->post_ev, cudaEventDisableTiming));
    else ABUB_CHECK_CUDA(ctx, cudaEventSynchronize(ctx->post_ev));     // the previous chunk's upload has left the staging copy
    for (int j = 0, k = 0; j < njobs; j++)
        if (hj[j].genesis == 0) ctx->h_post_list[k++] = j;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->d_post_list, ctx->h_post_list, (size_t)np * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaEventRecord(ctx->post_ev, ctx->stream));
    const int nxg = (ctx->cols + P_OW - 1) / P_OW;
    const int nstrips = (ctx->rows + P_TH - 1) / P_TH;
    const int h0 = ctx->rows - P_TH * (nstrips - 1);
    const size_t smem = (size_t)P_S * P_STAGE_BYTES + P_S * 8;
    ABUB_CHECK_CUDA(ctx, cudaFuncSetAttribute(k5_post_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const CUtensorMap* maps = reinterpret_cast<const CUtensorMap*>(ctx->d_tmaps);
    for (int p0 = 0; p0 < np; p0 += 65535) {
        const int n = np - p0 < 65535 ? np - p0 : 65535;
        ABUB_L(ctx), k5_post_tma_kernel<<<dim3(nstrips * nxg, n), 32, smem, ctx->stream>>>(
            d_jobs, ctx->d_post_list + p0, maps, maps + ctx->max_slots + ctx->ncams, maps + ctx->max_slots + 2 * ctx->ncams, ctx->rows, ctx->cols,
            nxg, nstrips, h0, 1u);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    return ABUB_OK;
}
