// K5a (TMA-fed) -- the trigger ("genesis") frame of L3Localizer::CalculateInitialBubbleParams (reference:
// L3Localizer.cpp:232 ProcessFrame(comparison frame, trigger frame), :252 THRESH_TOZERO at loc_thres) for every genesis
// job of a localizer chunk, plus the 256-bin histogram the Otsu step (K5c) needs.
//
// ProcessFrame is the arithmetic of the trigger kernel (AnalyzerUnit.cpp:346-377; k1_trigger_tma.cu): this kernel is that
// kernel's row pipeline -- one warp (= one CTA) per strip of 240 output columns, lane l holds 8 adjacent pixels, rows
// arrive as TMA boxes of 272 pixels x 4 rows (trigger frame and comparison frame from the slot's 3-D map, the constant
// plane K16 = -(255 + min(6 sigma, 255)) from the camera's 2-D map) through a two-stage ring completed on mbarriers;
// stage A on s16x2 with VIADDMNMX, the vertical [1 4 6 4 1] as a cascade on carried state, the horizontal pass with SHFL
// halos, |P - N| per byte -- with a different tail:
//   * THRESH_TOZERO at the job's loc_thres on the packed bytes (SIMD-in-register compare, no branches);
//   * the frame is stored SPARSELY with the segment-flag protocol of the post-trigger kernel (k5_post_tma.cu,
//     k5_localize.cu): a 240-pixel row segment is written only when it holds a pixel over the cut, or when the memory
//     may still hold older non-zero bytes; its flag then says which.  The dense store of the legacy path wrote 1.76 MB
//     per genesis frame and made the candidate scan (K6) read it all back;
//   * pixels that survive the cut are rare (bubble pixels): their histogram counts go straight to the job's table with
//     global atomics.
// Strips are 128 output rows high (9 strips of a 1050-row frame): a chunk holds only ~120 genesis jobs, and the
// trigger kernel's 350-row strips would leave the GPU under one wave of warps.
//
// Strips tile the output rows [2, rows-2); the four border rows, whose vertical reflection a streaming strip cannot
// express with contiguous boxes, go through the generic kernel (abub_k1_launch_genesis_border, k1_processframe.cu),
// which writes them densely and flags their segments as such.
#include <cuda.h>

#include "abub_internal.cuh"

namespace {

constexpr int G_R = 4;                    // rows per TMA box (fixed by the tensor maps, k1_trigger_tma.cu)
constexpr int G_S = 2;                    // ring stages
constexpr int G_OW = 240;                 // output columns per warp
constexpr int G_BW = 272;                 // box width in pixels: [x0-16, x0+256)
constexpr int G_TH = 128;                 // output rows per strip (a multiple of G_R)
constexpr int G_A_OFF = 0;
constexpr int G_B_OFF = ((G_R * G_BW + 127) / 128) * 128;
constexpr int G_K_OFF = 2 * G_B_OFF;
constexpr int G_TX_BYTES = G_R * G_BW * 4;
constexpr int G_STAGE_BYTES = ((G_K_OFF + G_R * G_BW * 2 + 127) / 128) * 128;
constexpr int G_MINCTA = 20;

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
__device__ __forceinline__ uint32_t relu_add_s16x2(uint32_t x, uint32_t k) { return __viaddmax_s16x2_relu(x, k, k); }

// horizontal [1 4 6 4 1] of one plane: W[0..5] = words -1..4 (two pixels each), out H[0..3] incl. +128
__device__ __forceinline__ void hpass(const uint32_t W[6], uint32_t H[4]) {
    uint32_t S[5];
#pragma unroll
    for (int k = 0; k < 5; k++) S[k] = __byte_perm(W[k], W[k + 1], 0x5432);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t e = W[k] + W[k + 2] + 0x00800080u;
        const uint32_t t = S[k] + S[k + 1];
        H[k] = e + 6u * W[k + 1] + 4u * t;
    }
}

// grid: x = strip * nxg + xg, y = index into the list of genesis jobs ; block = 32
__global__ void __launch_bounds__(32, G_MINCTA)
k5_genesis_tma_kernel(const abub_loc_job* __restrict__ jobs, const int* __restrict__ gen_list, const CUtensorMap* __restrict__ slot_maps,
                      const CUtensorMap* __restrict__ k16_maps, int rows, int cols, int nxg, int nstrips, int h0) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const int lane = threadIdx.x;
    const abub_loc_job job = jobs[gen_list[blockIdx.y]];
    const uint32_t ring = smem_u32(smem_raw);
    const uint32_t bars = ring + G_S * G_STAGE_BYTES;
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < G_S; s++) mbar_init(bars + s * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    // strips tile the output rows [2, rows-2); the first strip takes the remainder h0 so that the others are G_TH high
    const int strip = blockIdx.x / nxg, xg = blockIdx.x % nxg;
    const int y0 = strip == 0 ? 2 : 2 + h0 + (strip - 1) * G_TH;
    const int nout = strip == 0 ? h0 : G_TH;
    const int x0 = xg * G_OW;
    const int nloc = nout + 4;                          // image rows y0-2 .. y0+nout+1
    const int nst = (nloc + G_R - 1) / G_R;
    const int xe = (x0 - 16) / 2;
    const int gx = x0 - 8 + lane * 8;
    const bool out_lane = lane >= 1 && lane <= 30 && gx < cols;
    const bool left_edge = (gx == 0), right_edge = (gx + 8 == cols);
    const CUtensorMap* const smap = slot_maps + job.slot;
    const CUtensorMap* const kmap = k16_maps + job.cam;
    // ProcessFrame(comparison frame, trigger frame) (L3Localizer.cpp:232): a = job.a = the trigger frame, b = job.b
    const int za = job.zframe, zb = job.zb;
    // v > loc_thres per byte: a negative cut keeps every non-zero value (as v > 0 does), a cut >= 255 keeps nothing
    const uint32_t cut = (uint32_t)min(max(job.tozero, 0), 255) * 0x01010101u;
    auto issue = [&](int s, int stage) {
        const uint32_t base = ring + stage * G_STAGE_BYTES, bar = bars + stage * 8;
        const int ys = y0 - 2 + s * G_R;
        mbar_expect_tx(bar, G_TX_BYTES);
        tma_load_3d(base + G_A_OFF, smap, xe, ys, za, bar);
        tma_load_3d(base + G_B_OFF, smap, xe, ys, zb, bar);
        tma_load_2d(base + G_K_OFF, kmap, xe, ys, bar);
    };
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < G_S; s++)
            if (s < nst) issue(s, s);
    }
    uint32_t X[8], P[8], Q[8], Rr[8];
#pragma unroll
    for (int i = 0; i < 8; i++) X[i] = P[i] = Q[i] = Rr[i] = 0;
    const int nxw = (cols + G_OW - 1) / G_OW;
    uint8_t* const flags = job.segflags + xg;                           // [row][nxw]
    // segment flags of the four output rows a stage completes (lanes 0..3), fetched one stage ahead.  Local row rl of the
    // strip (image row y0-2+rl) completes output row y0+rl-4.
    auto flag_load = [&](int s) -> uint32_t {
        const int yf = y0 + s * G_R + (lane & 3) - 4;
        return (lane < 4 && s < nst && yf >= y0 && yf < y0 + nout) ? (uint32_t)flags[(size_t)yf * nxw] : 0u;
    };
    uint32_t fl_next = flag_load(0);
    int stage = 0;
    uint32_t parity = 0;
#pragma unroll 1
    for (int s = 0; s < nst; s++) {
        const int yf = y0 + s * G_R + (lane & 3) - 4;
        const bool frow = yf >= y0 && yf < y0 + nout;
        const uint32_t fl = fl_next;
        fl_next = flag_load(s + 1);
        mbar_wait(bars + stage * 8, parity);
        const uint8_t* const st = smem_raw + stage * G_STAGE_BYTES;
        uint32_t newfl = fl;
#pragma unroll
        for (int r = 0; r < G_R; r++) {
            const int rl = s * G_R + r;
            const uint2 a = *reinterpret_cast<const uint2*>(st + G_A_OFF + r * G_BW + 8 + lane * 8);
            const uint2 b = *reinterpret_cast<const uint2*>(st + G_B_OFF + r * G_BW + 8 + lane * 8);
            const uint4 k = *reinterpret_cast<const uint4*>(st + G_K_OFF + r * 2 * G_BW + 16 + lane * 16);
            uint32_t w[8];
            {
                const uint32_t a0 = __byte_perm(a.x, 0, 0x4140), b0 = __byte_perm(b.x, 0, 0x4140);
                const uint32_t a1 = __byte_perm(a.x, 0, 0x4342), b1 = __byte_perm(b.x, 0, 0x4342);
                const uint32_t a2 = __byte_perm(a.y, 0, 0x4140), b2 = __byte_perm(b.y, 0, 0x4140);
                const uint32_t a3 = __byte_perm(a.y, 0, 0x4342), b3 = __byte_perm(b.y, 0, 0x4342);
                w[0] = relu_add_s16x2(a0 - b0 + 0x00FF00FFu, k.x); w[4] = relu_add_s16x2(b0 - a0 + 0x00FF00FFu, k.x);
                w[1] = relu_add_s16x2(a1 - b1 + 0x00FF00FFu, k.y); w[5] = relu_add_s16x2(b1 - a1 + 0x00FF00FFu, k.y);
                w[2] = relu_add_s16x2(a2 - b2 + 0x00FF00FFu, k.z); w[6] = relu_add_s16x2(b2 - a2 + 0x00FF00FFu, k.z);
                w[3] = relu_add_s16x2(a3 - b3 + 0x00FF00FFu, k.w); w[7] = relu_add_s16x2(b3 - a3 + 0x00FF00FFu, k.w);
            }
            uint32_t v[8];
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint32_t p = w[i] + X[i];
                const uint32_t q = p + P[i];
                const uint32_t rr = q + Q[i];
                v[i] = rr + Rr[i];
                X[i] = w[i]; P[i] = p; Q[i] = q; Rr[i] = rr;
            }
            if (rl >= 4 && rl < nloc) {                                  // warp-uniform: the row completes an output row
                const int yo = y0 + rl - 4;
                uint32_t Wp[6], Wn[6];
                Wp[0] = __shfl_up_sync(0xffffffffu, v[3], 1);
                Wn[0] = __shfl_up_sync(0xffffffffu, v[7], 1);
                Wp[5] = __shfl_down_sync(0xffffffffu, v[0], 1);
                Wn[5] = __shfl_down_sync(0xffffffffu, v[4], 1);
#pragma unroll
                for (int i = 0; i < 4; i++) { Wp[1 + i] = v[i]; Wn[1 + i] = v[4 + i]; }
                // BORDER_REFLECT_101 in x: cols -2,-1 -> 2,1 ; cols, cols+1 -> cols-2, cols-3
                if (left_edge)  { Wp[0] = __byte_perm(v[0], v[1], 0x3254); Wn[0] = __byte_perm(v[4], v[5], 0x3254); }
                if (right_edge) { Wp[5] = __byte_perm(v[2], v[3], 0x3254); Wn[5] = __byte_perm(v[6], v[7], 0x3254); }
                uint32_t Hp[4], Hn[4];
                hpass(Wp, Hp);
                hpass(Wn, Hn);
                const uint32_t p03 = __byte_perm(Hp[0], Hp[1], 0x7531), n03 = __byte_perm(Hn[0], Hn[1], 0x7531);
                const uint32_t p47 = __byte_perm(Hp[2], Hp[3], 0x7531), n47 = __byte_perm(Hn[2], Hn[3], 0x7531);
                uint32_t o03 = __vabsdiffu4(p03, n03), o47 = __vabsdiffu4(p47, n47);
                o03 &= __vcmpgtu4(o03, cut);                             // THRESH_TOZERO (L3Localizer.cpp:252)
                o47 &= __vcmpgtu4(o47, cut);
                const bool hit = out_lane && (o03 | o47) != 0u;
                const bool nz = __any_sync(0xffffffffu, hit);
                if (hit) {                                               // rare: bubble pixels only
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const uint32_t v0 = (o03 >> (8 * c)) & 0xFFu, v1 = (o47 >> (8 * c)) & 0xFFu;
                        if (v0) atomicAdd(&job.hist[v0], 1u);
                        if (v1) atomicAdd(&job.hist[v1], 1u);
                    }
                }
                const uint32_t prev = __shfl_sync(0xffffffffu, fl, r);   // this output row's segment flag
                // the segment is stored when it holds something, or when what the memory holds is not known to be zero
                if ((nz || (prev & 2u)) && out_lane) *reinterpret_cast<uint2*>(job.tz + (size_t)yo * cols + gx) = make_uint2(o03, o47);
                if (lane == r) newfl = nz ? 3u : 0u;
            }
        }
        if (lane < 4 && frow && newfl != fl) flags[(size_t)yf * nxw] = (uint8_t)newfl;
        __syncwarp();                                                   // every lane is done with the stage
        if (lane == 0 && s + G_S < nst) issue(s + G_S, stage);
        if (++stage == G_S) { stage = 0; parity ^= 1u; }
    }
}

}  // namespace

bool abub_k5a_tma_supported(const abub_ctx* ctx) {
    return ctx->d_tmaps != nullptr && (ctx->cols % 16) == 0 && ctx->cols >= 16 && ctx->rows >= 9;
}

// TMA kernel over rows [2, rows-2) of every genesis job of the chunk; the caller adds the four border rows
// (abub_k1_launch_genesis_border).  Uses the host copy of the job table to list the genesis jobs.
int abub_k5a_launch_tma(abub_ctx* ctx, const abub_loc_job* d_jobs, int njobs, int* n_genesis) {
    const abub_loc_job* hj = abub_loc_host_jobs(ctx);
    int ng = 0;
    for (int j = 0; j < njobs; j++) ng += (hj[j].genesis == 1);
    *n_genesis = ng;
    if (ng == 0) return ABUB_OK;
    if (ng > ctx->gen_cap) {
        if (ctx->gen_ev) ABUB_CHECK_CUDA(ctx, cudaEventSynchronize(ctx->gen_ev));
        cudaFree(ctx->d_gen_list);
        if (ctx->h_gen_list) cudaFreeHost(ctx->h_gen_list);
        ctx->d_gen_list = nullptr;
        ctx->h_gen_list = nullptr;
        ctx->gen_cap = 0;
        const int cap = ng + ng / 2 + 64;
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_gen_list, (size_t)cap * sizeof(int)));
        ABUB_CHECK_CUDA(ctx, cudaHostAlloc(&ctx->h_gen_list, (size_t)cap * sizeof(int), cudaHostAllocDefault));
        ctx->gen_cap = cap;
    }
    if (!ctx->gen_ev) ABUB_CHECK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->gen_ev, cudaEventDisableTiming));
    else ABUB_CHECK_CUDA(ctx, cudaEventSynchronize(ctx->gen_ev));      // the previous chunk's upload has left the staging copy
    for (int j = 0, k = 0; j < njobs; j++)
        if (hj[j].genesis == 1) ctx->h_gen_list[k++] = j;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->d_gen_list, ctx->h_gen_list, (size_t)ng * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaEventRecord(ctx->gen_ev, ctx->stream));
    const int nxg = (ctx->cols + G_OW - 1) / G_OW;
    const int nmain = ctx->rows - 4;
    const int nstrips = (nmain + G_TH - 1) / G_TH;
    const int h0 = nmain - G_TH * (nstrips - 1);
    const size_t smem = (size_t)G_S * G_STAGE_BYTES + G_S * 8;
    ABUB_CHECK_CUDA(ctx, cudaFuncSetAttribute(k5_genesis_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const CUtensorMap* maps = reinterpret_cast<const CUtensorMap*>(ctx->d_tmaps);
    for (int p0 = 0; p0 < ng; p0 += 65535) {
        const int n = ng - p0 < 65535 ? ng - p0 : 65535;
        ABUB_L(ctx), k5_genesis_tma_kernel<<<dim3(nstrips * nxg, n), 32, smem, ctx->stream>>>(
            d_jobs, ctx->d_gen_list + p0, maps, maps + ctx->max_slots, ctx->rows, ctx->cols, nxg, nstrips, h0);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    return ABUB_OK;
}
