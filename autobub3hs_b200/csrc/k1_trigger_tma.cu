// K1 (trigger form, TMA-fed) -- AnalyzerUnit::ProcessFrame fused with the 256-bin calcHist
// (reference: AnalyzerUnit.cpp:346-377 and :441-456), batched over every (slot, camera, j).
//
// Same arithmetic as k1_processframe.cu (which stays as the generic-pitch / store-mode path);
// this file is the hot trigger path for pitches that TMA can address (cols % 16 == 0):
//
//   * one warp (= one CTA) owns a strip of 240 output columns x ~350 rows of one frame pair; lane l holds 8 adjacent
//     pixels x0-8+8l .. x0-1+8l (lanes 0 and 31 are the +-2-column halo; the neighbouring strip recomputes them);
//   * rows arrive through a two-stage ring of TMA boxes (cp.async.bulk.tensor) completed on mbarriers: frame j,
//     frame j-d (u8 frames described as a 3-D UINT16 tensor {cols/2, rows, cam*frame}) and the per-pixel constant
//     plane K16 (s16, described as a 2-D UINT32 tensor), each box 272 pixels x T_R rows starting at x0-16.  The
//     box limit of 256 counts ELEMENTS, hence the wider element types; the x origin must be 16-byte aligned (an
//     unaligned origin is an illegal instruction, tools/tma_probe.cu), hence the 16-pixel left margin.  No
//     per-lane global addressing at all; lane 0 re-arms a stage as soon as the warp has consumed it.  TMA
//     zero-fills coordinates outside the image; the two lanes on the image's left/right border reflect in
//     registers (BORDER_REFLECT_101);
//   * strips tile output rows [2, rows-2): the four border rows, whose vertical reflection a streaming strip
//     cannot express, are computed by k1_trigger_kernel (k1_processframe.cu) in border_only mode;
//   * pixels are packed two per 32-bit word, one word per plane.  Stage A:
//         x   = prmt(a) + prmt(~b)                               s16x2 lanes a-b-1+256k (two PRMTs per operand)
//         pos = __viaddmax_s16x2_relu(x, K16, K16)               one VIADDMNMX.S16x2.RELU, K16 = -(255+t) per lane
//         neg = __viaddmax_s16x2_relu(0x01FE01FE - x, K16, K16)
//     with t = min(6 sigma, 255) (2 bytes per pixel of L2-resident constants);
//   * the vertical [1 4 6 4 1] is a cascade of four [1 1] adds on carried state, updated unconditionally (rows beyond
//     the strip are computed and dropped, which keeps the loop branch-free); the horizontal pass takes the two halo
//     words from the neighbour lanes with SHFL and builds the odd-aligned pairs with PRMT 0x5432;
//   * (sum + 128) >> 8 per lane, packed to bytes, |P - N| with VABSDIFF4.U8.  A warp-wide vote skips the histogram
//     for all-zero rows; otherwise all 8 bytes of every lane are added to the shared histogram without branching
//     (bin 0 is a scratch bin) and the true bin 0 is recovered as rows*cols - sum(other bins) by k1_fill_bin0_kernel.
#include <cuda.h>

#include <type_traits>

#include "abub_internal.cuh"

namespace {

#ifndef ABUB_K1T_R
#define ABUB_K1T_R 4
#endif
#ifndef ABUB_K1T_S
#define ABUB_K1T_S 2
#endif
#ifndef ABUB_K1T_MINCTA
#define ABUB_K1T_MINCTA 20
#endif
constexpr int T_R = ABUB_K1T_R;           // rows per TMA box
constexpr int T_S = ABUB_K1T_S;           // ring stages per warp
constexpr int T_OW = 240;                 // output columns per warp (lanes 1..30 x 8 px)
constexpr int T_BW = 272;                 // box width in pixels: [x0-16, x0+256) -- TMA needs a 16-byte aligned start
constexpr int T_A_OFF = 0;                                        // frame j rows      (T_R x 272 B)
constexpr int T_B_OFF = ((T_R * T_BW + 127) / 128) * 128;         // frame j-d rows
constexpr int T_K_OFF = 2 * T_B_OFF;                              // K16 rows          (T_R x 544 B)
constexpr int T_TX_BYTES = T_R * T_BW * 4;                        // bytes one stage receives
constexpr int T_STAGE_BYTES = ((T_K_OFF + T_R * T_BW * 2 + 127) / 128) * 128;
constexpr int T_HB_ENTRIES = 64;          // pending histogram entries per warp: < 32 left over + one row of 30 lanes
constexpr int T_HB_BYTES = T_HB_ENTRIES * 8;
constexpr int T_MAX_WARPS = 1;            // one warp per CTA: nothing to wait for at the end of a strip, and every
                                          // TMA operand (ring, barrier, coordinates) lives in uniform registers
constexpr int T_TH_TARGET = 350;          // output rows per strip (4 halo rows per strip)

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
// relu(x + k) per s16 lane.  k < 0 in both lanes, so max(x + k, k) clamps nothing that the relu would
// not: passing k as the max operand saves materialising a zero register.
__device__ __forceinline__ uint32_t relu_add_s16x2(uint32_t x, uint32_t k) {
    return __viaddmax_s16x2_relu(x, k, k);
}

// Issue the TMA loads of ring stage `s`: image rows y0 - 2 + s*T_R .. +T_R-1 (always >= 0; rows past the
// image bottom are zero-filled and never consumed).  Called by lane 0 only.  xe = (x0-16)/2: the maps
// address the planes as 2-byte (frames) / 4-byte (K16) elements so that one box spans 272 pixels.
__device__ __forceinline__ void k1t_issue(const CUtensorMap* smap, const CUtensorMap* kmap, uint32_t stage_base,
                                          uint32_t bar, int s, int y0, int xe, int za, int zb) {
    const int ys = y0 - 2 + s * T_R;
    mbar_expect_tx(bar, T_TX_BYTES);
    tma_load_3d(stage_base + T_A_OFF, smap, xe, ys, za, bar);
    tma_load_3d(stage_base + T_B_OFF, smap, xe, ys, zb, bar);
    tma_load_2d(stage_base + T_K_OFF, kmap, xe, ys, bar);
}

// horizontal [1 4 6 4 1] of one plane: W[0..5] = words -1..4 (two pixels each), out H[0..3] incl. +128
__device__ __forceinline__ void k1t_hpass(const uint32_t W[6], uint32_t H[4]) {
    uint32_t S[5];
#pragma unroll
    for (int k = 0; k < 5; k++) S[k] = __byte_perm(W[k], W[k + 1], 0x5432);    // (hi of W[k], lo of W[k+1])
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t e = W[k] + W[k + 2] + 0x00800080u;
        const uint32_t t = S[k] + S[k + 1];
        H[k] = e + 6u * W[k + 1] + 4u * t;
    }
}

// grid: x = strip * nxg + xg, y = j-1, z = (slot-slot_first)*ncams + cam ; block = 32 * warps_per_cta
__global__ void __launch_bounds__(32 * T_MAX_WARPS, ABUB_K1T_MINCTA)
k1_trigger_tma_kernel(const CUtensorMap* __restrict__ slot_maps, const CUtensorMap* __restrict__ k16_maps,
                      const int* __restrict__ nframes, const int* __restrict__ tss, uint32_t* __restrict__ hist,
                      int slot_first, int rows, int cols, int ncams, int max_frames, int th, int nxg, int j_first) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const int warp = (T_MAX_WARPS == 1) ? 0 : (int)(threadIdx.x >> 5), lane = threadIdx.x & 31;
    const int wpc = (T_MAX_WARPS == 1) ? 1 : (int)(blockDim.x >> 5);
    // layout: [ring stages, 128-aligned][hist 256 u32][pending entries T_HB_ENTRIES x 8 B + counter per warp][mbarriers]
    const uint32_t ring = smem_u32(smem_raw);
    uint8_t* const ring_p = smem_raw;
    uint32_t* const sh_hist = reinterpret_cast<uint32_t*>(ring_p + (size_t)wpc * T_S * T_STAGE_BYTES);
    const uint32_t bars = ring + (uint32_t)wpc * T_S * T_STAGE_BYTES + 1024u + (uint32_t)wpc * T_HB_BYTES + (uint32_t)warp * T_S * 8u;
    const uint32_t my_ring = ring + (uint32_t)warp * T_S * T_STAGE_BYTES;
    const uint8_t* const my_ring_p = ring_p + (size_t)warp * T_S * T_STAGE_BYTES;

    const int rel = blockIdx.z;
    const int slot = slot_first + rel / ncams;
    const int cam = rel % ncams;
    const int j = blockIdx.y + j_first;
    const int nf = nframes[slot * ncams + cam];
    if (j >= nf) return;
    const int d = (tss[cam] < 6) ? 1 : 2;           // twoFrameOffset, AnalyzerUnit.cpp:182-188
    const int jp = (j - d < 0) ? 0 : j - d;         // j==1 uses frame 0 (:176)

    for (int i = threadIdx.x; i < 256; i += blockDim.x) sh_hist[i] = 0;
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < T_S; s++) mbar_init(bars + s * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    // strips tile the output rows [2, rows-2): no strip ever needs a reflected row (the two border
    // rows at the top and bottom go through the generic kernel, see abub_k1_launch_trigger)
    const int strip = blockIdx.x / nxg, xg = blockIdx.x % nxg;
    const int y0 = 2 + strip * th;
    const int x0 = (xg * wpc + warp) * T_OW;
    if (y0 < rows - 2 && x0 < cols) {
        const int nout = min(th, rows - 2 - y0);
        const int nloc = nout + 4;
        const int nst = (nloc + T_R - 1) / T_R;
        const int xe = (x0 - 16) / 2;
        const int gx = x0 - 8 + lane * 8;
        const bool out_lane = lane >= 1 && lane <= 30 && gx < cols;
        const bool left_edge = (gx == 0), right_edge = (gx + 8 == cols);
        const bool edge_warp = (x0 == 0) || (x0 + T_OW >= cols);
        const CUtensorMap* smap = slot_maps + slot;
        const CUtensorMap* kmap = k16_maps + cam;
        const int za = cam * max_frames + j, zb = cam * max_frames + jp;

        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < T_S; s++)
                if (s < nst) k1t_issue(smap, kmap, my_ring + s * T_STAGE_BYTES, bars + s * 8, s, y0, xe, za, zb);
        }
        // carried state of the vertical cascade: X = previous row, P,Q,Rr = partial binomials
        uint32_t X[8], P[8], Q[8], Rr[8];
#pragma unroll
        for (int i = 0; i < 8; i++) X[i] = P[i] = Q[i] = Rr[i] = 0;

        int stage = 0;
        uint32_t parity = 0;
        // Non-zero outputs are rare (about 2.6 lanes of 32 per row on the benchmark's noise): a lane whose 8 packed outputs
        // are not all zero appends them to a per-warp list (slot from a shared counter), and the list is histogrammed 32
        // entries at a time -- all eight bytes of an entry are added, zeros included (bin 0 of the shared table is scratch,
        // the true bin 0 is rows*cols - sum of the others): no branches or predicates in the byte-wise atomics.
        uint2* const hbuf = reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(sh_hist + 256) + (size_t)warp * T_HB_BYTES);
        int hcnt = 0;                                  // pending entries (warp-uniform)
        const uint32_t lt_mask = (1u << lane) - 1u;
        auto hist32 = [&](const uint2 e) {
#pragma unroll
            for (int c = 0; c < 4; c++) {
                atomicAdd(&sh_hist[__byte_perm(e.x, 0, 0x4440 + c)], 1u);
                atomicAdd(&sh_hist[__byte_perm(e.y, 0, 0x4440 + c)], 1u);
            }
        };
        // 32 pending entries go to the histogram, the rest (< 32) moves to the front
        auto drain = [&]() {
            __syncwarp();
            hist32(hbuf[lane]);
            hcnt -= 32;
            const uint2 mv = hbuf[32 + lane];
            __syncwarp();
            if (lane < hcnt) hbuf[lane] = mv;
        };
        // one ring stage = T_R image rows.  CHK = the stage may hold rows without an output (the first four rows of the
        // strip only prime the cascade, rows past nloc are the tail of the last box): interior stages skip the tests.
        auto do_stage = [&](auto chk, const int s, const uint8_t* const st) {
            constexpr bool CHK = decltype(chk)::value;
#pragma unroll
            for (int r = 0; r < T_R; r++) {
                const int rl = s * T_R + r;
                const uint2 a = *reinterpret_cast<const uint2*>(st + T_A_OFF + r * T_BW + 8 + lane * 8);
                const uint2 b = *reinterpret_cast<const uint2*>(st + T_B_OFF + r * T_BW + 8 + lane * 8);
                const uint4 k = *reinterpret_cast<const uint4*>(st + T_K_OFF + r * 2 * T_BW + 16 + lane * 16);
                uint32_t w[8];
                {
                    // s16x2 lanes a - b + 255 and b - a + 255 (both lanes >= 0: the 32-bit difference cannot borrow across)
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
                // rows without an output are computed and dropped: keeping the state update unconditional lets the
                // cascade registers rotate without moves
                if (!CHK || (rl >= 4 && rl < nloc)) {
                    uint32_t Wp[6], Wn[6];
                    Wp[0] = __shfl_up_sync(0xffffffffu, v[3], 1);
                    Wn[0] = __shfl_up_sync(0xffffffffu, v[7], 1);
                    Wp[5] = __shfl_down_sync(0xffffffffu, v[0], 1);
                    Wn[5] = __shfl_down_sync(0xffffffffu, v[4], 1);
#pragma unroll
                    for (int i = 0; i < 4; i++) { Wp[1 + i] = v[i]; Wn[1 + i] = v[4 + i]; }
                    if (edge_warp) {
                        // BORDER_REFLECT_101 in x: cols -2,-1 -> 2,1 ; cols, cols+1 -> cols-2, cols-3
                        if (left_edge)  { Wp[0] = __byte_perm(v[0], v[1], 0x3254); Wn[0] = __byte_perm(v[4], v[5], 0x3254); }
                        if (right_edge) { Wp[5] = __byte_perm(v[2], v[3], 0x3254); Wn[5] = __byte_perm(v[6], v[7], 0x3254); }
                    }
                    uint32_t Hp[4], Hn[4];
                    k1t_hpass(Wp, Hp);
                    k1t_hpass(Wn, Hn);
                    // P = byte 1 / byte 3 of each word
                    const uint32_t p03 = __byte_perm(Hp[0], Hp[1], 0x7531), n03 = __byte_perm(Hn[0], Hn[1], 0x7531);
                    const uint32_t p47 = __byte_perm(Hp[2], Hp[3], 0x7531), n47 = __byte_perm(Hn[2], Hn[3], 0x7531);
                    const uint32_t o03 = __vabsdiffu4(p03, n03), o47 = __vabsdiffu4(p47, n47);
                    const bool hit = out_lane && (o03 | o47) != 0u;
                    const uint32_t hm = __ballot_sync(0xffffffffu, hit);
                    if (hm) {
                        if (hit) hbuf[hcnt + __popc(hm & lt_mask)] = make_uint2(o03, o47);
                        hcnt += __popc(hm);
                        if (hcnt >= 32) drain();
                    }
                }
            }
        };
#pragma unroll 1
        for (int s = 0; s < nst; s++) {
            mbar_wait(bars + stage * 8, parity);
            const uint8_t* st = my_ring_p + stage * T_STAGE_BYTES;
            if (s >= 1 && (s + 1) * T_R <= nloc) do_stage(std::false_type{}, s, st);
            else do_stage(std::true_type{}, s, st);
            __syncwarp();       // every lane is done with the stage
            if (lane == 0 && s + T_S < nst)
                k1t_issue(smap, kmap, my_ring + stage * T_STAGE_BYTES, bars + stage * 8, s + T_S, y0, xe, za, zb);
            if (++stage == T_S) { stage = 0; parity ^= 1u; }
        }
        __syncwarp();
        hist32((lane < hcnt) ? hbuf[lane] : make_uint2(0u, 0u));
    }
    __syncthreads();
    uint32_t* H = hist + (((size_t)slot * ncams + cam) * max_frames + j) * 256;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        const uint32_t c = sh_hist[i];
        if (c && i) atomicAdd(&H[i], c);       // bin 0 of the shared table is scratch
    }
}

// K16 = -(255 + t) as s16, t = min(6*sigma, 255)
__global__ void k1_build_k16_kernel(const uint8_t* __restrict__ sigma, int16_t* __restrict__ k16, int16_t* __restrict__ nt16, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int t = min(6 * (int)sigma[i], 255);
    k16[i] = (int16_t)(-(255 + t));
    nt16[i] = (int16_t)(-t);
}

}  // namespace

// ---- host side: tensor maps -----------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
    static PFN_encodeTiled fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(p);
    }
    return fn;
}

bool abub_k1_tma_supported(const abub_ctx* ctx) { return (ctx->cols % 16) == 0 && ctx->cols >= 16 && ctx->rows >= 5; }

// Builds the device array of tensor maps: [slot] then [cam].  Called once from abub_ctx_create.
// The u8 frames are described as 2-byte elements and the s16 constant plane as 4-byte elements: a box
// is limited to 256 elements per dimension and one box must span 272 pixels.
int abub_k1_tma_init(abub_ctx* ctx) {
    ctx->d_tmaps = nullptr;
    if (!abub_k1_tma_supported(ctx)) return ABUB_OK;
    PFN_encodeTiled enc = get_encode();
    if (!enc) { ctx->err = "cuTensorMapEncodeTiled not available from the driver"; return ABUB_E_CUDA; }
    const int nmaps = ctx->max_slots + 3 * ctx->ncams;
    std::vector<CUtensorMap> maps(nmaps);
    const cuuint32_t ones[3] = {1, 1, 1};
    for (int s = 0; s < ctx->max_slots; s++) {
        const cuuint64_t dims[3] = {(cuuint64_t)ctx->cols / 2, (cuuint64_t)ctx->rows, (cuuint64_t)ctx->ncams * ctx->max_frames};
        const cuuint64_t strides[2] = {(cuuint64_t)ctx->cols, (cuuint64_t)ctx->npix};
        const cuuint32_t box[3] = {(cuuint32_t)T_BW / 2, (cuuint32_t)T_R, 1};
        CUresult r = enc(&maps[s], CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, ctx->slots[s].d_frames, dims, strides, box, ones,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { ctx->err = "cuTensorMapEncodeTiled(slot) failed: " + std::to_string((int)r); return ABUB_E_CUDA; }
    }
    for (int c = 0; c < ctx->ncams; c++) {
        const cuuint64_t dims[2] = {(cuuint64_t)ctx->cols / 2, (cuuint64_t)ctx->rows};
        const cuuint64_t strides[1] = {(cuuint64_t)ctx->cols * 2};
        const cuuint32_t box[2] = {(cuuint32_t)T_BW / 2, (cuuint32_t)T_R};
        CUresult r = enc(&maps[ctx->max_slots + c], CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, ctx->cam[c].d_k16, dims, strides, box,
                         ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { ctx->err = "cuTensorMapEncodeTiled(k16) failed: " + std::to_string((int)r); return ABUB_E_CUDA; }
    }
    // localizer post-trigger kernel (k5_post_tma.cu): TrainedAvgImage as 2-byte elements, -t (s16) as 4-byte elements
    for (int c = 0; c < ctx->ncams; c++) {
        const cuuint64_t dims[2] = {(cuuint64_t)ctx->cols / 2, (cuuint64_t)ctx->rows};
        const cuuint32_t box[2] = {(cuuint32_t)T_BW / 2, (cuuint32_t)T_R};
        const cuuint64_t s16[1] = {(cuuint64_t)ctx->cols};
        CUresult r = enc(&maps[ctx->max_slots + ctx->ncams + c], CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, ctx->cam[c].d_mean, dims, s16, box,
                         ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { ctx->err = "cuTensorMapEncodeTiled(mean) failed: " + std::to_string((int)r); return ABUB_E_CUDA; }
        const cuuint64_t s32[1] = {(cuuint64_t)ctx->cols * 2};
        r = enc(&maps[ctx->max_slots + 2 * ctx->ncams + c], CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, ctx->cam[c].d_nt16, dims, s32, box,
                ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { ctx->err = "cuTensorMapEncodeTiled(nt16) failed: " + std::to_string((int)r); return ABUB_E_CUDA; }
    }
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_tmaps, (size_t)nmaps * sizeof(CUtensorMap)));
    ABUB_CHECK_CUDA(ctx, cudaMemcpy(ctx->d_tmaps, maps.data(), (size_t)nmaps * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
    return ABUB_OK;
}

int abub_k1_build_k16(abub_ctx* ctx, int cam) {
    abub_cam_state& c = ctx->cam[cam];
    const size_t n = ctx->npix;
    ABUB_L(ctx), k1_build_k16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(c.d_sigma, c.d_k16, c.d_nt16, n);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

static inline int k1t_wpc(int cols) {
    const int nw = (cols + T_OW - 1) / T_OW;
    return nw < T_MAX_WARPS ? nw : T_MAX_WARPS;
}

// launches the TMA trigger kernel over [slot_first, slot_first+nslots); the caller zeroes the
// histogram table before and runs k1_fill_bin0_kernel after.
int abub_k1_launch_trigger_tma(abub_ctx* ctx, int slot_first, int nslots, int j_first, int j_last, const int* d_limit) {
    const int wpc = k1t_wpc(ctx->cols);
    const int nw = (ctx->cols + T_OW - 1) / T_OW;
    const int nxg = (nw + wpc - 1) / wpc;
    const int nmain = ctx->rows - 4;              // output rows [2, rows-2)
    const int nstrips = (nmain + T_TH_TARGET - 1) / T_TH_TARGET;
    const int th = (nmain + nstrips - 1) / nstrips;
    const size_t smem = (size_t)wpc * T_S * T_STAGE_BYTES + 1024 + (size_t)wpc * T_HB_BYTES + (size_t)wpc * T_S * 8;
    ABUB_CHECK_CUDA(ctx, cudaFuncSetAttribute(k1_trigger_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const CUtensorMap* maps = reinterpret_cast<const CUtensorMap*>(ctx->d_tmaps);
    const int per = 65535 / ctx->ncams;
    for (int s0 = 0; s0 < nslots; s0 += per) {
        const int ns = (nslots - s0 < per) ? nslots - s0 : per;
        dim3 grid(nstrips * nxg, j_last - j_first, ns * ctx->ncams);
        ABUB_L(ctx), k1_trigger_tma_kernel<<<grid, 32 * wpc, smem, ctx->stream>>>(
            maps, maps + ctx->max_slots, d_limit, ctx->d_tss, ctx->d_hist, slot_first + s0, ctx->rows,
            ctx->cols, ctx->ncams, ctx->max_frames, th, nxg, j_first);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    return ABUB_OK;
}
