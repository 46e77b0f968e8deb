// K1 -- AnalyzerUnit::ProcessFrame fused with the 256-bin calcHist
// (reference: AnalyzerUnit.cpp:346-377 and :441-456).
//
//   pos = sat_u8(a - b - 6*sigma), neg = sat_u8(b - a - 6*sigma)        (:351-352)
//   P = GaussianBlur5x5(pos), N = GaussianBlur5x5(neg)                  (:359-360)
//       = (sum_ij k_i k_j p + 128) >> 8, k = [1 4 6 4 1], BORDER_REFLECT_101
//   out = |P - N|                                                        (:370)
//   H[out] += 1 over the whole frame, no mask                            (:456)
//
// Layout of the computation (HBM-bound by design, integer-issue-bound in practice):
//   * one warp owns a strip of 120 output columns x TH rows; lane l holds the 4 pixels
//     x0-4+4l .. x0-1+4l (lanes 0 and 31 are the 2-px halo, rounded up to a word);
//   * every pixel is one 32-bit word  pos | neg<<16 : the binomial sums of both planes
//     stay below 2^16 (255*256 + 128), so plain IADD/IMAD blur both planes at once;
//   * stage A (a,b,6sigma -> word) uses dp4a on [a_i,b_i,a_j,b_j] byte pairs and one IMAD
//     against a per-pixel constant plane K32 = (255-t) | (765-t)<<16, t = min(6 sigma,255):
//        y = (a-b+255) * 0xFFFF0001 + K32  ->  lanes [a-b-t+510 , b-a-t+510]
//        word = max.u16x2(y, 510) - 510
//   * vertical 5-tap pass in registers (5-row sliding window), horizontal pass after a
//     conflict-free 8-byte neighbour exchange through shared memory (one __syncwarp);
//   * only non-zero outputs touch the shared-memory histogram; bin 0 is recovered as
//     rows*cols - sum(other bins) by a finalize kernel.
#include <stdlib.h>

#include "abub_internal.cuh"

namespace {

constexpr uint32_t BIAS2 = 0x01FE01FEu;   // 510 | 510<<16
constexpr uint32_t RND2  = 0x00800080u;   // 128 | 128<<16

__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ int reflect101(int p, int n) {
    if (p < 0) p = -p;
    if (p >= n) p = 2 * (n - 1) - p;
    return min(max(p, 0), n - 1);
}

__device__ __forceinline__ uint32_t stageA1(int s255, uint32_t k) {
    uint32_t y = (uint32_t)s255 * 0xFFFF0001u + k;
    return __vmaxu2(y, BIAS2) - BIAS2;
}

// words of 4 neighbouring pixels from packed bytes a4,b4 and their K32 constants
__device__ __forceinline__ void stageA(uint32_t a4, uint32_t b4, const uint4 k, uint32_t w[4]) {
    const uint32_t p01 = __byte_perm(a4, b4, 0x5140);   // [a0,b0,a1,b1]
    const uint32_t p23 = __byte_perm(a4, b4, 0x7362);   // [a2,b2,a3,b3]
    w[0] = stageA1(dp4a_us(p01, 0x0000FF01, 255), k.x);
    w[1] = stageA1(dp4a_us(p01, (int)0xFF010000, 255), k.y);
    w[2] = stageA1(dp4a_us(p23, 0x0000FF01, 255), k.z);
    w[3] = stageA1(dp4a_us(p23, (int)0xFF010000, 255), k.w);
}

struct RowIn { uint32_t a4, b4; uint4 k; };

__device__ __forceinline__ RowIn load_row(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B,
                                          const uint32_t* __restrict__ K, int y, int gx, int cols, bool fast) {
    RowIn r;
    const size_t ro = (size_t)y * cols;
    if (fast) {
        r.a4 = __ldg(reinterpret_cast<const uint32_t*>(A + ro + gx));
        r.b4 = __ldg(reinterpret_cast<const uint32_t*>(B + ro + gx));
        r.k  = __ldg(reinterpret_cast<const uint4*>(K + ro + gx));
    } else {
        uint32_t a = 0, b = 0, kk[4];
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const int x = reflect101(gx + c, cols);
            a |= (uint32_t)A[ro + x] << (8 * c);
            b |= (uint32_t)B[ro + x] << (8 * c);
            kk[c] = K[ro + x];
        }
        r.a4 = a; r.b4 = b; r.k = make_uint4(kk[0], kk[1], kk[2], kk[3]);
    }
    return r;
}

// One warp: strip of ABUB_K1_TILE_W output columns starting at x0, output rows [y0, y0+th).
// MODE 0: histogram of the outputs only.  MODE 1: also store the (optionally TOZERO'd) output.
template <int MODE>
__device__ __forceinline__ void k1_strip(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B,
                                         const uint32_t* __restrict__ K, int rows, int cols,
                                         int y0, int th, int x0, uint32_t* sh_hist, uint32_t* xch,
                                         uint8_t* __restrict__ out, int tozero) {
    const int lane = threadIdx.x & 31;
    const int gx = x0 - 4 + lane * 4;
    const bool fast = ((cols & 3) == 0) && gx >= 0 && gx + 4 <= cols;
    const bool out_lane = lane >= 1 && lane <= 30 && gx < cols;
    const int nout = min(th, rows - y0);
    const int nloc = nout + 4;
    uint32_t win[5][4];
    RowIn nxt = load_row(A, B, K, reflect101(y0 - 2, rows), gx, cols, fast);

#pragma unroll 1
    for (int rb = 0; rb < nloc; rb += 5) {
#pragma unroll
        for (int r = 0; r < 5; r++) {
            const int rl = rb + r;
            if (rl < nloc) {
                const RowIn cur = nxt;
                if (rl + 1 < nloc) nxt = load_row(A, B, K, reflect101(y0 - 2 + rl + 1, rows), gx, cols, fast);
                stageA(cur.a4, cur.b4, cur.k, win[r]);
                if (rl >= 4) {
                    // vertical pass: rows rl-4..rl live in win[(r+1)%5] .. win[r]
                    uint32_t v[4];
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const uint32_t e = win[(r + 1) % 5][c] + win[r][c];
                        const uint32_t o = win[(r + 2) % 5][c] + win[(r + 4) % 5][c];
                        v[c] = e + 4u * o + 6u * win[(r + 3) % 5][c];
                    }
                    // neighbour exchange (lo pair = cols 0,1 ; hi pair = cols 2,3)
                    uint2* xlo = reinterpret_cast<uint2*>(xch + (rl & 1) * 128);
                    uint2* xhi = xlo + 32;
                    xlo[lane] = make_uint2(v[0], v[1]);
                    xhi[lane] = make_uint2(v[2], v[3]);
                    __syncwarp();
                    const uint2 L = xhi[(lane + 31) & 31];
                    const uint2 R = xlo[(lane + 1) & 31];
                    const uint32_t e0 = L.x, e1 = L.y, e6 = R.x, e7 = R.y;
                    uint32_t h[4];
                    h[0] = (e0 + v[2]) + 4u * (e1 + v[1]) + 6u * v[0] + RND2;
                    h[1] = (e1 + v[3]) + 4u * (v[0] + v[2]) + 6u * v[1] + RND2;
                    h[2] = (v[0] + e6) + 4u * (v[1] + v[3]) + 6u * v[2] + RND2;
                    h[3] = (v[1] + e7) + 4u * (v[2] + e6) + 6u * v[3] + RND2;
                    const int y = y0 + rl - 4;
                    const uint32_t z = (h[0] | h[1] | h[2] | h[3]) & 0xFF00FF00u;
                    uint32_t packed = 0;
                    if (z != 0 && out_lane) {
#pragma unroll
                        for (int c = 0; c < 4; c++) {
                            const uint32_t P = (h[c] >> 8) & 0xFFu, N = h[c] >> 24;
                            uint32_t o = __usad(P, N, 0u);
                            if (MODE == 1 && (int)o <= tozero) o = 0;
                            if (o != 0 && gx + c < cols) {
                                atomicAdd(&sh_hist[o], 1u);
                                packed |= o << (8 * c);
                            }
                        }
                    }
                    if (MODE == 1 && out_lane) {
                        uint8_t* op = out + (size_t)y * cols + gx;
                        if (fast) {
                            *reinterpret_cast<uint32_t*>(op) = packed;
                        } else {
#pragma unroll
                            for (int c = 0; c < 4; c++)
                                if (gx + c < cols) op[c] = (uint8_t)(packed >> (8 * c));
                        }
                    }
                }
            }
        }
    }
}


// ---- fast strip (cols % 4 == 0) -------------------------------------------------------
// Same arithmetic as k1_strip, restructured for instruction count:
//  * word loads only (lanes outside the image load a clamped address and are ignored); column
//    reflection is a register select on the two lanes that touch the image border;
//  * row pointers step by +-pitch (the sign flips inside the reflected rows when EDGE_Y);
//  * stage A keeps the +510 lane bias (word' = max.u16x2(y,510)); the vertical pass removes
//    16*510 from both lanes with one constant folded into an IADD3 -- plain 32-bit adds are
//    exact mod 2^32, so transient lane borrows cancel as long as the final lanes are in range;
//  * 10-row unroll: the 5-row window, the 2-deep load prefetch and the exchange double buffer
//    all get static register names;
//  * histogram updates are red.shared on a precomputed shared-window address.
__device__ __forceinline__ uint32_t stageA1_biased(int s255, uint32_t k) {
    return __vmaxu2((uint32_t)s255 * 0xFFFF0001u + k, BIAS2);
}
__device__ __forceinline__ void stageA_biased(uint32_t a4, uint32_t b4, const uint4 k, uint32_t w[4]) {
    const uint32_t p01 = __byte_perm(a4, b4, 0x5140);
    const uint32_t p23 = __byte_perm(a4, b4, 0x7362);
    w[0] = stageA1_biased(dp4a_us(p01, 0x0000FF01, 255), k.x);
    w[1] = stageA1_biased(dp4a_us(p01, (int)0xFF010000, 255), k.y);
    w[2] = stageA1_biased(dp4a_us(p23, 0x0000FF01, 255), k.z);
    w[3] = stageA1_biased(dp4a_us(p23, (int)0xFF010000, 255), k.w);
}
__device__ __forceinline__ void cp_async4(uint32_t saddr, const void* g) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(saddr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t saddr, const void* g) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(saddr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait4() { asm volatile("cp.async.wait_group 4;" ::: "memory"); }
__device__ __forceinline__ void red_shared_inc(uint32_t saddr) {
    asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(saddr) : "memory");
}

template <int MODE, bool EDGE_Y>
__device__ __forceinline__ void k1_strip_fast(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B,
                                              const uint32_t* __restrict__ K, int rows, int cols, int y0, int th,
                                              int x0, uint32_t* sh_hist, uint32_t* xch, uint8_t* __restrict__ out,
                                              int tozero) {
    constexpr uint32_t FIX = RND2 - 256u * 510u * 0x00010001u;          // mod 2^32
    const int lane = threadIdx.x & 31;
    const int gx = x0 - 4 + lane * 4;
    const int gxl = min(max(gx, 0), cols - 4);                 // clamped load column
    const bool out_lane = lane >= 1 && lane <= 30 && gx < cols;
    const bool left_edge = (gx == 0), right_edge = (gx + 4 == cols);
    const int nout = min(th, rows - y0);
    const int nloc = nout + 4;
    const long long pitch = cols;
    const int yfirst = EDGE_Y ? reflect101(y0 - 2, rows) : y0 - 2;
    const uint8_t* pa = A + (size_t)yfirst * cols + gxl;
    const uint8_t* pb = B + (size_t)yfirst * cols + gxl;
    const uint32_t* pk = K + (size_t)yfirst * cols + gxl;
    uint8_t* po = (MODE == 1) ? out + (size_t)y0 * cols + gxl : nullptr;
    const uint32_t hist_s = (uint32_t)__cvta_generic_to_shared(sh_hist);
    uint2* const xlo = reinterpret_cast<uint2*>(xch) + lane;        // (v0,v1) of this lane, buffer 0
    uint2* const xhi = xlo + 32;                                    // (v2,v3)
    const uint2* const xL = reinterpret_cast<const uint2*>(xch) + 32 + ((lane + 31) & 31);
    const uint2* const xR = reinterpret_cast<const uint2*>(xch) + ((lane + 1) & 31);
    uint32_t win[5][4];
    // 5-stage cp.async ring per warp (after the 256 exchange words): stage s holds one row of this
    // warp: a[32] | b[32] | k32[128] words.  Every lane copies and later reads only its own words,
    // so cp.async.wait_group is the only synchronisation the ring needs.
    const uint32_t* const stg = xch + 256;
    const uint32_t stg_a = (uint32_t)__cvta_generic_to_shared(stg) + lane * 4u;
    const uint32_t stg_k = (uint32_t)__cvta_generic_to_shared(stg) + 256u + lane * 16u;
#pragma unroll
    for (int s0 = 0; s0 < 4; s0++) {
        if (s0 < nloc) {
            if (s0 > 0) {
                long long step = pitch;
                if (EDGE_Y) { const int yn = y0 - 2 + s0; if (yn <= 0 || yn >= rows) step = -pitch; }
                pa += step; pb += step; pk += step;
            }
            cp_async4(stg_a + s0 * 768u, pa);
            cp_async4(stg_a + s0 * 768u + 128u, pb);
            cp_async16(stg_k + s0 * 768u, pk);
        }
        cp_async_commit();
    }
#pragma unroll 1
    for (int rb0 = 0; rb0 < nloc; rb0 += 10) {
#pragma unroll
        for (int r = 0; r < 10; r++) {
            const int rl = rb0 + r;
            if (rl < nloc) {
                if (rl + 4 < nloc) {
                    long long step = pitch;
                    if (EDGE_Y) {
                        const int yn = y0 - 2 + rl + 4;             // unreflected row being fetched
                        if (yn <= 0 || yn >= rows) step = -pitch;   // BORDER_REFLECT_101 walks backwards there
                    }
                    pa += step; pb += step; pk += step;
                    cp_async4(stg_a + ((r + 4) % 5) * 768u, pa);
                    cp_async4(stg_a + ((r + 4) % 5) * 768u + 128u, pb);
                    cp_async16(stg_k + ((r + 4) % 5) * 768u, pk);
                }
                cp_async_commit();
                cp_async_wait4();
                const uint32_t ra_ = stg[(r % 5) * 192 + lane];
                const uint32_t rb_ = stg[(r % 5) * 192 + 32 + lane];
                const uint4 rk_ = *reinterpret_cast<const uint4*>(stg + (r % 5) * 192 + 64 + lane * 4);
                stageA_biased(ra_, rb_, rk_, win[r % 5]);
                if (rl >= 4) {
                    // vertical pass on biased words (each lane carries +16*510 afterwards)
                    uint32_t v[4];
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const uint32_t e = win[(r + 1) % 5][c] + win[r % 5][c];
                        const uint32_t o = win[(r + 2) % 5][c] + win[(r + 4) % 5][c];
                        v[c] = e + 4u * o + 6u * win[(r + 3) % 5][c];
                    }
                    xlo[(r & 1) * 64] = make_uint2(v[0], v[1]);
                    xhi[(r & 1) * 64] = make_uint2(v[2], v[3]);
                    __syncwarp();
                    const uint2 L = xL[(r & 1) * 64];
                    const uint2 R = xR[(r & 1) * 64];
                    // BORDER_REFLECT_101 in x: col -2 -> 2, -1 -> 1 ; cols -> cols-2, cols+1 -> cols-3
                    const uint32_t e0 = left_edge ? v[2] : L.x, e1 = left_edge ? v[1] : L.y;
                    const uint32_t e6 = right_edge ? v[2] : R.x, e7 = right_edge ? v[1] : R.y;
                    // horizontal pass; FIX = +128 rounding - 256*510 bias in both lanes (mod 2^32: the
                    // transient lane overflow of the biased sums cancels exactly)
                    uint32_t h[4];
                    h[0] = (e0 + v[2]) + 4u * (e1 + v[1]) + 6u * v[0] + FIX;
                    h[1] = (e1 + v[3]) + 4u * (v[0] + v[2]) + 6u * v[1] + FIX;
                    h[2] = (v[0] + e6) + 4u * (v[1] + v[3]) + 6u * v[2] + FIX;
                    h[3] = (v[1] + e7) + 4u * (v[2] + e6) + 6u * v[3] + FIX;
                    const uint32_t z = (h[0] | h[1] | h[2] | h[3]) & 0xFF00FF00u;
                    uint32_t packed = 0;
                    if (z != 0 && out_lane) {
                        // P = byte 1, N = byte 3 of each h: gather the four P's and N's, one VABSDIFF4
                        const uint32_t p4 = __byte_perm(__byte_perm(h[0], h[1], 0x0051), __byte_perm(h[2], h[3], 0x0051), 0x5410);
                        const uint32_t n4 = __byte_perm(__byte_perm(h[0], h[1], 0x0073), __byte_perm(h[2], h[3], 0x0073), 0x5410);
                        uint32_t o4 = __vabsdiffu4(p4, n4);
                        if (MODE == 1) {
#pragma unroll
                            for (int c = 0; c < 4; c++)
                                if ((int)((o4 >> (8 * c)) & 0xFFu) <= tozero) o4 &= ~(0xFFu << (8 * c));
                        }
                        packed = o4;
#pragma unroll
                        for (int c = 0; c < 4; c++) {
                            const uint32_t o = (o4 >> (8 * c)) & 0xFFu;
                            if (o != 0) red_shared_inc(hist_s + o * 4u);
                        }
                    }
                    if (MODE == 1) {
                        if (out_lane) *reinterpret_cast<uint32_t*>(po) = packed;
                        po += pitch;
                    }
                }
            }
        }
    }
}

// strip dispatcher: fast path when the pitch allows aligned word access
template <int MODE>
__device__ __forceinline__ void k1_strip_any(const uint8_t* A, const uint8_t* B, const uint32_t* K, int rows, int cols,
                                             int y0, int th, int x0, uint32_t* sh_hist, uint32_t* xch, uint8_t* out,
                                             int tozero) {
    if ((cols & 3) == 0 && cols >= 8) {
        const bool edge_y = (y0 < 2) || (y0 + th + 2 > rows);
        if (edge_y) k1_strip_fast<MODE, true>(A, B, K, rows, cols, y0, th, x0, sh_hist, xch, out, tozero);
        else        k1_strip_fast<MODE, false>(A, B, K, rows, cols, y0, th, x0, sh_hist, xch, out, tozero);
    } else {
        k1_strip<MODE>(A, B, K, rows, cols, y0, th, x0, sh_hist, xch, out, tozero);
    }
}

constexpr int K1_MAX_THREADS = 512;             // 16 warps = 1920 columns per CTA; wider images add CTAs in x
constexpr int K1_WARP_WORDS = 256 + 5 * 192;   // exchange double buffer + 5-stage cp.async ring
constexpr int K1_TH_TARGET = 105;   // output rows per strip (1050 = 10 x 105): 4 halo rows per 105

// grid: x = strip, y = j-1, z = (slot-slot_first)*ncams + cam ; block = 32 * ceil(cols/120)
__global__ void __launch_bounds__(K1_MAX_THREADS, 2) k1_trigger_kernel(uint8_t* const* __restrict__ slot_ptrs, const int* __restrict__ nframes,
                                  const int* __restrict__ tss, uint32_t* const* __restrict__ k32_ptrs,
                                  uint32_t* __restrict__ hist, int slot_first, int rows, int cols,
                                  int ncams, int max_frames, int th, int nxg, int border_only, int j_first) {
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* sh_hist = smem;                       // 256
    const int warp = threadIdx.x >> 5;
    uint32_t* xch = smem + 256 + warp * K1_WARP_WORDS;        // 2 x 128 words per warp
    const int rel = blockIdx.z;
    const int slot = slot_first + rel / ncams;
    const int cam = rel % ncams;
    const int j = blockIdx.y + j_first;
    const int nf = nframes[slot * ncams + cam];
    if (j >= nf) return;
    const int d = (tss[cam] < 6) ? 1 : 2;           // twoFrameOffset, AnalyzerUnit.cpp:182-188
    const int jp = (j - d < 0) ? 0 : j - d;         // j==1 uses frame 0 (:176)
    const size_t npix = (size_t)rows * cols;
    const uint8_t* base = slot_ptrs[slot] + (size_t)cam * max_frames * npix;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sh_hist[i] = 0;
    __syncthreads();
    // border_only: two strips of `th` (= 2) rows at the top and bottom of the frame -- the rows that need
    // BORDER_REFLECT_101 in y; the TMA kernel (k1_trigger_tma.cu) covers [2, rows-2)
    const int y0 = border_only ? ((blockIdx.x / nxg) == 0 ? 0 : rows - th) : (blockIdx.x / nxg) * th;
    const int x0 = ((blockIdx.x % nxg) * (blockDim.x >> 5) + warp) * ABUB_K1_TILE_W;
    if (y0 < rows && x0 < cols)
        k1_strip_any<0>(base + (size_t)j * npix, base + (size_t)jp * npix, k32_ptrs[cam], rows, cols, y0, th,
                        x0, sh_hist, xch, nullptr, -1);
    __syncthreads();
    uint32_t* H = hist + (((size_t)slot * ncams + cam) * max_frames + j) * 256;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        const uint32_t c = sh_hist[i];
        if (c) atomicAdd(&H[i], c);
    }
}

// bin 0 = rows*cols - sum(bins 1..255); one warp per (slot,cam,j)
__global__ void k1_fill_bin0_kernel(uint32_t* __restrict__ hist, const int* __restrict__ nframes, int slot_first,
                                    int nslots, int ncams, int max_frames, uint32_t npix, int j_first, int j_last) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int total = nslots * ncams * max_frames;
    if (w >= total) return;
    const int j = w % max_frames;
    const int sc = w / max_frames;
    const int slot = slot_first + sc / ncams, cam = sc % ncams;
    if (j < j_first || j >= j_last || j >= nframes[slot * ncams + cam]) return;
    uint32_t* H = hist + (((size_t)slot * ncams + cam) * max_frames + j) * 256;
    uint32_t s = 0;
    for (int i = lane; i < 256; i += 32) s += (i == 0) ? 0u : H[i];
    s = __reduce_add_sync(0xffffffffu, s);
    if (lane == 0) H[0] = npix - s;
}

// single frame pair (parity tap + localizer genesis frame)
__global__ void __launch_bounds__(K1_MAX_THREADS, 2) k1_single_kernel(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B,
                                 const uint32_t* __restrict__ K, uint8_t* __restrict__ out,
                                 uint32_t* __restrict__ hist, int rows, int cols, int tozero, int th, int nxg) {
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* sh_hist = smem;
    const int warp = threadIdx.x >> 5;
    uint32_t* xch = smem + 256 + warp * K1_WARP_WORDS;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sh_hist[i] = 0;
    __syncthreads();
    const int y0 = (blockIdx.x / nxg) * th;
    const int x0 = ((blockIdx.x % nxg) * (blockDim.x >> 5) + warp) * ABUB_K1_TILE_W;
    if (y0 < rows && x0 < cols)
        k1_strip_any<1>(A, B, K, rows, cols, y0, th, x0, sh_hist, xch, out, tozero);
    __syncthreads();
    if (hist)
        for (int i = threadIdx.x; i < 256; i += blockDim.x) {
            const uint32_t c = sh_hist[i];
            if (c) atomicAdd(&hist[i], c);
        }
}

// localizer genesis jobs (CalculateInitialBubbleParams, L3Localizer.cpp:232,252): ProcessFrame,
// THRESH_TOZERO at loc_thres, histogram of the result; grid.y = job (non-genesis jobs return)
__global__ void __launch_bounds__(K1_MAX_THREADS, 2) k1_genesis_jobs_kernel(const abub_loc_job* __restrict__ jobs, int rows, int cols, int th, int nxg) {
    extern __shared__ __align__(16) uint32_t smem[];
    const abub_loc_job job = jobs[blockIdx.y];
    if (job.genesis != 1) return;
    uint32_t* sh_hist = smem;
    const int warp = threadIdx.x >> 5;
    uint32_t* xch = smem + 256 + warp * K1_WARP_WORDS;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sh_hist[i] = 0;
    __syncthreads();
    const int y0 = (blockIdx.x / nxg) * th;
    const int x0 = ((blockIdx.x % nxg) * (blockDim.x >> 5) + warp) * ABUB_K1_TILE_W;
    if (y0 < rows && x0 < cols)
        k1_strip_any<1>(job.a, job.b, job.k32, rows, cols, y0, th, x0, sh_hist, xch, job.tz, job.tozero);
    __syncthreads();
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        const uint32_t c = sh_hist[i];
        if (c) atomicAdd(&job.hist[i], c);
    }
}

// the four border rows (0, 1, rows-2, rows-1) of the genesis jobs listed in gen_list, for the TMA genesis kernel
// (k5_genesis_tma.cu) that covers rows [2, rows-2): two strips of two rows with reflected vertical taps, written densely;
// the rows' 240-pixel segments are flagged "holds something / memory not known to be zero" (3).
__global__ void __launch_bounds__(K1_MAX_THREADS, 2) k1_genesis_border_kernel(const abub_loc_job* __restrict__ jobs, const int* __restrict__ gen_list,
                                                                              int rows, int cols, int nxg) {
    extern __shared__ __align__(16) uint32_t smem[];
    const abub_loc_job job = jobs[gen_list[blockIdx.y]];
    uint32_t* sh_hist = smem;
    const int warp = threadIdx.x >> 5;
    uint32_t* xch = smem + 256 + warp * K1_WARP_WORDS;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sh_hist[i] = 0;
    __syncthreads();
    const int y0 = (blockIdx.x / nxg) == 0 ? 0 : rows - 2;
    const int x0 = ((blockIdx.x % nxg) * (blockDim.x >> 5) + warp) * ABUB_K1_TILE_W;
    if (x0 < cols)
        k1_strip_any<1>(job.a, job.b, job.k32, rows, cols, y0, 2, x0, sh_hist, xch, job.tz, job.tozero);
    const int nxw = (cols + 239) / 240;
    if ((blockIdx.x % nxg) == 0 && (int)threadIdx.x < 2 * nxw) job.segflags[(size_t)(y0 + (int)threadIdx.x / nxw) * nxw + (int)threadIdx.x % nxw] = 3;
    __syncthreads();
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        const uint32_t c = sh_hist[i];
        if (c && i) atomicAdd(&job.hist[i], c);
    }
}

__global__ void k1_fill_bin0_single(uint32_t* hist, uint32_t npix) {
    const int lane = threadIdx.x;
    uint32_t s = 0;
    for (int i = lane; i < 256; i += 32) s += (i == 0) ? 0u : hist[i];
    s = __reduce_add_sync(0xffffffffu, s);
    if (lane == 0) hist[0] = npix - s;
}

// K32 = (255-t) | (765-t)<<16, t = min(6*sigma, 255); also t8
__global__ void k1_build_k32_kernel(const uint8_t* __restrict__ sigma, uint32_t* __restrict__ k32,
                                    uint8_t* __restrict__ t8, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t t = min(6u * sigma[i], 255u);
    k32[i] = (255u - t) | ((765u - t) << 16);
    t8[i] = (uint8_t)t;
}

}  // namespace

static inline void k1_strips(int rows, int* nstrips, int* th) {
    *nstrips = (rows + K1_TH_TARGET - 1) / K1_TH_TARGET;
    *th = (rows + *nstrips - 1) / *nstrips;
}
template <typename KernelT>
static int k1_allow_smem(abub_ctx* ctx, KernelT kernel, size_t bytes) {
    if (bytes > 48 * 1024)
        ABUB_CHECK_CUDA(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return ABUB_OK;
}
static inline int k1_block_threads(int cols) {
    const int ntx = (cols + ABUB_K1_TILE_W - 1) / ABUB_K1_TILE_W;
    return 32 * (ntx < K1_MAX_THREADS / 32 ? ntx : K1_MAX_THREADS / 32);
}
static inline int k1_xgroups(int cols) {
    const int ntx = (cols + ABUB_K1_TILE_W - 1) / ABUB_K1_TILE_W, wpc = k1_block_threads(cols) / 32;
    return (ntx + wpc - 1) / wpc;
}
static inline size_t k1_smem_bytes(int threads) { return (256 + (threads / 32) * K1_WARP_WORDS) * sizeof(uint32_t); }

int abub_k1_build_k32(abub_ctx* ctx, int cam) {
    abub_cam_state& c = ctx->cam[cam];
    const size_t n = ctx->npix;
    ABUB_L(ctx), k1_build_k32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(c.d_sigma, c.d_k32, c.d_t8, n);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

// Histograms of the frame pairs j in [j_first, j_last) of every (slot, camera) of the range; a pair is skipped
// when j >= d_limit[slot*ncams+cam] (d_limit = the frame counts, or the number of frames resident so far in the
// streamed form, 0 = this camera is finished).  zero_hist clears the whole table of the range first.
int abub_k1_launch_trigger_range(abub_ctx* ctx, int slot_first, int nslots, int j_first, int j_last, const int* d_limit,
                                 bool zero_hist) {
    const int threads = k1_block_threads(ctx->cols);
    if (zero_hist) {
        const size_t hist_off = (size_t)slot_first * ctx->ncams * ctx->max_frames * 256;
        const size_t hist_n = (size_t)nslots * ctx->ncams * ctx->max_frames * 256;
        ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(ctx->d_hist + hist_off, 0, hist_n * sizeof(uint32_t), ctx->stream));
    }
    if (j_first < 1) j_first = 1;
    if (j_last > ctx->max_frames) j_last = ctx->max_frames;
    if (j_last <= j_first) return ABUB_OK;
    abub_prof_scope prof(ctx, ABUB_PROF_K1);
    int nstrips, th;
    k1_strips(ctx->rows, &nstrips, &th);
    if (int rc = k1_allow_smem(ctx, k1_trigger_kernel, k1_smem_bytes(threads))) return rc;
    // z is limited to 65535: chunk the (slot,cam) range
    const int per = 65535 / ctx->ncams;
    static const bool legacy = getenv("ABUB_K1_LEGACY") != nullptr;     // A/B switch for profiling
    const bool tma = ctx->d_tmaps && !legacy;
    if (tma)
        if (int rc = abub_k1_launch_trigger_tma(ctx, slot_first, nslots, j_first, j_last, d_limit)) return rc;
    for (int s0 = 0; s0 < nslots; s0 += per) {
        const int ns = (nslots - s0 < per) ? nslots - s0 : per;
        // with the TMA kernel on rows [2, rows-2) this kernel only does the 2+2 border rows
        dim3 grid((tma ? 2 : nstrips) * k1_xgroups(ctx->cols), j_last - j_first, ns * ctx->ncams);
        ABUB_L(ctx), k1_trigger_kernel<<<grid, threads, k1_smem_bytes(threads), ctx->stream>>>(
            ctx->d_slot_ptrs, d_limit, ctx->d_tss, ctx->d_k32_ptrs, ctx->d_hist, slot_first + s0,
            ctx->rows, ctx->cols, ctx->ncams, ctx->max_frames, tma ? 2 : th, k1_xgroups(ctx->cols), tma ? 1 : 0, j_first);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    const int total_w = nslots * ctx->ncams * ctx->max_frames;
    ABUB_L(ctx), k1_fill_bin0_kernel<<<(total_w * 32 + 255) / 256, 256, 0, ctx->stream>>>(
        ctx->d_hist, d_limit, slot_first, nslots, ctx->ncams, ctx->max_frames, (uint32_t)ctx->npix, j_first, j_last);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

int abub_k1_launch_trigger(abub_ctx* ctx, int slot_first, int nslots) {
    return abub_k1_launch_trigger_range(ctx, slot_first, nslots, 1, ctx->max_frames, ctx->d_nframes, true);
}

int abub_k1_launch_single(abub_ctx* ctx, const uint8_t* d_a, const uint8_t* d_b, const uint32_t* d_k32,
                          uint8_t* d_out, uint32_t* d_hist, int tozero_thres) {
    const int threads = k1_block_threads(ctx->cols);
    if (d_hist) ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(d_hist, 0, 256 * sizeof(uint32_t), ctx->stream));
    int nstrips, th;
    k1_strips(ctx->rows, &nstrips, &th);
    if (int rc = k1_allow_smem(ctx, k1_single_kernel, k1_smem_bytes(threads))) return rc;
    const int nxg = k1_xgroups(ctx->cols);
    ABUB_L(ctx), k1_single_kernel<<<nstrips * nxg, threads, k1_smem_bytes(threads), ctx->stream>>>(d_a, d_b, d_k32, d_out, d_hist, ctx->rows,
                                                                                       ctx->cols, tozero_thres, th, nxg);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    if (d_hist) {
        ABUB_L(ctx), k1_fill_bin0_single<<<1, 32, 0, ctx->stream>>>(d_hist, (uint32_t)ctx->npix);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    return ABUB_OK;
}

int abub_k1_launch_genesis_border(abub_ctx* ctx, const abub_loc_job* d_jobs, const int* d_gen_list, int ngen) {
    const int threads = k1_block_threads(ctx->cols);
    if (int rc = k1_allow_smem(ctx, k1_genesis_border_kernel, k1_smem_bytes(threads))) return rc;
    const int nxg = k1_xgroups(ctx->cols);
    for (int j0 = 0; j0 < ngen; j0 += 65535) {
        const int nj = (ngen - j0 < 65535) ? ngen - j0 : 65535;
        ABUB_L(ctx), k1_genesis_border_kernel<<<dim3(2 * nxg, nj), threads, k1_smem_bytes(threads), ctx->stream>>>(
            d_jobs, d_gen_list + j0, ctx->rows, ctx->cols, nxg);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    return ABUB_OK;
}

int abub_k1_launch_genesis_jobs(abub_ctx* ctx, const abub_loc_job* d_jobs, int njobs) {
    abub_prof_scope prof(ctx, ABUB_PROF_K5A);
    // cols % 16 == 0: the TMA-fed kernel (k5_genesis_tma.cu) on rows [2, rows-2), sparse store; the four border rows here.
    // A/B switch for profiling: ABUB_K5A_LEGACY
    static const bool legacy = getenv("ABUB_K5A_LEGACY") != nullptr;
    if (!legacy && abub_k5a_tma_supported(ctx)) {
        int ngen = 0;
        if (int rc = abub_k5a_launch_tma(ctx, d_jobs, njobs, &ngen)) return rc;
        if (ngen == 0) return ABUB_OK;
        return abub_k1_launch_genesis_border(ctx, d_jobs, ctx->d_gen_list, ngen);
    }
    const int threads = k1_block_threads(ctx->cols);
    int nstrips, th;
    k1_strips(ctx->rows, &nstrips, &th);
    if (int rc = k1_allow_smem(ctx, k1_genesis_jobs_kernel, k1_smem_bytes(threads))) return rc;
    for (int j0 = 0; j0 < njobs; j0 += 65535) {
        const int nj = (njobs - j0 < 65535) ? njobs - j0 : 65535;
        const int nxg = k1_xgroups(ctx->cols);
        ABUB_L(ctx), k1_genesis_jobs_kernel<<<dim3(nstrips * nxg, nj), threads, k1_smem_bytes(threads), ctx->stream>>>(
            d_jobs + j0, ctx->rows, ctx->cols, th, nxg);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    return ABUB_OK;
}
