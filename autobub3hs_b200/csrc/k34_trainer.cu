// K3 / K4 -- Trainer background statistics.
// Reference: Trainer::calculateEntropyFrame on frame1-frame0 (Trainer.cpp:279-280,341-376)
// and Trainer::CalculateMeanSigmaImageVector (Trainer.cpp:144-216).
//
// K3: 16-bin histogram of the saturating difference frame1 - frame0 (bin = value >> 4).
//     Quiet pairs put every pixel in bin 0, so only bytes >= 16 touch shared-memory atomics
//     and bin 0 is recovered by subtraction.  The entropy itself (16 float terms) is host work.
// K4: per pixel, the exact sequential float32 Welford recurrence over the kept frames in
//     push order (frame index innermost, as Trainer.cpp:180-189) -- each thread owns 4
//     pixels and streams one 32-bit word per frame, so a frame byte is read exactly once
//     and the only writes are the two u8 result planes.  _rn intrinsics forbid FMA
//     contraction (the reference is compiled at -O0).
// K4p: exact integer partials (u32 sum x, u64 sum x^2) for the multi-GPU all-reduce, and
//     the fp64 finish from reduced partials with boundary flagging.
#include <stdlib.h>

#include "abub_internal.cuh"

namespace {

__global__ void k3_gate_hist16_kernel(const uint8_t* __restrict__ f0, const uint8_t* __restrict__ f1,
                                      uint32_t* __restrict__ hist16, size_t npix) {
    __shared__ uint32_t sh[16];
    if (threadIdx.x < 16) sh[threadIdx.x] = 0;
    __syncthreads();
    const bool aligned = (((uintptr_t)f0 | (uintptr_t)f1) & 15) == 0;
    const size_t nvec = aligned ? npix / 16 : 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(f1) + i);
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(f0) + i);
        const uint32_t d[4] = {__vsubus4(a.x, b.x), __vsubus4(a.y, b.y), __vsubus4(a.z, b.z), __vsubus4(a.w, b.w)};
        if (((d[0] | d[1] | d[2] | d[3]) & 0xF0F0F0F0u) != 0) {
#pragma unroll
            for (int w = 0; w < 4; w++)
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const uint32_t bin = (d[w] >> (8 * c + 4)) & 0xFu;
                    if (bin) atomicAdd(&sh[bin], 1u);
                }
        }
    }
    // tail (npix not a multiple of 16, or unaligned planes)
    for (size_t p = nvec * 16 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += stride) {
        const int v = (int)f1[p] - (int)f0[p];
        const uint32_t bin = (uint32_t)(v > 0 ? v : 0) >> 4;
        if (bin) atomicAdd(&sh[bin], 1u);
    }
    __syncthreads();
    if (threadIdx.x > 0 && threadIdx.x < 16 && sh[threadIdx.x]) atomicAdd(&hist16[threadIdx.x], sh[threadIdx.x]);
}

__global__ void k3_fill_bin0(uint32_t* hist16, uint32_t npix) {
    uint32_t s = 0;
    for (int i = 1; i < 16; i++) s += hist16[i];
    hist16[0] = npix - s;
}

// Batched gate: blockIdx.y = event, pairs[2e] = frame 0, pairs[2e+1] = frame 1, hist16 + 16e its histogram.
__global__ void k3_gate_batch_kernel(const uint8_t* const* __restrict__ pairs, uint32_t* __restrict__ hist16, size_t npix) {
    __shared__ uint32_t sh[16];
    const uint8_t* __restrict__ f0 = pairs[2 * blockIdx.y];
    const uint8_t* __restrict__ f1 = pairs[2 * blockIdx.y + 1];
    if (threadIdx.x < 16) sh[threadIdx.x] = 0;
    __syncthreads();
    const bool aligned = (((uintptr_t)f0 | (uintptr_t)f1) & 15) == 0;
    const size_t nvec = aligned ? npix / 16 : 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(f1) + i);
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(f0) + i);
        // bin = sat(f1-f0) >> 4 is non-zero iff f1 >= f0 + 16: absdiff keeps that visible in the high nibbles
        // whenever it can be true, so the (5-instruction) saturating subtract runs only on words that need it
        const uint32_t q[4] = {__vabsdiffu4(a.x, b.x), __vabsdiffu4(a.y, b.y), __vabsdiffu4(a.z, b.z), __vabsdiffu4(a.w, b.w)};
        if (((q[0] | q[1] | q[2] | q[3]) & 0xF0F0F0F0u) != 0) {
            const uint32_t d[4] = {__vsubus4(a.x, b.x), __vsubus4(a.y, b.y), __vsubus4(a.z, b.z), __vsubus4(a.w, b.w)};
#pragma unroll
            for (int w = 0; w < 4; w++)
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const uint32_t bin = (d[w] >> (8 * c + 4)) & 0xFu;
                    if (bin) atomicAdd(&sh[bin], 1u);
                }
        }
    }
    for (size_t p = nvec * 16 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += stride) {
        const int v = (int)f1[p] - (int)f0[p];
        const uint32_t bin = (uint32_t)(v > 0 ? v : 0) >> 4;
        if (bin) atomicAdd(&sh[bin], 1u);
    }
    __syncthreads();
    if (threadIdx.x > 0 && threadIdx.x < 16 && sh[threadIdx.x]) atomicAdd(&hist16[16 * blockIdx.y + threadIdx.x], sh[threadIdx.x]);
}

__global__ void k3_fill_bin0_batch(uint32_t* hist16, int n, uint32_t npix) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    uint32_t s = 0;
    for (int i = 1; i < 16; i++) s += hist16[16 * e + i];
    hist16[16 * e] = npix - s;
}

__device__ __forceinline__ void welford_step(float x, float fdiv, float& mean, float& m2) {
    const float delta = __fsub_rn(x, mean);                          // Trainer.cpp:185
    mean = __fadd_rn(mean, __fdiv_rn(delta, fdiv));                  // :186
    m2 = __fadd_rn(m2, __fmul_rn(delta, __fsub_rn(x, mean)));        // :187
}

__device__ __forceinline__ void welford_finish(float mean, float m2, float nm1, uint8_t& mu8, uint8_t& su8) {
    const float var = __fdiv_rn(m2, nm1);                            // :191  (N==1 -> 0/0 = NaN)
    const float sg = __fsqrt_rn(var);                                // :192
    // (int)NaN is INT_MIN on x86 -> low byte 0; CUDA float->int of NaN is 0 -> same byte
    su8 = (uint8_t)(int)sg;                                          // :195
    mu8 = (uint8_t)mean;                                             // :196 (truncation)
}

// frames: device table of n frame pointers (push order).  4 pixels per thread.
__global__ void k4_welford_kernel(const uint8_t* const* __restrict__ frames, int n, uint8_t* __restrict__ mean_out,
                                  uint8_t* __restrict__ sigma_out, size_t npix) {
    const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t p = g * 4;
    if (p >= npix) return;
    float mean[4] = {0.f, 0.f, 0.f, 0.f}, m2[4] = {0.f, 0.f, 0.f, 0.f};
    const bool full = (p + 4 <= npix) && ((npix & 3) == 0);
    int f = 0;
    if (full) {
        for (; f + 4 <= n; f += 4) {                                 // 4 independent loads in flight
            uint32_t w[4];
#pragma unroll
            for (int u = 0; u < 4; u++) w[u] = __ldg(reinterpret_cast<const uint32_t*>(frames[f + u] + p));
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float fd = (float)(f + u + 1);
#pragma unroll
                for (int c = 0; c < 4; c++) welford_step((float)((w[u] >> (8 * c)) & 0xFFu), fd, mean[c], m2[c]);
            }
        }
    }
    for (; f < n; f++) {
        const float fd = (float)(f + 1);
#pragma unroll
        for (int c = 0; c < 4; c++)
            if (p + c < npix) welford_step((float)frames[f][p + c], fd, mean[c], m2[c]);
    }
    const float nm1 = (float)(size_t)(n - 1);                        // size_t(N-1) -> float, :191
#pragma unroll
    for (int c = 0; c < 4; c++)
        if (p + c < npix) welford_finish(mean[c], m2[c], nm1, mean_out[p + c], sigma_out[p + c]);
}

// exact integer partials of this rank's kept frames
__global__ void k4_partials_kernel(const uint8_t* const* __restrict__ frames, int n, uint32_t* __restrict__ sum,
                                   uint64_t* __restrict__ sumsq, size_t npix) {
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    uint32_t s = 0;
    uint64_t q = 0;
    for (int f = 0; f < n; f++) {
        const uint32_t x = frames[f][p];
        s += x;
        q += (uint64_t)(x * x);
    }
    sum[p] = s;
    sumsq[p] = q;
}

// the same, 16 pixels per thread (one 16-byte load per frame, four frames in flight) with 32-bit sums of squares:
// 255^2 * n < 2^32 for n <= 66051 frames; the launcher uses this kernel up to 60000 frames and 16-byte aligned planes
__global__ void __launch_bounds__(128) k4_partials16_kernel(const uint8_t* const* __restrict__ frames, int n, uint32_t* __restrict__ sum,
                                                            uint64_t* __restrict__ sumsq, size_t nvec) {
    const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nvec) return;
    uint32_t s[16], q[16];
#pragma unroll
    for (int i = 0; i < 16; i++) s[i] = q[i] = 0;
    auto add = [&](const uint4 v) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; k++)
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const uint32_t x = (w[k] >> (8 * c)) & 0xFFu;
                s[4 * k + c] += x;
                q[4 * k + c] += x * x;
            }
    };
    int f = 0;
    for (; f + 4 <= n; f += 4) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) v[u] = __ldg(reinterpret_cast<const uint4*>(frames[f + u]) + g);
#pragma unroll
        for (int u = 0; u < 4; u++) add(v[u]);
    }
    for (; f < n; f++) add(__ldg(reinterpret_cast<const uint4*>(frames[f]) + g));
    uint4* so = reinterpret_cast<uint4*>(sum) + 4 * g;
#pragma unroll
    for (int k = 0; k < 4; k++) so[k] = make_uint4(s[4 * k], s[4 * k + 1], s[4 * k + 2], s[4 * k + 3]);
    ulonglong2* qo = reinterpret_cast<ulonglong2*>(sumsq) + 8 * g;
#pragma unroll
    for (int k = 0; k < 8; k++) qo[k] = make_ulonglong2((unsigned long long)q[2 * k], (unsigned long long)q[2 * k + 1]);
}

// mean/sigma from reduced integer partials in fp64; flags every pixel whose truncated planes COULD differ from the
// sequential float32 recurrence of K4 (Trainer.cpp:185-196).  The flag test is a worst-case bound of that recurrence's
// rounding error, u = 2^-24, samples in [0, 255], k = 2 .. n (step 1 is exact: mean = x, m2 = 0):
//   mean:  e_k = e_{k-1} (k-1)/k + l_k with |l_k| <= u' (2 |delta| / k + |mean_k|) <= 256 u' (2 / k + 1)
//          => n |e_n| <= 256 u' sum_k (2 + k)   =>   |e_n| <= 256 u' (n / 2 + 5 / 2) =: e_mean   (increasing in n, so it
//          also bounds every intermediate e_k)
//   m2:    the exact terms P_k = delta_k (x_k - mean_k) = delta_k^2 (k-1)/k are >= 0 and sum to M2.  Computed terms
//          differ by the three roundings (<= 3 u' P_k) and by the mean errors inside both factors
//          (<= e_mean (|delta_k| + |x_k - mean_k|) + e_mean^2 <= 2 e_mean |delta_k| + e_mean^2); the n-1 additions
//          each lose <= u' M2.  With sum |delta_k| <= sqrt(n sum delta_k^2) <= sqrt(2 n M2):
//          |dM2| <= u' M2 (n + 3) + 2 e_mean sqrt(2 n M2) + n e_mean^2 =: bM,  bv = bM / (n-1)
//   sigma: sqrt(var + bv) - sqrt(var) <= min(bv / (2 sigma), sqrt(bv))   (computed too large: matters just BELOW an integer)
//          sqrt(var) - sqrt(var - bv) <= min(bv / sigma, sqrt(bv))       (computed too small: matters just ABOVE an integer)
//          plus the division's and the square root's own roundings (<= 3 u' sigma).
// u' = 1.01 u absorbs the second-order terms.  A flagged pixel is recomputed with the float32 recurrence itself
// (k4_fixup_kernel), so a generous bound costs time, never exactness; an unflagged pixel provably needs no repair.
// tests/test_gpu_train_localize.py::test_partials_flag_bound_covers_every_float32_difference attacks it (sparse 255
// spikes, two-level and drifting data, n up to 2000).
__global__ void k4_finish_partials_kernel(const uint32_t* __restrict__ sum, const uint64_t* __restrict__ sumsq, int n,
                                          uint8_t* __restrict__ mean_out, uint8_t* __restrict__ sigma_out,
                                          int* __restrict__ nflag, uint8_t* __restrict__ flag, size_t npix) {
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    const double s = (double)sum[p], q = (double)sumsq[p];
    const double mean = s / n;
    double M2 = (n > 1) ? (q - s * s / n) : 0.0;
    if (M2 < 0) M2 = 0;
    const double var = (n > 1) ? M2 / (n - 1) : 0.0;
    const double sg = sqrt(var);
    const double u = 1.01 * 5.9604644775390625e-08;
    const double e_mean = 256.0 * u * (0.5 * (double)n + 2.5);
    const double bM = u * M2 * ((double)n + 3.0) + 2.0 * e_mean * sqrt(2.0 * (double)n * M2) + (double)n * e_mean * e_mean;
    const double bv = bM / (double)(n > 1 ? n - 1 : 1) + 1e-9;       // + the fp64 cancellation in M2 (q, s*s/n < 2^38)
    const double rb = sqrt(bv), tail = 3.0 * u * sg + 1e-12;
    const double e_up = (sg > 0 ? fmin(bv / (2.0 * sg), rb) : rb) + tail;
    const double e_dn = (sg > 0 ? fmin(bv / sg, rb) : rb) + tail;
    const double fm = mean - floor(mean), fs = sg - floor(sg);
    // all samples equal (M2 == 0): the recurrence is exact (delta = 0 from step 2 on), nothing to repair
    const bool risky_m = M2 > 0 && (fm < e_mean || 1.0 - fm < e_mean);
    const bool risky_s = M2 > 0 && (fs < e_dn || 1.0 - fs < e_up);
    const bool risky = risky_m || risky_s || n < 2;
    mean_out[p] = (uint8_t)mean;
    sigma_out[p] = (n > 1) ? (uint8_t)(int)sg : 0;
    flag[p] = risky ? 1 : 0;
    if (risky) atomicAdd(nflag, 1);
}

// sequential float32 recurrence restricted to flagged pixels (fix-up after the all-gather
// of their samples is replaced, on one node, by reading the full frame table)
__global__ void k4_fixup_kernel(const uint8_t* const* __restrict__ frames, int n, const uint8_t* __restrict__ flag,
                                uint8_t* __restrict__ mean_out, uint8_t* __restrict__ sigma_out, size_t npix) {
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix || !flag[p]) return;
    float mean = 0.f, m2 = 0.f;
    for (int f = 0; f < n; f++) welford_step((float)frames[f][p], (float)(f + 1), mean, m2);
    welford_finish(mean, m2, (float)(size_t)(n - 1), mean_out[p], sigma_out[p]);
}

// ---- K4 fast path ---------------------------------------------------------------------------------
// The generic kernel above spends ~28 instructions per sample: __fdiv_rn expands to MUFU.RCP + FCHK + a
// refinement chain + a slow-path call per sample, and the u8 -> f32 conversions are quarter-rate I2F.  Both
// are avoided here without changing one bit of the result:
//   * the divisor (frame index + 1) is the same for every pixel, so its correctly rounded reciprocal
//     r = RN(1/k) is formed once per frame per thread and the quotient is
//         q0 = RN(a*r);  e = RN(a - q0*k) (one FMA, exact);  q = RN(q0 + e*r)
//     -- Markstein's correction step.  One step returns RN(a/k) whenever q0 is within one ulp of a/k; that it
//     does so for every value a Welford delta can take (+0 and 2^-50 <= |a| <= 256, see k4_fdiv_check_kernel)
//     and every k <= ABUB_FDIV1_MAX is checked exhaustively against __fdiv_rn on the device by abub_selftest_fdiv
//     (tests/test_gpu_train_localize.py).
//     Above that bound a second step is applied (q0 is then faithful, for which the step is proven exact).
//   * byte c of a word becomes a float by PRMT into the mantissa of 2^23 and one exact FADD of -2^23.
// Each thread owns 16 pixels (one 16-byte load per frame, 16 independent recurrences), 4 frames in flight.
#define ABUB_FDIV1_MAX 4096

template <int STEPS>
__device__ __forceinline__ float div_by_count(float a, float k, float r) {
    float q = __fmul_rn(a, r);
    float e = __fmaf_rn(-q, k, a);
    q = __fmaf_rn(e, r, q);
    if (STEPS == 2) {
        e = __fmaf_rn(-q, k, a);
        q = __fmaf_rn(e, r, q);
    }
    return q;
}

__device__ __forceinline__ float byte_to_float(uint32_t w, int c) {
    return __fadd_rn(__uint_as_float(__byte_perm(w, 0x4B000000u, 0x7540 + c)), -8388608.0f);
}

template <int STEPS>
__device__ __forceinline__ void welford_word(uint32_t w, float k, float r, float* mean, float* m2) {
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const float x = byte_to_float(w, c);
        const float delta = __fsub_rn(x, mean[c]);                                   // Trainer.cpp:185
        mean[c] = __fadd_rn(mean[c], div_by_count<STEPS>(delta, k, r));              // :186
        m2[c] = __fadd_rn(m2[c], __fmul_rn(delta, __fsub_rn(x, mean[c])));           // :187
    }
}

template <int STEPS>
__global__ void __launch_bounds__(64) k4_welford16_kernel(const uint8_t* const* __restrict__ frames, int n,
                                                          uint8_t* __restrict__ mean_out, uint8_t* __restrict__ sigma_out,
                                                          size_t nvec) {
    const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nvec) return;
    float mean[16], m2[16];
#pragma unroll
    for (int i = 0; i < 16; i++) mean[i] = m2[i] = 0.f;
    int f = 0;
    // groups of 4 frames, the next group's loads issued before the current group's arithmetic
    uint4 w[4], nx[4];
    if (n >= 4) {
#pragma unroll
        for (int u = 0; u < 4; u++) w[u] = __ldg(reinterpret_cast<const uint4*>(frames[u]) + g);
    }
#pragma unroll 1
    for (; f + 4 <= n; f += 4) {
        if (f + 8 <= n) {
#pragma unroll
            for (int u = 0; u < 4; u++) nx[u] = __ldg(reinterpret_cast<const uint4*>(frames[f + 4 + u]) + g);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const float k = (float)(f + u + 1);
            const float r = __frcp_rn(k);
            welford_word<STEPS>(w[u].x, k, r, mean + 0, m2 + 0);
            welford_word<STEPS>(w[u].y, k, r, mean + 4, m2 + 4);
            welford_word<STEPS>(w[u].z, k, r, mean + 8, m2 + 8);
            welford_word<STEPS>(w[u].w, k, r, mean + 12, m2 + 12);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) w[u] = nx[u];
    }
    for (; f < n; f++) {
        const uint4 w = __ldg(reinterpret_cast<const uint4*>(frames[f]) + g);
        const float k = (float)(f + 1);
        const float r = __frcp_rn(k);
        welford_word<STEPS>(w.x, k, r, mean + 0, m2 + 0);
        welford_word<STEPS>(w.y, k, r, mean + 4, m2 + 4);
        welford_word<STEPS>(w.z, k, r, mean + 8, m2 + 8);
        welford_word<STEPS>(w.w, k, r, mean + 12, m2 + 12);
    }
    const float nm1 = (float)(size_t)(n - 1);                                       // :191
    uint32_t mo[4], so[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        mo[q] = so[q] = 0;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            uint8_t mu, su;
            welford_finish(mean[4 * q + c], m2[4 * q + c], nm1, mu, su);
            mo[q] |= (uint32_t)mu << (8 * c);
            so[q] |= (uint32_t)su << (8 * c);
        }
    }
    reinterpret_cast<uint4*>(mean_out)[g] = make_uint4(mo[0], mo[1], mo[2], mo[3]);
    reinterpret_cast<uint4*>(sigma_out)[g] = make_uint4(so[0], so[1], so[2], so[3]);
}

// Exhaustive check of div_by_count<STEPS> against __fdiv_rn over the whole set a Welford delta can come from:
// a = +0 and every float with 2^-50 <= |a| <= 256, both signs (blockIdx.y picks the divisor).
// Why nothing smaller can occur (x = sample, an integer 0..255; m = running mean >= 0; delta = RN(x - m)):
//   * m becomes non-zero only from m < 1/2 with x >= 1, where delta >= 1/2 and the new m >= (1/2k)(1 - 2^-23);
//     afterwards a step shrinks it by at most (1 - 1/k)(1 - 2^-23) (x = 0), so after K <= 2^24 frames a non-zero m
//     is still above 2^-25 e^-2 > 2^-28 -- except k = 1 with x = 0, which gives m = 0 exactly;
//   * hence for x = 0, |delta| = m is 0 or > 2^-28; for x >= 1, |delta| >= 1/2 when m < 1/2, and otherwise x - m is a
//     multiple of ulp(1/2)/2 = 2^-25: 0 or >= 2^-25.  RN(x - x) is +0, never -0.
// 2^-50 leaves 22 binades of margin; quotients and residuals of this set stay far from the subnormal range, which
// is where the shortcut (like any multiply-by-reciprocal scheme) would stop being exact.
// mismatches[0] counts, mismatches[1..2] keep one failing (a bits, k).
template <int STEPS>
__global__ void k4_fdiv_check_kernel(int k_first, unsigned long long* __restrict__ mismatches) {
    const float k = (float)(k_first + (int)blockIdx.y);
    const float r = __frcp_rn(k);
    const uint32_t lo = 0x26800000u, hi = 0x43800000u;                              // bits of 2^-50 and of 256
    unsigned bad = 0;
    uint32_t bad_bits = 0;
    for (uint64_t i = (uint64_t)lo + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= hi; i += (uint64_t)gridDim.x * blockDim.x) {
        const float a = __uint_as_float((uint32_t)i);
        const float want = __fdiv_rn(a, k), got = div_by_count<STEPS>(a, k, r);
        const float wantn = __fdiv_rn(-a, k), gotn = div_by_count<STEPS>(-a, k, r);
        if (__float_as_uint(want) != __float_as_uint(got) || __float_as_uint(wantn) != __float_as_uint(gotn)) {
            bad++;
            bad_bits = (uint32_t)i;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && __float_as_uint(div_by_count<STEPS>(0.0f, k, r)) != 0u) {
        bad++;
        bad_bits = 0;
    }
    if (bad) {
        atomicAdd(&mismatches[0], (unsigned long long)bad);
        mismatches[1] = bad_bits;
        mismatches[2] = (unsigned long long)(k_first + (int)blockIdx.y);
    }
}

}  // namespace

int abub_k3_gate_hist16(abub_ctx* ctx, const uint8_t* d_f0, const uint8_t* d_f1, uint32_t* d_hist16) {
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(d_hist16, 0, 16 * sizeof(uint32_t), ctx->stream));
    ABUB_L(ctx), k3_gate_hist16_kernel<<<148 * 4, 256, 0, ctx->stream>>>(d_f0, d_f1, d_hist16, ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    ABUB_L(ctx), k3_fill_bin0<<<1, 1, 0, ctx->stream>>>(d_hist16, (uint32_t)ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

int abub_k3_gate_batch(abub_ctx* ctx, const uint8_t* const* d_pairs, int n, uint32_t* d_hist16) {
    abub_prof_scope prof(ctx, ABUB_PROF_K3);
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(d_hist16, 0, (size_t)n * 16 * sizeof(uint32_t), ctx->stream));
    // 32 CTAs x 256 threads x 16 B per sweep of one pair; the event index rides on grid.y
    for (int e0 = 0; e0 < n; e0 += 65535) {
        const int ne = n - e0 < 65535 ? n - e0 : 65535;
        ABUB_L(ctx), k3_gate_batch_kernel<<<dim3(32, ne), 256, 0, ctx->stream>>>(d_pairs + 2 * (size_t)e0, d_hist16 + 16 * (size_t)e0, ctx->npix);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    ABUB_L(ctx), k3_fill_bin0_batch<<<(n + 127) / 128, 128, 0, ctx->stream>>>(d_hist16, n, (uint32_t)ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

// d_frame_table holds frames parked by train.cu (cudaMalloc'd chunks, frame stride npix): 16-byte aligned whenever
// npix is a multiple of 16
int abub_k4_welford(abub_ctx* ctx, const uint8_t* const* d_frame_table, int n, uint8_t* d_mean, uint8_t* d_sigma) {
    abub_prof_scope prof(ctx, ABUB_PROF_K4);
    const bool fast = n <= (1 << 24) && ctx->npix % 16 == 0 && (((uintptr_t)d_mean | (uintptr_t)d_sigma) & 15) == 0 &&
                      !getenv("ABUB_K4_GENERIC");
    if (fast) {
        const size_t nvec = ctx->npix / 16;
        const unsigned nb = (unsigned)((nvec + 63) / 64);
        if (n <= ABUB_FDIV1_MAX)
            ABUB_L(ctx), k4_welford16_kernel<1><<<nb, 64, 0, ctx->stream>>>(d_frame_table, n, d_mean, d_sigma, nvec);
        else
            ABUB_L(ctx), k4_welford16_kernel<2><<<nb, 64, 0, ctx->stream>>>(d_frame_table, n, d_mean, d_sigma, nvec);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
        return ABUB_OK;
    }
    const size_t nthreads = (ctx->npix + 3) / 4;
    ABUB_L(ctx), k4_welford_kernel<<<(unsigned)((nthreads + 127) / 128), 128, 0, ctx->stream>>>(d_frame_table, n, d_mean, d_sigma,
                                                                                  ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

int abub_k4_partials(abub_ctx* ctx, const uint8_t* const* d_frame_table, int n, uint32_t* d_sum, uint64_t* d_sumsq, bool aligned16) {
    abub_prof_scope prof(ctx, ABUB_PROF_K4);
    if (aligned16 && n <= 60000 && ctx->npix % 16 == 0 && !getenv("ABUB_K4_GENERIC")) {
        const size_t nvec = ctx->npix / 16;
        ABUB_L(ctx), k4_partials16_kernel<<<(unsigned)((nvec + 127) / 128), 128, 0, ctx->stream>>>(d_frame_table, n, d_sum, d_sumsq, nvec);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
        return ABUB_OK;
    }
    ABUB_L(ctx), k4_partials_kernel<<<(unsigned)((ctx->npix + 255) / 256), 256, 0, ctx->stream>>>(d_frame_table, n, d_sum, d_sumsq,
                                                                                     ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

int abub_k4_finish_from_partials(abub_ctx* ctx, const uint32_t* d_sum, const uint64_t* d_sumsq, int n, uint8_t* d_mean,
                                 uint8_t* d_sigma, int* d_nflag, uint8_t* d_flag) {
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(d_nflag, 0, sizeof(int), ctx->stream));
    ABUB_L(ctx), k4_finish_partials_kernel<<<(unsigned)((ctx->npix + 255) / 256), 256, 0, ctx->stream>>>(d_sum, d_sumsq, n, d_mean,
                                                                                            d_sigma, d_nflag, d_flag,
                                                                                            ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

int abub_k4_fixup(abub_ctx* ctx, const uint8_t* const* d_frame_table, int n, const uint8_t* d_flag, uint8_t* d_mean,
                  uint8_t* d_sigma) {
    ABUB_L(ctx), k4_fixup_kernel<<<(unsigned)((ctx->npix + 255) / 256), 256, 0, ctx->stream>>>(d_frame_table, n, d_flag, d_mean,
                                                                                  d_sigma, ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

// Self-test of the K4 division shortcut: compares div_by_count<steps> with __fdiv_rn for a = +0 and every float a,
// 2^-50 <= |a| <= 256, and every divisor in [k_first, k_last].  n_mismatch = 0 is the pass condition.
extern "C" int abub_selftest_fdiv(abub_ctx* ctx, int steps, int k_first, int k_last, unsigned long long* n_mismatch,
                                  unsigned long long* example) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, (steps == 1 || steps == 2) && k_first >= 1 && k_last >= k_first && k_last - k_first < 65535 && n_mismatch,
                 "abub_selftest_fdiv");
    cudaSetDevice(ctx->device);
    unsigned long long* d = nullptr;
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&d, 3 * sizeof(unsigned long long)));
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(d, 0, 3 * sizeof(unsigned long long), ctx->stream));
    const dim3 grid(148 * 2, (unsigned)(k_last - k_first + 1));
    if (steps == 1) ABUB_L(ctx), k4_fdiv_check_kernel<1><<<grid, 256, 0, ctx->stream>>>(k_first, d);
    else ABUB_L(ctx), k4_fdiv_check_kernel<2><<<grid, 256, 0, ctx->stream>>>(k_first, d);
    unsigned long long h[3] = {0, 0, 0};
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(h, d, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    ABUB_CHECK_CUDA(ctx, e);
    *n_mismatch = h[0];
    if (example) { example[0] = h[1]; example[1] = h[2]; }
    return ABUB_OK;
}

int abub_k4_fdiv1_max(void) { return ABUB_FDIV1_MAX; }
