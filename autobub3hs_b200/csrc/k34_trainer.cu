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

// mean/sigma from reduced integer partials in fp64; flags pixels whose truncation could
// differ from the sequential float32 recurrence (value within eps of an integer boundary).
__global__ void k4_finish_partials_kernel(const uint32_t* __restrict__ sum, const uint64_t* __restrict__ sumsq, int n,
                                          uint8_t* __restrict__ mean_out, uint8_t* __restrict__ sigma_out,
                                          int* __restrict__ nflag, uint8_t* __restrict__ flag, size_t npix) {
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    const double s = (double)sum[p], q = (double)sumsq[p];
    const double mean = s / n;
    double var = (n > 1) ? (q - s * s / n) / (n - 1) : 0.0;
    if (var < 0) var = 0;
    const double sg = sqrt(var);
    const double eps = (double)n * 1.1920928955078125e-07 * 4.0 * (1.0 + mean);
    const double fm = mean - floor(mean), fs = sg - floor(sg);
    const bool risky = (fm < eps) || (1.0 - fm < eps) || (fs < eps) || (1.0 - fs < eps) || n < 2;
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

}  // namespace

int abub_k3_gate_hist16(abub_ctx* ctx, const uint8_t* d_f0, const uint8_t* d_f1, uint32_t* d_hist16) {
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(d_hist16, 0, 16 * sizeof(uint32_t), ctx->stream));
    k3_gate_hist16_kernel<<<148 * 4, 256, 0, ctx->stream>>>(d_f0, d_f1, d_hist16, ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    k3_fill_bin0<<<1, 1, 0, ctx->stream>>>(d_hist16, (uint32_t)ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

int abub_k4_welford(abub_ctx* ctx, const uint8_t* const* d_frame_table, int n, uint8_t* d_mean, uint8_t* d_sigma) {
    const size_t nthreads = (ctx->npix + 3) / 4;
    k4_welford_kernel<<<(unsigned)((nthreads + 127) / 128), 128, 0, ctx->stream>>>(d_frame_table, n, d_mean, d_sigma,
                                                                                  ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

int abub_k4_partials(abub_ctx* ctx, const uint8_t* const* d_frame_table, int n, uint32_t* d_sum, uint64_t* d_sumsq) {
    k4_partials_kernel<<<(unsigned)((ctx->npix + 255) / 256), 256, 0, ctx->stream>>>(d_frame_table, n, d_sum, d_sumsq,
                                                                                     ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

int abub_k4_finish_from_partials(abub_ctx* ctx, const uint32_t* d_sum, const uint64_t* d_sumsq, int n, uint8_t* d_mean,
                                 uint8_t* d_sigma, int* d_nflag, uint8_t* d_flag) {
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(d_nflag, 0, sizeof(int), ctx->stream));
    k4_finish_partials_kernel<<<(unsigned)((ctx->npix + 255) / 256), 256, 0, ctx->stream>>>(d_sum, d_sumsq, n, d_mean,
                                                                                            d_sigma, d_nflag, d_flag,
                                                                                            ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}

int abub_k4_fixup(abub_ctx* ctx, const uint8_t* const* d_frame_table, int n, const uint8_t* d_flag, uint8_t* d_mean,
                  uint8_t* d_sigma) {
    k4_fixup_kernel<<<(unsigned)((ctx->npix + 255) / 256), 256, 0, ctx->stream>>>(d_frame_table, n, d_flag, d_mean,
                                                                                  d_sigma, ctx->npix);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    return ABUB_OK;
}
