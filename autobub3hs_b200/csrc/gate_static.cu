// abub_gate_entropy_static -- the Trainer entropy gate (K3) for callers that must not allocate: code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
// The arithmetic is the gate of Trainer.cpp:279-287 / calculateEntropyFrame
// (Trainer.cpp:341-376), the same as k3_gate_batch_kernel (k34_trainer.cu) behind abub_train_add_events; what differs:
//
//   * the scratch of the batched gate -- the [n][16] table of bin counts that the CTAs of a pair accumulate into, which
//     abub_train_add_* keeps in a cudaMalloc'd arena of the context -- lives here in a file-scope
//     `static __device__ __align__(256)` array of ABUB_GATE_STATIC_MAX_PAIRS x 16 counters (64 KiB).  Every GPU that
//     loads the library holds it until the process exits;
//   * no context, no pointer table, no device->host read: inputs and outputs are the caller's device buffers, the
//     entropy's 16 float terms are formed on the device, everything is queued on the caller's stream and the call
//     returns without waiting;
//   * it calls no cudaMalloc / cudaFree / cudaHostAlloc and no allocating library routine.
//
// In host code `g_gate_scratch` is only a stand-in for the symbol, not a device address: its address is taken with
// cudaGetSymbolAddress once per device, with that device current, and handed to the kernels as an argument.
//
// Every call on a device shares the one array, and the array keeps the counts of the previous call: the slice a call
// uses is cleared on the stream before the gate kernel runs.  Two calls whose work overlaps on the GPU (different
// streams without an ordering between them, or two host threads) race on it -- callers serialise.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "../../include/abub_cuda.h"

// the scratch: [pair][16] bin counts, 64 KiB, held by every GPU that loads the library until the process exits
static __device__ __align__(256) uint32_t g_gate_scratch[ABUB_GATE_STATIC_MAX_PAIRS * 16];

namespace {

constexpr int GS_MAX_DEVICES = 64;

std::mutex g_mu;
uint32_t* g_scratch_ptr[GS_MAX_DEVICES] = {nullptr};   // device address of g_gate_scratch on device d (null = not asked yet)

// grid (32, n): blockIdx.y = pair; pair e = frame 0 at pairs + e*2*npix, frame 1 right behind it.  bin = sat(f1-f0) >> 4;
// bin 0 is not counted (quiet pairs put nearly every pixel there) and recovered by the finishing kernel.
__global__ void gate_static_kernel(const uint8_t* __restrict__ pairs, uint32_t* __restrict__ scratch, uint32_t npix) {
    __shared__ uint32_t sh[16];
    const uint8_t* __restrict__ f0 = pairs + (size_t)blockIdx.y * 2 * npix;
    const uint8_t* __restrict__ f1 = f0 + npix;
    if (threadIdx.x < 16) sh[threadIdx.x] = 0;
    __syncthreads();
    const bool aligned = (((uintptr_t)f0 | (uintptr_t)f1) & 15) == 0;
    const uint32_t nvec = aligned ? npix / 16 : 0;
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(f1) + i);
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(f0) + i);
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
    for (uint32_t p = nvec * 16 + blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += stride) {
        const int v = (int)f1[p] - (int)f0[p];
        const uint32_t bin = (uint32_t)(v > 0 ? v : 0) >> 4;
        if (bin) atomicAdd(&sh[bin], 1u);
    }
    __syncthreads();
    if (threadIdx.x > 0 && threadIdx.x < 16 && sh[threadIdx.x]) atomicAdd(&scratch[16 * blockIdx.y + threadIdx.x], sh[threadIdx.x]);
}

// one thread per pair: bin 0 by subtraction, then Trainer::calculateEntropyFrame's tail (Trainer.cpp:367-373):
// p = h * (float)(1/N), E -= p * log2(p) in float, and the gate E <= 0.0005 (Trainer.cpp:287)
__global__ void gate_static_finish_kernel(uint32_t* __restrict__ scratch, int n, uint32_t npix, float scale,
                                          float* __restrict__ entropy, int* __restrict__ kept, uint32_t* __restrict__ hist16) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    uint32_t h[16];
    uint32_t s = 0;
    for (int i = 1; i < 16; i++) { h[i] = scratch[16 * e + i]; s += h[i]; }
    h[0] = npix - s;
    float ent = 0.0f;
    for (int i = 0; i < 16; i++) {
        const float p = __fmul_rn((float)h[i], scale);
        if (p != 0.0f) ent = __fsub_rn(ent, __fmul_rn(p, log2f(p)));
        if (hist16) hist16[16 * e + i] = h[i];
    }
    entropy[e] = ent;
    if (kept) kept[e] = ent <= 0.0005f ? 1 : 0;
}

int fail(char* err, int cap, int code, const char* msg, const char* detail = nullptr) {
    if (err && cap > 0) snprintf(err, (size_t)cap, detail ? "abub_gate_entropy_static: %s (%s)" : "abub_gate_entropy_static: %s", msg, detail);
    return code;
}

}  // namespace

extern "C" int abub_gate_entropy_static(const uint8_t* d_pairs, int n_pairs, int rows, int cols, float* d_entropy, int* d_kept,
                                        uint32_t* d_hist16, void* stream, char* err, int err_cap) {
    if (err && err_cap > 0) err[0] = 0;
    if (n_pairs < 0 || rows <= 0 || cols <= 0) return fail(err, err_cap, ABUB_E_INVALID, "n_pairs, rows and cols must be positive");
    if (n_pairs > ABUB_GATE_STATIC_MAX_PAIRS) {
        char d[96];
        snprintf(d, sizeof d, "%d pairs asked, the static scratch holds %d", n_pairs, ABUB_GATE_STATIC_MAX_PAIRS);
        return fail(err, err_cap, ABUB_E_CAPACITY, "too many frame pairs for one call: split the batch", d);
    }
    if ((long long)rows * cols > 0x7fffffffLL) return fail(err, err_cap, ABUB_E_CAPACITY, "rows*cols must stay below 2^31 pixels");
    if (n_pairs == 0) return ABUB_OK;
    if (!d_pairs || !d_entropy) return fail(err, err_cap, ABUB_E_INVALID, "d_pairs and d_entropy must be device pointers, not NULL");

    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(err, err_cap, ABUB_E_CUDA, "cudaGetDevice", cudaGetErrorString(e));
    if (dev < 0 || dev >= GS_MAX_DEVICES) return fail(err, err_cap, ABUB_E_INVALID, "device ordinal out of range");
    uint32_t* scratch = nullptr;
    {
        // set-up, once per device, with that device current (it is: it is the caller's current device)
        std::lock_guard<std::mutex> lk(g_mu);
        if (!g_scratch_ptr[dev]) {
            uint32_t* p = nullptr;
            e = cudaGetSymbolAddress((void**)&p, g_gate_scratch);
            if (e != cudaSuccess || !p) return fail(err, err_cap, ABUB_E_CUDA, "cudaGetSymbolAddress(g_gate_scratch)", cudaGetErrorString(e));
            g_scratch_ptr[dev] = p;
        }
        scratch = g_scratch_ptr[dev];
    }
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t npix = (uint32_t)((long long)rows * cols);
    // the array keeps the previous call's counts: clear this call's slice first (stream-ordered, allocates nothing)
    e = cudaMemsetAsync(scratch, 0, (size_t)n_pairs * 16 * sizeof(uint32_t), st);
    if (e != cudaSuccess) return fail(err, err_cap, ABUB_E_CUDA, "cudaMemsetAsync(scratch)", cudaGetErrorString(e));
    gate_static_kernel<<<dim3(32, (unsigned)n_pairs), 256, 0, st>>>(d_pairs, scratch, npix);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(err, err_cap, ABUB_E_CUDA, "gate_static_kernel launch", cudaGetErrorString(e));
    const float scale = (float)(1.0 / (double)npix);
    gate_static_finish_kernel<<<(n_pairs + 127) / 128, 128, 0, st>>>(scratch, n_pairs, npix, scale, d_entropy, d_kept, d_hist16);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(err, err_cap, ABUB_E_CUDA, "gate_static_finish_kernel launch", cudaGetErrorString(e));
    return ABUB_OK;
}
