// Standalone probe for the TMA features k1_trigger_tma.cu relies on (run on the GPU box):
// tensor maps in global memory vs kernel params, 3-D u8 / 2-D u16 boxes, negative coordinates.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int DIMS>
__device__ void run(const CUtensorMap* map, int x, int y, int z, int bytes, uint8_t* out) {
    extern __shared__ __align__(1024) uint8_t sm[];
    __shared__ __align__(8) uint64_t bar;
    const uint32_t b = smem_u32(&bar), dst = smem_u32(sm);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
        if (DIMS == 3)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(z), "r"(b) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(b) : "memory");
    }
    asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(b) : "memory");
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = sm[i];
}
__global__ void k_param3(const __grid_constant__ CUtensorMap map, int x, int y, int z, int bytes, uint8_t* out) { run<3>(&map, x, y, z, bytes, out); }
__global__ void k_glob3(const CUtensorMap* map, int x, int y, int z, int bytes, uint8_t* out) { run<3>(map, x, y, z, bytes, out); }
__global__ void k_glob2(const CUtensorMap* map, int x, int y, int bytes, uint8_t* out) { run<2>(map, x, y, 0, bytes, out); }

#include <stdlib.h>
int main(int argc, char** argv) {
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaFree(0));
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    PFN_encodeTiled enc = (PFN_encodeTiled)p;
    const int cols = 400, rows = 300, nz = 12;
    std::vector<uint8_t> h((size_t)cols * rows * nz);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint8_t)((i * 7 + i / cols) & 0xFF);
    uint8_t* d; CK(cudaMalloc(&d, h.size())); CK(cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
    std::vector<uint16_t> h16((size_t)cols * rows);
    for (size_t i = 0; i < h16.size(); i++) h16[i] = (uint16_t)(i * 3);
    uint16_t* d16; CK(cudaMalloc(&d16, h16.size() * 2)); CK(cudaMemcpy(d16, h16.data(), h16.size() * 2, cudaMemcpyHostToDevice));
    uint8_t* out; CK(cudaMalloc(&out, 8192));
    CUtensorMap m3, m2; const cuuint32_t ones[3] = {1, 1, 1};
    {
        const cuuint64_t dims[3] = {cols, rows, nz}; const cuuint64_t str[2] = {cols, (cuuint64_t)cols * rows};
        const cuuint32_t box[3] = {256, 4, 1};
        CUresult r = enc(&m3, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, str, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("encode 3d u8: %d\n", (int)r);
    }
    {
        const cuuint64_t dims[2] = {cols, rows}; const cuuint64_t str[1] = {cols * 2};
        const cuuint32_t box[2] = {256, 4};
        CUresult r = enc(&m2, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, d16, dims, str, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("encode 2d u16: %d\n", (int)r);
    }
    CUtensorMap* dm; CK(cudaMalloc(&dm, 2 * sizeof(CUtensorMap)));
    CK(cudaMemcpy(dm, &m3, sizeof(m3), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dm + 1, &m2, sizeof(m2), cudaMemcpyHostToDevice));
    std::vector<uint8_t> o(8192);
    auto check3 = [&](const char* name, int x, int y, int z) {
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
        cudaMemcpy(o.data(), out, 1024, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int r = 0; r < 4; r++) for (int c = 0; c < 256; c++) {
            int gx = x + c, gy = y + r; uint8_t want = 0;
            if (gx >= 0 && gx < cols && gy >= 0 && gy < rows) want = h[((size_t)z * rows + gy) * cols + gx];
            bad += (o[r * 256 + c] != want);
        }
        printf("%s (x=%d y=%d z=%d): %d mismatches\n", name, x, y, z, bad);
    };
    int X = argc > 1 ? atoi(argv[1]) : 16, Y = argc > 2 ? atoi(argv[2]) : 8;
    k_glob3<<<1, 64, 4096>>>(dm, X, Y, 3, 1024, out); check3("global 3d", X, Y, 3);
    k_glob2<<<1, 64, 4096>>>(dm + 1, X, 5, 2048, out);
    {
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) printf("global 2d u16: %s\n", cudaGetErrorString(e));
        else {
            cudaMemcpy(o.data(), out, 2048, cudaMemcpyDeviceToHost);
            int bad = 0; const uint16_t* o16 = (const uint16_t*)o.data();
            for (int r = 0; r < 4; r++) for (int c = 0; c < 256; c++) {
                int gx = X + c, gy = 5 + r; uint16_t want = 0;
                if (gx >= 0 && gx < cols) want = h16[(size_t)gy * cols + gx];
                bad += (o16[r * 256 + c] != want);
            }
            printf("global 2d u16 neg x: %d mismatches\n", bad);
        }
    }
    return 0;
}
