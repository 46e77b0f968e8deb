// NCCL side of the path: the one collective is the all-reduce of the Trainer's exact
// integer partials; this file adds what a process-per-GPU host needs around it -- communicator set-up
// through an id file, a barrier, and the exact boundary fix-up (all-gather of the flagged pixels'
// samples, then the reference's sequential float32 recurrence in global push order).
// libnccl is resolved with dlopen/dlsym: libabub_cuda.so has no link-time NCCL dependency.
#include <dlfcn.h>
#include <unistd.h>

#include <cub/cub.cuh>

#include "abub_internal.cuh"

namespace {

struct NcclId { char internal[128]; };      // ncclUniqueId
typedef int (*fn_get_id)(NcclId*);
typedef int (*fn_init_rank)(void**, int, NcclId, int);
typedef int (*fn_destroy)(void*);
typedef int (*fn_count)(void*, int*);
typedef int (*fn_allreduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_allgather)(const void*, void*, size_t, int, void*, cudaStream_t);

void* nccl_sym(const char* name) {
    static void* handle = nullptr;
    void* s = dlsym(RTLD_DEFAULT, name);      // already loaded by the host process (e.g. torch)
    if (s) return s;
    if (!handle) handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!handle) handle = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    return handle ? dlsym(handle, name) : nullptr;
}

enum { NCCL_U8 = 1, NCCL_I32 = 2, NCCL_SUM = 0 };

__device__ __forceinline__ void wf_step(float x, float cnt, float& mean, float& m2) {   // Trainer.cpp:185-187
    const float delta = __fsub_rn(x, mean);
    mean = __fadd_rn(mean, __fdiv_rn(delta, cnt));
    m2 = __fadd_rn(m2, __fmul_rn(delta, __fsub_rn(x, mean)));
}

// samples of the flagged pixels from this rank's kept frames: out[i*stride + f].  A block moves 32 pixels x 32 frames
// through shared memory: the reads run along the (ascending) pixel list inside one frame, the writes along the frames
// of one pixel, both contiguous.  grid.x = ceil(n_flagged / 32), block = (32, 8).
__global__ void gather_flagged_kernel(const uint8_t* const* __restrict__ frames, int n_local, const int* __restrict__ idx,
                                      int n_flagged, int stride, uint8_t* __restrict__ out) {
    __shared__ uint8_t tile[32][33];
    const int i0 = blockIdx.x * 32;
    const int pi = i0 + threadIdx.x;
    const int p = pi < n_flagged ? idx[pi] : -1;
    for (int f0 = 0; f0 < n_local; f0 += 32) {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int f = f0 + threadIdx.y + 8 * k;
            tile[threadIdx.y + 8 * k][threadIdx.x] = (f < n_local && p >= 0) ? frames[f][p] : (uint8_t)0;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int i = i0 + threadIdx.y + 8 * k, f = f0 + threadIdx.x;
            if (i < n_flagged && f < n_local) out[(size_t)i * stride + f] = tile[threadIdx.x][threadIdx.y + 8 * k];
        }
        __syncthreads();
    }
}

// sequential float32 recurrence over the ranks' samples in rank order (= global push order)
__global__ void fixup_gathered_kernel(const uint8_t* __restrict__ all, const int* __restrict__ counts, int world,
                                      const int* __restrict__ idx, int n_flagged, int stride, int n_total,
                                      uint8_t* __restrict__ mean_out, uint8_t* __restrict__ sigma_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_flagged) return;
    float mean = 0.f, m2 = 0.f;
    int k = 0;
    for (int r = 0; r < world; r++) {
        const uint8_t* src = all + ((size_t)r * n_flagged + i) * stride;
        for (int f = 0; f < counts[r]; f++) wf_step((float)src[f], (float)(++k), mean, m2);
    }
    const float var = __fdiv_rn(m2, (float)(n_total - 1));                 // Trainer.cpp:191 (0/0 -> NaN when N == 1)
    const float sg = __fsqrt_rn(var);
    const int p = idx[i];
    sigma_out[p] = (sg == sg) ? (uint8_t)(int)sg : (uint8_t)0;             // (int)NaN is INT_MIN on x86 -> u8 0
    mean_out[p] = (uint8_t)mean;
}

}  // namespace

extern "C" int abub_nccl_comm_create(abub_ctx* ctx, const char* id_file, int rank, int world, void** comm) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, id_file && comm && world >= 1 && rank >= 0 && rank < world, "abub_nccl_comm_create");
    fn_get_id get_id = (fn_get_id)nccl_sym("ncclGetUniqueId");
    fn_init_rank init_rank = (fn_init_rank)nccl_sym("ncclCommInitRank");
    if (!get_id || !init_rank) { ctx->err = "libnccl not found"; return ABUB_E_NCCL; }
    cudaSetDevice(ctx->device);
    NcclId id;
    const std::string tmp = std::string(id_file) + ".tmp";
    if (rank == 0) {
        if (get_id(&id)) { ctx->err = "ncclGetUniqueId failed"; return ABUB_E_NCCL; }
        FILE* f = fopen(tmp.c_str(), "wb");
        if (!f || fwrite(&id, 1, sizeof(id), f) != sizeof(id)) { if (f) fclose(f); ctx->err = "cannot write the NCCL id file"; return ABUB_E_NCCL; }
        fclose(f);
        rename(tmp.c_str(), id_file);        // atomic publish
    } else {
        bool got = false;
        for (int t = 0; t < 6000 && !got; t++) {           // up to 10 minutes
            FILE* f = fopen(id_file, "rb");
            if (f) { got = fread(&id, 1, sizeof(id), f) == sizeof(id); fclose(f); }
            if (!got) usleep(100000);
        }
        if (!got) { ctx->err = "timed out waiting for the NCCL id file"; return ABUB_E_NCCL; }
    }
    void* c = nullptr;
    const int r = init_rank(&c, world, id, rank);
    if (r) { ctx->err = "ncclCommInitRank failed with code " + std::to_string(r); return ABUB_E_NCCL; }
    *comm = c;
    if (rank == 0) {
        // every rank has read the id once the communicator exists
        unlink(id_file);
    }
    return ABUB_OK;
}

extern "C" int abub_nccl_comm_destroy(void* comm) {
    fn_destroy d = (fn_destroy)nccl_sym("ncclCommDestroy");
    if (!d || !comm) return ABUB_E_NCCL;
    return d(comm) ? ABUB_E_NCCL : ABUB_OK;
}

extern "C" int abub_nccl_barrier(abub_ctx* ctx, void* comm) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, comm != nullptr, "abub_nccl_barrier");
    fn_allreduce ar = (fn_allreduce)nccl_sym("ncclAllReduce");
    if (!ar) { ctx->err = "libnccl not found"; return ABUB_E_NCCL; }
    cudaSetDevice(ctx->device);
    int* d = reinterpret_cast<int*>(ctx->d_hist_tmp);
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(d, 0, sizeof(int), ctx->stream));
    if (ar(d, d, 1, NCCL_I32, NCCL_SUM, comm, ctx->stream)) { ctx->err = "ncclAllReduce (barrier) failed"; return ABUB_E_NCCL; }
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ABUB_OK;
}

// After abub_train_partials + abub_train_allreduce: planes from the reduced partials, then the flagged
// pixels are recomputed exactly from every rank's samples.  Result replicated on all ranks.
extern "C" int abub_train_finish_allreduced(abub_ctx* ctx, int cam, void* comm, uint8_t* mean, uint8_t* sigma,
                                            int* set_size, int* status, int* n_flagged) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams && comm, "abub_train_finish_allreduced");
    abub_cam_state& cs = ctx->cam[cam];
    ABUB_REQUIRE(ctx, cs.d_sum != nullptr, "abub_train_partials not called");
    fn_count cnt_fn = (fn_count)nccl_sym("ncclCommCount");
    fn_allgather ag = (fn_allgather)nccl_sym("ncclAllGather");
    if (!cnt_fn || !ag) { ctx->err = "libnccl not found"; return ABUB_E_NCCL; }
    cudaSetDevice(ctx->device);
    int world = 1;
    if (cnt_fn(comm, &world)) { ctx->err = "ncclCommCount failed"; return ABUB_E_NCCL; }
    int cnt[4];
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(cnt, cs.d_count, sizeof(cnt), cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int n = cnt[0];
    if (set_size) *set_size = n;
    if (n_flagged) *n_flagged = 0;
    cs.training = false;
    if (n == 0) { if (status) *status = -7; return ABUB_OK; }
    int rc = abub_k4_finish_from_partials(ctx, cs.d_sum, cs.d_sumsq, n, cs.d_mean, cs.d_sigma, cs.d_count + 1, cs.d_flag);
    if (rc) return rc;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(cnt, cs.d_count, sizeof(cnt), cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int F = cnt[1];                      // identical on every rank: flags derive from the reduced partials
    if (n_flagged) *n_flagged = F;
    // scratch kept from call to call (one camera after the other reuses it): [0] counts, [1] flagged indices, [2] count of
    // selected, [3] cub temp, [4] this rank's samples, [5] every rank's samples
    auto need = [&](int i, size_t bytes) -> int {
        if (bytes <= ctx->fix_cap[i]) return ABUB_OK;
        cudaFree(ctx->fix_buf[i]);
        ctx->fix_buf[i] = nullptr;
        ctx->fix_cap[i] = 0;
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->fix_buf[i], bytes + bytes / 4));
        ctx->fix_cap[i] = bytes + bytes / 4;
        return ABUB_OK;
    };
    // kept-frame counts of every rank
    const int n_local = (int)cs.kept.size();
    if (int rc2 = need(0, (size_t)(world + 1) * sizeof(int))) return rc2;
    int* d_counts = (int*)ctx->fix_buf[0];
    // 4 bytes from the stack: a pageable source is staged by the runtime before cudaMemcpyAsync returns
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(d_counts + world, &n_local, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    int nr = ag(d_counts + world, d_counts, 1, NCCL_I32, comm, ctx->stream);
    std::vector<int> counts(world);
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(counts.data(), d_counts, world * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (nr) { ctx->err = "ncclAllGather(counts) failed"; return ABUB_E_NCCL; }
    if (F > 0) {
        int stride = 1;
        for (int c : counts) stride = c > stride ? c : stride;
        stride = (stride + 15) & ~15;
        size_t tmp_bytes = 0;
        cub::CountingInputIterator<int> it(0);
        cub::DeviceSelect::Flagged(nullptr, tmp_bytes, it, cs.d_flag, (int*)nullptr, (int*)nullptr, (int)ctx->npix, ctx->stream);
        // the flagged pixels go through the gather in tiles so that the gathered samples (world x tile x stride bytes on
        // every rank) stay under 256 MB however many frames were kept and however many pixels are flagged; F, world and
        // stride are the same on every rank, so every rank walks the same tiles
        const size_t budget = (size_t)256 << 20;
        size_t tile = budget / ((size_t)world * stride);
        tile = tile < 1024 ? 1024 : tile;
        if (tile > (size_t)F) tile = (size_t)F;
        int rc2;
        if ((rc2 = need(1, (size_t)F * sizeof(int))) || (rc2 = need(2, sizeof(int))) || (rc2 = need(3, tmp_bytes)) ||
            (rc2 = need(4, tile * stride)) || (rc2 = need(5, (size_t)world * tile * stride)))
            return rc2;
        int* d_idx = (int*)ctx->fix_buf[1];
        int* d_nsel = (int*)ctx->fix_buf[2];
        uint8_t* d_loc = (uint8_t*)ctx->fix_buf[4];
        uint8_t* d_all = (uint8_t*)ctx->fix_buf[5];
        cub::DeviceSelect::Flagged(ctx->fix_buf[3], tmp_bytes, it, cs.d_flag, d_idx, d_nsel, (int)ctx->npix, ctx->stream);
        for (size_t t0 = 0; t0 < (size_t)F && !nr; t0 += tile) {
            const int ft = (int)((size_t)F - t0 < tile ? (size_t)F - t0 : tile);
            ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(d_loc, 0, (size_t)ft * stride, ctx->stream));
            ABUB_L(ctx), gather_flagged_kernel<<<(ft + 31) / 32, dim3(32, 8), 0, ctx->stream>>>(cs.d_table, n_local, d_idx + t0, ft, stride, d_loc);
            nr = ag(d_loc, d_all, (size_t)ft * stride, NCCL_U8, comm, ctx->stream);
            if (!nr)
                ABUB_L(ctx), fixup_gathered_kernel<<<(ft + 127) / 128, 128, 0, ctx->stream>>>(d_all, d_counts, world, d_idx + t0, ft, stride, n,
                                                                                   cs.d_mean, cs.d_sigma);
        }
        if (nr) { cudaStreamSynchronize(ctx->stream); ctx->err = "ncclAllGather(samples) failed"; return ABUB_E_NCCL; }
    }
    cs.training_set_size = n;
    cs.has_background = true;
    if (status) *status = 0;
    if (mean) ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(mean, cs.d_mean, ctx->npix, cudaMemcpyDeviceToHost, ctx->stream));
    if (sigma) ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(sigma, cs.d_sigma, ctx->npix, cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ABUB_OK;
}
