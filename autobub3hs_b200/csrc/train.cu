// C-ABI of the Trainer path: Trainer::MakeAvgSigmaImage (Trainer.cpp:221-331).
// Kept frames (TrainingSequence {0,1} of every event that passes the entropy gate,
// Trainer.hpp:54, Trainer.cpp:287-292) are parked in device chunks in push order; the
// statistics are one pass of K4 over that table at abub_train_finish.
#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>

#include "abub_internal.cuh"

static int ensure_room(abub_ctx* ctx, abub_cam_state& cs, int nframes_more) {
    while ((int)cs.chunks.size() * ABUB_TRAIN_CHUNK < cs.n_stored + nframes_more) {
        uint8_t* p = nullptr;
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&p, (size_t)ABUB_TRAIN_CHUNK * ctx->npix));
        cs.chunks.push_back(p);
    }
    return ABUB_OK;
}

static uint8_t* store_ptr(abub_ctx* ctx, abub_cam_state& cs, int idx) {
    return cs.chunks[idx / ABUB_TRAIN_CHUNK] + (size_t)(idx % ABUB_TRAIN_CHUNK) * ctx->npix;
}

extern "C" int abub_train_begin(abub_ctx* ctx, int cam) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams, "cam");
    cudaSetDevice(ctx->device);
    abub_cam_state& cs = ctx->cam[cam];
    cs.kept.clear();
    cs.n_stored = 0;
    cs.training = true;
    return ABUB_OK;
}

// Trainer::calculateEntropyFrame tail (Trainer.cpp:367-373): p = h * (float)(1/N) (MatExpr scale
// applied by convertTo in float), E -= p*log2(p) in float.
static float entropy16(const uint32_t* h, size_t npix) {
    const float scale = (float)(1.0 / (double)npix);
    float e = 0.0f;
    for (int i = 0; i < 16; i++) {
        const float p = (float)h[i] * scale;
        if (p != 0) e -= p * log2f(p);
    }
    return e;
}

static int gate_and_keep(abub_ctx* ctx, int cam, float* entropy, int* kept) {
    abub_cam_state& cs = ctx->cam[cam];
    uint8_t* f0 = store_ptr(ctx, cs, cs.n_stored);
    uint8_t* f1 = store_ptr(ctx, cs, cs.n_stored + 1);
    int rc = abub_k3_gate_hist16(ctx, f0, f1, ctx->d_hist_tmp);
    if (rc) return rc;
    uint32_t h[16];
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(h, ctx->d_hist_tmp, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const float e = entropy16(h, ctx->npix);
    const bool keep = e <= 0.0005;                                   // Trainer.cpp:287
    if (keep) {
        cs.kept.push_back(f0);
        cs.kept.push_back(f1);
        cs.n_stored += 2;
    }
    if (entropy) *entropy = e;
    if (kept) *kept = keep ? 1 : 0;
    return ABUB_OK;
}

extern "C" int abub_train_add_event(abub_ctx* ctx, int cam, const uint8_t* frames01, float* entropy, int* kept) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams && frames01, "abub_train_add_event");
    ABUB_REQUIRE(ctx, ctx->cam[cam].training, "abub_train_begin not called");
    cudaSetDevice(ctx->device);
    abub_cam_state& cs = ctx->cam[cam];
    int rc = ensure_room(ctx, cs, 2);
    if (rc) return rc;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(store_ptr(ctx, cs, cs.n_stored), frames01, ctx->npix, cudaMemcpyHostToDevice,
                                         ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(store_ptr(ctx, cs, cs.n_stored + 1), frames01 + ctx->npix, ctx->npix,
                                         cudaMemcpyHostToDevice, ctx->stream));
    return gate_and_keep(ctx, cam, entropy, kept);
}

extern "C" int abub_train_add_slot(abub_ctx* ctx, int cam, int slot, float* entropy, int* kept) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams && slot >= 0 && slot < ctx->max_slots, "cam/slot");
    ABUB_REQUIRE(ctx, ctx->cam[cam].training, "abub_train_begin not called");
    ABUB_REQUIRE(ctx, ctx->slots[slot].nframes[cam] >= 2, "slot has fewer than 2 frames for this camera");
    cudaSetDevice(ctx->device);
    abub_cam_state& cs = ctx->cam[cam];
    int rc = ensure_room(ctx, cs, 2);
    if (rc) return rc;
    const uint8_t* src = ctx->slots[slot].d_frames + (size_t)cam * ctx->max_frames * ctx->npix;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(store_ptr(ctx, cs, cs.n_stored), src, ctx->npix, cudaMemcpyDeviceToDevice,
                                         ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(store_ptr(ctx, cs, cs.n_stored + 1), src + ctx->npix, ctx->npix,
                                         cudaMemcpyDeviceToDevice, ctx->stream));
    return gate_and_keep(ctx, cam, entropy, kept);
}


// ---- batched forms: one K3 launch and one device->host read for n events --------------------------
static int ensure_gate(abub_ctx* ctx, int n) {
    if (n <= ctx->gate_cap) return ABUB_OK;
    cudaFree((void*)ctx->d_gate_tab);
    cudaFree(ctx->d_gate_hist);
    ctx->d_gate_tab = nullptr;
    ctx->d_gate_hist = nullptr;
    ctx->gate_cap = 0;
    const int cap = n + n / 2 + 64;
    ABUB_CHECK_CUDA(ctx, cudaMalloc((void**)&ctx->d_gate_tab, (size_t)cap * 2 * sizeof(uint8_t*)));
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_gate_hist, (size_t)cap * 16 * sizeof(uint32_t)));
    ctx->gate_cap = cap;
    return ABUB_OK;
}

// gate of n pairs given by host table `pairs` (device pointers); fills e[] and keep[]
static int gate_batch(abub_ctx* ctx, const std::vector<const uint8_t*>& pairs, int n, std::vector<float>& e,
                      std::vector<char>& keep) {
    int rc = ensure_gate(ctx, n);
    if (rc) return rc;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync((void*)ctx->d_gate_tab, pairs.data(), (size_t)n * 2 * sizeof(uint8_t*),
                                         cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = abub_k3_gate_batch(ctx, ctx->d_gate_tab, n, ctx->d_gate_hist))) return rc;
    std::vector<uint32_t> h((size_t)n * 16);
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(h.data(), ctx->d_gate_hist, h.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                         ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    e.resize(n);
    keep.resize(n);
    for (int i = 0; i < n; i++) {
        e[i] = entropy16(&h[(size_t)i * 16], ctx->npix);
        keep[i] = e[i] <= 0.0005;                                   // Trainer.cpp:287
    }
    return ABUB_OK;
}

extern "C" int abub_train_add_slots(abub_ctx* ctx, int cam, const int* slots, int n, float* entropy, int* kept) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams && slots && n >= 0, "abub_train_add_slots");
    ABUB_REQUIRE(ctx, ctx->cam[cam].training, "abub_train_begin not called");
    if (n == 0) return ABUB_OK;
    cudaSetDevice(ctx->device);
    abub_cam_state& cs = ctx->cam[cam];
    std::vector<const uint8_t*> pairs((size_t)n * 2);
    for (int i = 0; i < n; i++) {
        ABUB_REQUIRE(ctx, slots[i] >= 0 && slots[i] < ctx->max_slots, "slot");
        ABUB_REQUIRE(ctx, ctx->slots[slots[i]].nframes[cam] >= 2, "slot has fewer than 2 frames for this camera");
        const uint8_t* src = ctx->slots[slots[i]].d_frames + (size_t)cam * ctx->max_frames * ctx->npix;
        pairs[2 * i] = src;
        pairs[2 * i + 1] = src + ctx->npix;
    }
    std::vector<float> e;
    std::vector<char> keep;
    int rc = gate_batch(ctx, pairs, n, e, keep);
    if (rc) return rc;
    // the kept pairs stay where they are: K4 / the partial sums read them through the frame table in place (no parking
    // copy, no wait) -- the caller keeps these slots untouched until abub_train_finish* / abub_train_partials has run
    for (int i = 0; i < n; i++) {
        if (entropy) entropy[i] = e[i];
        if (kept) kept[i] = keep[i];
        if (!keep[i]) continue;
        cs.kept.push_back(pairs[2 * i]);
        cs.kept.push_back(pairs[2 * i + 1]);
    }
    return ABUB_OK;
}

extern "C" int abub_train_add_events(abub_ctx* ctx, int cam, const uint8_t* frames01, int n, float* entropy, int* kept) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams && frames01 && n >= 0, "abub_train_add_events");
    ABUB_REQUIRE(ctx, ctx->cam[cam].training, "abub_train_begin not called");
    if (n == 0) return ABUB_OK;
    cudaSetDevice(ctx->device);
    abub_cam_state& cs = ctx->cam[cam];
    int rc = ensure_room(ctx, cs, 2 * n);
    if (rc) return rc;
    // every pair is uploaded to the next free positions; a rejected pair's position is simply not listed in
    // the frame table (ABUB_TRAIN_CHUNK is even, so a pair never straddles two chunks)
    std::vector<const uint8_t*> pairs((size_t)n * 2);
    for (int i = 0; i < n; i++) {
        uint8_t* f0 = store_ptr(ctx, cs, cs.n_stored + 2 * i);
        pairs[2 * i] = f0;
        pairs[2 * i + 1] = f0 + ctx->npix;
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(f0, frames01 + (size_t)i * 2 * ctx->npix, 2 * ctx->npix, cudaMemcpyHostToDevice,
                                             ctx->stream));
    }
    std::vector<float> e;
    std::vector<char> keep;
    if ((rc = gate_batch(ctx, pairs, n, e, keep))) return rc;
    // compact: a kept pair moves down over the holes left by rejected ones (device copy, rarely needed)
    int w = cs.n_stored;
    for (int i = 0; i < n; i++) {
        if (entropy) entropy[i] = e[i];
        if (kept) kept[i] = keep[i];
        if (!keep[i]) continue;
        uint8_t* dst = store_ptr(ctx, cs, w);
        if (dst != pairs[2 * i])
            ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(dst, pairs[2 * i], 2 * ctx->npix, cudaMemcpyDeviceToDevice, ctx->stream));
        cs.kept.push_back(dst);
        cs.kept.push_back(dst + ctx->npix);
        w += 2;
    }
    cs.n_stored = w;
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ABUB_OK;
}

static int upload_table(abub_ctx* ctx, abub_cam_state& cs) {
    const int n = (int)cs.kept.size();
    if (n > cs.table_cap) {
        cudaFree((void*)cs.d_table);
        cs.d_table = nullptr;
        cs.table_cap = n + 64;
        ABUB_CHECK_CUDA(ctx, cudaMalloc((void**)&cs.d_table, cs.table_cap * sizeof(uint8_t*)));
    }
    if (n > 0) {
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync((void*)cs.d_table, cs.kept.data(), n * sizeof(uint8_t*),
                                             cudaMemcpyHostToDevice, ctx->stream));
        ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return ABUB_OK;
}

static int read_planes(abub_ctx* ctx, abub_cam_state& cs, uint8_t* mean, uint8_t* sigma) {
    if (mean) ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(mean, cs.d_mean, ctx->npix, cudaMemcpyDeviceToHost, ctx->stream));
    if (sigma) ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(sigma, cs.d_sigma, ctx->npix, cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ABUB_OK;
}

extern "C" int abub_train_finish(abub_ctx* ctx, int cam, uint8_t* mean, uint8_t* sigma, int* set_size, int* status) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams, "cam");
    ABUB_REQUIRE(ctx, ctx->cam[cam].training, "abub_train_begin not called");
    cudaSetDevice(ctx->device);
    abub_cam_state& cs = ctx->cam[cam];
    const int n = (int)cs.kept.size();
    if (set_size) *set_size = n;
    if (n == 0) {                                                   // Trainer.cpp:308-313
        if (status) *status = -7;
        cs.training = false;
        return ABUB_OK;
    }
    int rc = upload_table(ctx, cs);
    if (rc) return rc;
    rc = abub_k4_welford(ctx, cs.d_table, n, cs.d_mean, cs.d_sigma);
    if (rc) return rc;
    cs.training_set_size = n;                                       // Trainer.cpp:161
    cs.has_background = true;
    cs.training = false;
    if (status) *status = 0;
    return read_planes(ctx, cs, mean, sigma);
}

// ---- multi-GPU: exact integer partials + all-reduce ----------------------------------
static int ensure_partials(abub_ctx* ctx, abub_cam_state& cs) {
    if (!cs.d_sum) {
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_sum, ctx->npix * sizeof(uint32_t)));
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_sumsq, ctx->npix * sizeof(uint64_t)));
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_count, 4 * sizeof(int)));
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_flag, ctx->npix));
    }
    return ABUB_OK;
}

extern "C" int abub_train_partials(abub_ctx* ctx, int cam, uint32_t** d_sum, uint64_t** d_sumsq, int** d_count) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams, "cam");
    cudaSetDevice(ctx->device);
    abub_cam_state& cs = ctx->cam[cam];
    int rc = ensure_partials(ctx, cs);
    if (rc) return rc;
    rc = upload_table(ctx, cs);
    if (rc) return rc;
    const int n = (int)cs.kept.size();
    bool aligned16 = true;
    for (const uint8_t* f : cs.kept) aligned16 = aligned16 && (((uintptr_t)f) & 15) == 0;
    rc = abub_k4_partials(ctx, cs.d_table, n, cs.d_sum, cs.d_sumsq, aligned16);
    if (rc) return rc;
    const int cnt[4] = {n, 0, 0, 0};
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(cs.d_count, cnt, sizeof(cnt), cudaMemcpyHostToDevice, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (d_sum) *d_sum = cs.d_sum;
    if (d_sumsq) *d_sumsq = cs.d_sumsq;
    if (d_count) *d_count = cs.d_count;
    return ABUB_OK;
}

// libnccl is resolved at run time so that the library has no link-time NCCL dependency.
typedef int (*nccl_allreduce_fn)(const void*, void*, size_t, int, int, void*, cudaStream_t);
static nccl_allreduce_fn resolve_nccl() {
    static nccl_allreduce_fn fn = nullptr;
    static bool tried = false;
    if (tried) return fn;
    tried = true;
    void* sym = dlsym(RTLD_DEFAULT, "ncclAllReduce");   // already loaded by the host (e.g. torch)
    if (!sym) {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (h) sym = dlsym(h, "ncclAllReduce");
    }
    fn = (nccl_allreduce_fn)sym;
    return fn;
}

extern "C" int abub_train_allreduce(abub_ctx* ctx, int cam, void* nccl_comm) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams && nccl_comm, "abub_train_allreduce");
    abub_cam_state& cs = ctx->cam[cam];
    ABUB_REQUIRE(ctx, cs.d_sum != nullptr, "abub_train_partials not called");
    cudaSetDevice(ctx->device);
    nccl_allreduce_fn ar = resolve_nccl();
    if (!ar) { ctx->err = "libnccl not found (ncclAllReduce)"; return ABUB_E_NCCL; }
    // ncclDataType_t: ncclInt32=2, ncclUint32=3, ncclUint64=5 ; ncclRedOp_t: ncclSum=0
    int r = ar(cs.d_sum, cs.d_sum, ctx->npix, 3, 0, nccl_comm, ctx->stream);
    if (!r) r = ar(cs.d_sumsq, cs.d_sumsq, ctx->npix, 5, 0, nccl_comm, ctx->stream);
    if (!r) r = ar(cs.d_count, cs.d_count, 1, 2, 0, nccl_comm, ctx->stream);
    if (r) { ctx->err = "ncclAllReduce failed with code " + std::to_string(r); return ABUB_E_NCCL; }
    return ABUB_OK;
}

extern "C" int abub_train_finish_partials(abub_ctx* ctx, int cam, uint8_t* mean, uint8_t* sigma, int* set_size,
                                          int* status, int* n_flagged) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams, "cam");
    abub_cam_state& cs = ctx->cam[cam];
    ABUB_REQUIRE(ctx, cs.d_sum != nullptr, "abub_train_partials not called");
    cudaSetDevice(ctx->device);
    int cnt[4];
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(cnt, cs.d_count, sizeof(cnt), cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int n = cnt[0];
    if (set_size) *set_size = n;
    if (n == 0) {
        if (status) *status = -7;
        if (n_flagged) *n_flagged = 0;
        cs.training = false;
        return ABUB_OK;
    }
    int rc = abub_k4_finish_from_partials(ctx, cs.d_sum, cs.d_sumsq, n, cs.d_mean, cs.d_sigma, cs.d_count + 1, cs.d_flag);
    if (rc) return rc;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(cnt, cs.d_count, sizeof(cnt), cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (n_flagged) *n_flagged = cnt[1];
    // single-rank case: the full frame table is local, repair flagged pixels exactly.  ABUB_TRAIN_NO_FIXUP (test hook)
    // leaves the fp64 planes as they are so that a test can count what the repair had to change.
    if ((int)cs.kept.size() == n && !getenv("ABUB_TRAIN_NO_FIXUP")) {
        rc = abub_k4_fixup(ctx, cs.d_table, n, cs.d_flag, cs.d_mean, cs.d_sigma);
        if (rc) return rc;
    }
    cs.training_set_size = n;
    cs.has_background = true;
    cs.training = false;
    if (status) *status = 0;
    return read_planes(ctx, cs, mean, sigma);
}
