// Context management and the C-ABI entry points that are not a kernel by themselves
// (include/abub_cuda.h).  No CPU fallback anywhere: every arithmetic entry point launches
// the sm_100a kernels or fails with ABUB_E_CUDA.
#include <dlfcn.h>
#include <math.h>

#include "abub_internal.cuh"

static const char* kNoErr = "";

extern "C" const char* abub_last_error(const abub_ctx* ctx) { return ctx ? ctx->err.c_str() : kNoErr; }

extern "C" int abub_ctx_create(int device, int rows, int cols, int ncams, int max_frames, int max_slots,
                               abub_ctx** out) {
    if (!out) return ABUB_E_INVALID;
    *out = nullptr;
    if (rows < 5 || cols < 5 || ncams < 1 || ncams > ABUB_MAX_CAMS || max_frames < 2 || max_slots < 1)
        return ABUB_E_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) return ABUB_E_CUDA;
    abub_ctx* ctx = new abub_ctx();
    ctx->device = device; ctx->rows = rows; ctx->cols = cols; ctx->ncams = ncams;
    ctx->max_frames = max_frames; ctx->max_slots = max_slots;
    ctx->npix = (size_t)rows * cols;
    ctx->slots.resize(max_slots);
    *out = ctx;     // returned even on failure so that abub_last_error works; caller destroys
    ABUB_CHECK_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    const size_t n = ctx->npix;
    for (int c = 0; c < ncams; c++) {
        abub_cam_state& cs = ctx->cam[c];
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_mean, n));
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_sigma, n));
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_t8, n));
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_k32, n * sizeof(uint32_t)));
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_k16, n * sizeof(int16_t)));
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&cs.d_nt16, n * sizeof(int16_t)));
    }
    const size_t frames_per_slot = (size_t)ncams * max_frames;
    std::vector<uint8_t*> sp(max_slots);
    for (int s = 0; s < max_slots; s++) {
        ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->slots[s].d_frames, frames_per_slot * n));
        sp[s] = ctx->slots[s].d_frames;
    }
    const size_t sc = (size_t)max_slots * ncams;
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_hist, sc * max_frames * 256 * sizeof(uint32_t)));
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_nframes, sc * sizeof(int)));
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_start, sc * sizeof(int)));
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_avail, sc * sizeof(int)));
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_trig, sc * 4 * sizeof(int)));
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_sig, sc * max_frames * sizeof(double)));
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_tss, ABUB_MAX_CAMS * sizeof(int)));
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_k32_ptrs, ABUB_MAX_CAMS * sizeof(uint32_t*)));
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_slot_ptrs, max_slots * sizeof(uint8_t*)));
    ABUB_CHECK_CUDA(ctx, cudaMemset(ctx->d_nframes, 0, sc * sizeof(int)));
    ABUB_CHECK_CUDA(ctx, cudaMemset(ctx->d_tss, 0, ABUB_MAX_CAMS * sizeof(int)));
    ABUB_CHECK_CUDA(ctx, cudaMemset(ctx->d_hist, 0, sc * max_frames * 256 * sizeof(uint32_t)));
    ABUB_CHECK_CUDA(ctx, cudaMemcpy(ctx->d_slot_ptrs, sp.data(), max_slots * sizeof(uint8_t*), cudaMemcpyHostToDevice));
    uint32_t* kp[ABUB_MAX_CAMS] = {nullptr};
    for (int c = 0; c < ncams; c++) kp[c] = ctx->cam[c].d_k32;
    ABUB_CHECK_CUDA(ctx, cudaMemcpy(ctx->d_k32_ptrs, kp, sizeof(kp), cudaMemcpyHostToDevice));
    for (int i = 0; i < 4; i++) ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_tmp8[i], n));
    ABUB_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_hist_tmp, 16 * 256 * sizeof(uint32_t)));
    int rc = abub_loc_init(ctx);
    if (rc != ABUB_OK) return rc;
    return abub_k1_tma_init(ctx);
}

static void train_release(abub_cam_state& cs) {
    for (uint8_t* p : cs.chunks) cudaFree(p);
    cs.chunks.clear();
    cs.kept.clear();
    cs.n_stored = 0;
}

extern "C" void abub_ctx_destroy(abub_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    abub_loc_free(ctx);
    abub_bel_free(ctx);
    abub_png_free(ctx);
    for (int c = 0; c < ABUB_MAX_CAMS; c++) {
        abub_cam_state& cs = ctx->cam[c];
        cudaFree(cs.d_mean); cudaFree(cs.d_sigma); cudaFree(cs.d_t8); cudaFree(cs.d_k32); cudaFree(cs.d_k16); cudaFree(cs.d_nt16);
        cudaFree(cs.d_sum); cudaFree(cs.d_sumsq); cudaFree(cs.d_count); cudaFree(cs.d_flag);
        cudaFree((void*)cs.d_table);
        cudaFree(cs.d_templ);
        train_release(cs);
    }
    for (auto& s : ctx->slots) {
        cudaFree(s.d_frames);
        if (s.h_pinned) cudaFreeHost(s.h_pinned);
    }
    cudaFree(ctx->d_hist); cudaFree(ctx->d_nframes); cudaFree(ctx->d_start); cudaFree(ctx->d_avail); cudaFree(ctx->d_trig);
    cudaFree(ctx->d_sig); cudaFree(ctx->d_k2_state); cudaFree(ctx->d_tss); cudaFree(ctx->d_k32_ptrs); cudaFree(ctx->d_slot_ptrs);
    for (int i = 0; i < 4; i++) cudaFree(ctx->d_tmp8[i]);
    cudaFree(ctx->d_hist_tmp);
    cudaFree((void*)ctx->d_gate_tab); cudaFree(ctx->d_gate_hist);
    cudaFree(ctx->d_tmaps);
    cudaFree(ctx->d_post_list);
    if (ctx->h_post_list) cudaFreeHost(ctx->h_post_list);
    if (ctx->post_ev) cudaEventDestroy(ctx->post_ev);
    cudaFree(ctx->d_gen_list);
    for (void* b : ctx->fix_buf) cudaFree(b);
    if (ctx->h_gen_list) cudaFreeHost(ctx->h_gen_list);
    if (ctx->gen_ev) cudaEventDestroy(ctx->gen_ev);
    if (ctx->h_raw) cudaFreeHost(ctx->h_raw);
    for (cudaEvent_t e : ctx->prof.pool) cudaEventDestroy(e);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    if (ctx->ev_stream) cudaEventDestroy(ctx->ev_stream);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" int abub_sync(abub_ctx* ctx) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ABUB_OK;
}

extern "C" void* abub_stream(abub_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

extern "C" long long abub_launch_count(const abub_ctx* ctx) { return ctx ? ctx->n_launches : 0; }

static const char* const kProfNames[ABUB_PROF_N] = {
    "K1 k1_trigger_tma_kernel + border rows + bin0", "K2 k2_trigger_scan_kernel", "K5a k5_genesis_tma_kernel + border rows (generic pitch: k1_genesis_jobs_kernel)",
    "K5b k5_post_tma_kernel (generic pitch: k5_post_stream_kernel)", "K5c k5_otsu_kernel + flags", "K6 k6_candidates_seg_kernel", "cub sort/scan (library)",
    "K7a k7_trace1_kernel", "K7b k7_trace2_kernel", "K7c k7_external_kernel + k7_emit_kernel", "K3 k3_gate_batch_kernel",
    "K4 k4_welford16_kernel"};

static void prof_resolve(abub_ctx* ctx) {
    abub_prof_state& p = ctx->prof;
    for (const auto& r : p.recs) {
        float ms = 0.f;
        if (cudaEventSynchronize(p.pool[r.e1]) == cudaSuccess && cudaEventElapsedTime(&ms, p.pool[r.e0], p.pool[r.e1]) == cudaSuccess) {
            p.ms[r.id] += ms;
            p.calls[r.id]++;
        }
    }
    p.recs.clear();
    p.used = 0;
}

extern "C" int abub_profile(abub_ctx* ctx, int enable) {
    if (!ctx) return ABUB_E_INVALID;
    cudaSetDevice(ctx->device);
    prof_resolve(ctx);
    for (int i = 0; i < ABUB_PROF_N; i++) { ctx->prof.ms[i] = 0; ctx->prof.calls[i] = 0; }
    ctx->prof.on = enable != 0;
    return ABUB_OK;
}

extern "C" int abub_profile_read(abub_ctx* ctx, int cap, const char** names, double* ms, long long* calls) {
    if (!ctx) return ABUB_E_INVALID;
    cudaSetDevice(ctx->device);
    prof_resolve(ctx);
    const int n = cap < ABUB_PROF_N ? cap : ABUB_PROF_N;
    for (int i = 0; i < n; i++) {
        if (names) names[i] = kProfNames[i];
        if (ms) ms[i] = ctx->prof.ms[i];
        if (calls) calls[i] = ctx->prof.calls[i];
    }
    return ABUB_PROF_N;
}

// ---- background -------------------------------------------------------------------
static int install_background(abub_ctx* ctx, int cam, int tss) {
    abub_cam_state& cs = ctx->cam[cam];
    cs.training_set_size = tss;
    cs.has_background = true;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->d_tss + cam, &cs.training_set_size, sizeof(int), cudaMemcpyHostToDevice,
                                         ctx->stream));
    int rc = abub_k1_build_k32(ctx, cam);
    if (rc) return rc;
    if ((rc = abub_k1_build_k16(ctx, cam))) return rc;
    for (auto& s : ctx->slots) s.hist_valid = false;
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // &training_set_size must outlive the copy
    return ABUB_OK;
}

extern "C" int abub_set_background(abub_ctx* ctx, int cam, const uint8_t* mean, const uint8_t* sigma,
                                   int training_set_size) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams && mean && sigma, "abub_set_background");
    cudaSetDevice(ctx->device);
    abub_cam_state& cs = ctx->cam[cam];
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(cs.d_mean, mean, ctx->npix, cudaMemcpyHostToDevice, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(cs.d_sigma, sigma, ctx->npix, cudaMemcpyHostToDevice, ctx->stream));
    return install_background(ctx, cam, training_set_size);
}

extern "C" int abub_set_background_from_training(abub_ctx* ctx, int cam) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams, "cam");
    ABUB_REQUIRE(ctx, ctx->cam[cam].has_background, "no trained planes for this camera");
    cudaSetDevice(ctx->device);
    return install_background(ctx, cam, ctx->cam[cam].training_set_size);
}

// ---- event frames -----------------------------------------------------------------
static size_t slot_bytes(const abub_ctx* ctx) { return (size_t)ctx->ncams * ctx->max_frames * ctx->npix; }

// The frame counts travel as a kernel argument: a cudaMemcpyAsync from pageable memory may wait for the stream, and this
// call sits between the large uploads of the staging loop.
struct NframesArg { int n[ABUB_MAX_CAMS]; };
__global__ void set_nframes_kernel(int* __restrict__ dst, const NframesArg a, int ncams) {
    if ((int)threadIdx.x < ncams) dst[threadIdx.x] = a.n[threadIdx.x];
}

static int push_nframes(abub_ctx* ctx, int slot) {
    NframesArg a;
    for (int c = 0; c < ABUB_MAX_CAMS; c++) a.n[c] = ctx->slots[slot].nframes[c];
    ABUB_L(ctx), set_nframes_kernel<<<1, 32, 0, ctx->stream>>>(ctx->d_nframes + (size_t)slot * ctx->ncams, a, ctx->ncams);
    ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    ctx->slots[slot].hist_valid = false;
    return ABUB_OK;
}

// the same for a range of slots: up to 64 slots' counts per launch
struct NframesRangeArg { int n[64 * ABUB_MAX_CAMS]; };
__global__ void set_nframes_range_kernel(int* __restrict__ dst, const NframesRangeArg a, int count) {
    for (int i = threadIdx.x; i < count; i += blockDim.x) dst[i] = a.n[i];
}

extern "C" int abub_event_set_nframes_range(abub_ctx* ctx, int slot_first, int nslots, const int* nframes) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot_first >= 0 && nslots >= 1 && slot_first + nslots <= ctx->max_slots && nframes, "abub_event_set_nframes_range");
    cudaSetDevice(ctx->device);
    for (int i = 0; i < nslots * ctx->ncams; i++)
        ABUB_REQUIRE(ctx, nframes[i] >= 0 && nframes[i] <= ctx->max_frames, "nframes out of range");
    const int per = (64 * ABUB_MAX_CAMS) / ctx->ncams;               // slots per launch
    for (int s0 = 0; s0 < nslots; s0 += per) {
        const int ns = nslots - s0 < per ? nslots - s0 : per;
        NframesRangeArg a;
        for (int s = 0; s < ns; s++) {
            abub_slot_state& st = ctx->slots[slot_first + s0 + s];
            for (int c = 0; c < ctx->ncams; c++) a.n[s * ctx->ncams + c] = st.nframes[c] = nframes[(size_t)(s0 + s) * ctx->ncams + c];
            st.hist_valid = false;
        }
        ABUB_L(ctx), set_nframes_range_kernel<<<1, 256, 0, ctx->stream>>>(ctx->d_nframes + (size_t)(slot_first + s0) * ctx->ncams, a,
                                                                         ns * ctx->ncams);
        ABUB_CHECK_CUDA(ctx, cudaGetLastError());
    }
    return ABUB_OK;
}

extern "C" int abub_pinned_acquire(abub_ctx* ctx, int slot, uint8_t** host_ptr) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && host_ptr, "abub_pinned_acquire");
    cudaSetDevice(ctx->device);
    abub_slot_state& s = ctx->slots[slot];
    if (!s.h_pinned) ABUB_CHECK_CUDA(ctx, cudaHostAlloc(&s.h_pinned, slot_bytes(ctx), cudaHostAllocDefault));
    *host_ptr = s.h_pinned;
    return ABUB_OK;
}

extern "C" int abub_event_submit(abub_ctx* ctx, int slot, const int* nframes) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && nframes, "abub_event_submit");
    abub_slot_state& s = ctx->slots[slot];
    ABUB_REQUIRE(ctx, s.h_pinned != nullptr, "slot has no pinned buffer (call abub_pinned_acquire)");
    cudaSetDevice(ctx->device);
    for (int c = 0; c < ctx->ncams; c++) {
        ABUB_REQUIRE(ctx, nframes[c] >= 0 && nframes[c] <= ctx->max_frames, "nframes out of range");
        s.nframes[c] = nframes[c];
        const size_t off = (size_t)c * ctx->max_frames * ctx->npix;
        if (nframes[c] > 0)
            ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(s.d_frames + off, s.h_pinned + off, (size_t)nframes[c] * ctx->npix,
                                                 cudaMemcpyHostToDevice, ctx->stream));
    }
    return push_nframes(ctx, slot);
}

extern "C" int abub_event_upload(abub_ctx* ctx, int slot, int cam, const uint8_t* frames, int nframes) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && cam >= 0 && cam < ctx->ncams, "slot/cam");
    ABUB_REQUIRE(ctx, nframes >= 0 && nframes <= ctx->max_frames && (frames || nframes == 0), "frames");
    cudaSetDevice(ctx->device);
    abub_slot_state& s = ctx->slots[slot];
    s.nframes[cam] = nframes;
    if (nframes > 0) {
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(s.d_frames + (size_t)cam * ctx->max_frames * ctx->npix, frames,
                                             (size_t)nframes * ctx->npix, cudaMemcpyHostToDevice, ctx->stream));
        ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // pageable source may be reused by the caller
    }
    return push_nframes(ctx, slot);
}

// frames of one camera already decoded into the slot's pinned staging buffer: asynchronous H2D on the ctx stream
extern "C" int abub_event_submit_cam(abub_ctx* ctx, int slot, int cam, int nframes) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && cam >= 0 && cam < ctx->ncams, "slot/cam");
    ABUB_REQUIRE(ctx, nframes >= 0 && nframes <= ctx->max_frames, "nframes");
    cudaSetDevice(ctx->device);
    abub_slot_state& s = ctx->slots[slot];
    ABUB_REQUIRE(ctx, s.h_pinned != nullptr, "slot has no pinned buffer (call abub_pinned_acquire)");
    s.nframes[cam] = nframes;
    s.hist_valid = false;
    if (nframes > 0) {
        const size_t off = (size_t)cam * ctx->max_frames * ctx->npix;
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(s.d_frames + off, s.h_pinned + off, (size_t)nframes * ctx->npix, cudaMemcpyHostToDevice,
                                             ctx->stream));
    }
    return push_nframes(ctx, slot);
}

// one frame, from its region of the pinned staging buffer (asynchronous)
extern "C" int abub_event_upload_frame(abub_ctx* ctx, int slot, int cam, int frame) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && cam >= 0 && cam < ctx->ncams && frame >= 0 && frame < ctx->max_frames,
                 "slot/cam/frame");
    cudaSetDevice(ctx->device);
    abub_slot_state& s = ctx->slots[slot];
    ABUB_REQUIRE(ctx, s.h_pinned != nullptr, "slot has no pinned buffer (call abub_pinned_acquire)");
    const size_t off = ((size_t)cam * ctx->max_frames + frame) * ctx->npix;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(s.d_frames + off, s.h_pinned + off, ctx->npix, cudaMemcpyHostToDevice, ctx->stream));
    s.hist_valid = false;
    return ABUB_OK;
}

// frames [frame_first, frame_first + nframes) of one camera, from their regions of the pinned staging buffer (asynchronous).
// The frame count of the camera is not changed.
extern "C" int abub_event_submit_range(abub_ctx* ctx, int slot, int cam, int frame_first, int nframes) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && cam >= 0 && cam < ctx->ncams, "slot/cam");
    ABUB_REQUIRE(ctx, frame_first >= 0 && nframes >= 0 && frame_first + nframes <= ctx->max_frames, "frame range");
    cudaSetDevice(ctx->device);
    abub_slot_state& s = ctx->slots[slot];
    ABUB_REQUIRE(ctx, s.h_pinned != nullptr, "slot has no pinned buffer (call abub_pinned_acquire)");
    if (nframes > 0) {
        const size_t off = ((size_t)cam * ctx->max_frames + frame_first) * ctx->npix;
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(s.d_frames + off, s.h_pinned + off, (size_t)nframes * ctx->npix, cudaMemcpyHostToDevice,
                                             ctx->stream));
        s.hist_valid = false;
    }
    return ABUB_OK;
}

extern "C" int abub_event_device_ptr(abub_ctx* ctx, int slot, uint8_t** dev_ptr) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && dev_ptr, "abub_event_device_ptr");
    *dev_ptr = ctx->slots[slot].d_frames;
    return ABUB_OK;
}

extern "C" int abub_event_set_nframes(abub_ctx* ctx, int slot, const int* nframes) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && nframes, "abub_event_set_nframes");
    cudaSetDevice(ctx->device);
    for (int c = 0; c < ctx->ncams; c++) {
        ABUB_REQUIRE(ctx, nframes[c] >= 0 && nframes[c] <= ctx->max_frames, "nframes out of range");
        ctx->slots[slot].nframes[c] = nframes[c];
    }
    return push_nframes(ctx, slot);                                  // counts travel as a kernel argument: nothing to wait for
}

// ---- trigger ------------------------------------------------------------------------
static int require_backgrounds(abub_ctx* ctx) {
    for (int c = 0; c < ctx->ncams; c++)
        ABUB_REQUIRE(ctx, ctx->cam[c].has_background, "background not installed for every camera");
    return ABUB_OK;
}

extern "C" int abub_trigger_hist(abub_ctx* ctx, int slot_first, int nslots) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot_first >= 0 && nslots >= 1 && slot_first + nslots <= ctx->max_slots, "slot range");
    int rc = require_backgrounds(ctx);
    if (rc) return rc;
    cudaSetDevice(ctx->device);
    rc = abub_k1_launch_trigger(ctx, slot_first, nslots);
    if (rc) return rc;
    for (int s = slot_first; s < slot_first + nslots; s++) ctx->slots[s].hist_valid = true;
    return ABUB_OK;
}

extern "C" int abub_trigger_hist_read(abub_ctx* ctx, int slot, int cam, uint32_t* hist) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && cam >= 0 && cam < ctx->ncams && hist, "args");
    cudaSetDevice(ctx->device);
    const size_t n = (size_t)ctx->max_frames * 256;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(hist, ctx->d_hist + ((size_t)slot * ctx->ncams + cam) * n, n * sizeof(uint32_t),
                                         cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ABUB_OK;
}

extern "C" int abub_trigger_hist_write(abub_ctx* ctx, int slot, int cam, const uint32_t* hist, int nframes) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && cam >= 0 && cam < ctx->ncams && hist, "args");
    ABUB_REQUIRE(ctx, nframes >= 0 && nframes <= ctx->max_frames, "nframes");
    cudaSetDevice(ctx->device);
    const size_t n = (size_t)ctx->max_frames * 256;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->d_hist + ((size_t)slot * ctx->ncams + cam) * n, hist, n * sizeof(uint32_t),
                                         cudaMemcpyHostToDevice, ctx->stream));
    ctx->slots[slot].nframes[cam] = nframes;
    int rc = push_nframes(ctx, slot);
    if (rc) return rc;
    ctx->slots[slot].hist_valid = true;
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ABUB_OK;
}

extern "C" int abub_trigger_scan(abub_ctx* ctx, int slot_first, int nslots, const int* start_frames) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot_first >= 0 && nslots >= 1 && slot_first + nslots <= ctx->max_slots, "slot range");
    cudaSetDevice(ctx->device);
    const size_t n = (size_t)nslots * ctx->ncams;
    std::vector<int> st(n, 1);
    if (start_frames) st.assign(start_frames, start_frames + n);
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->d_start + (size_t)slot_first * ctx->ncams, st.data(), n * sizeof(int),
                                         cudaMemcpyHostToDevice, ctx->stream));
    int rc = abub_k2_launch(ctx, slot_first, nslots);
    // the start vector is a local: make sure the copy has been consumed
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return rc;
}

extern "C" int abub_trigger_results(abub_ctx* ctx, int slot_first, int nslots, int* trig, int* status, int* loc_thres,
                                    double* sig) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot_first >= 0 && nslots >= 1 && slot_first + nslots <= ctx->max_slots, "slot range");
    cudaSetDevice(ctx->device);
    const size_t n = (size_t)nslots * ctx->ncams;
    std::vector<int> r(n * 4);
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(r.data(), ctx->d_trig + (size_t)slot_first * ctx->ncams * 4, n * 4 * sizeof(int),
                                         cudaMemcpyDeviceToHost, ctx->stream));
    if (sig)
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(sig, ctx->d_sig + (size_t)slot_first * ctx->ncams * ctx->max_frames,
                                             n * ctx->max_frames * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t i = 0; i < n; i++) {
        if (trig) trig[i] = r[i * 4 + 0];
        if (status) status[i] = r[i * 4 + 1];
        if (loc_thres) loc_thres[i] = r[i * 4 + 2];
    }
    return ABUB_OK;
}

extern "C" int abub_trigger(abub_ctx* ctx, int slot, int cam, int start_frame, int* trig, int* status, int* loc_thres,
                            double* sig) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot >= 0 && slot < ctx->max_slots && cam >= 0 && cam < ctx->ncams, "slot/cam");
    int rc;
    if (!ctx->slots[slot].hist_valid) {
        rc = abub_trigger_hist(ctx, slot, 1);
        if (rc) return rc;
    }
    std::vector<int> st(ctx->ncams, 1), t(ctx->ncams), s(ctx->ncams), l(ctx->ncams);
    st[cam] = start_frame;
    rc = abub_trigger_scan(ctx, slot, 1, st.data());
    if (rc) return rc;
    std::vector<double> sg(sig ? (size_t)ctx->ncams * ctx->max_frames : 0);
    rc = abub_trigger_results(ctx, slot, 1, t.data(), s.data(), l.data(), sig ? sg.data() : nullptr);
    if (rc) return rc;
    if (trig) *trig = t[cam];
    if (status) *status = s[cam];
    if (loc_thres) *loc_thres = l[cam];
    if (sig) memcpy(sig, sg.data() + (size_t)cam * ctx->max_frames, ctx->max_frames * sizeof(double));
    return ABUB_OK;
}

// Streamed FindTriggerFrame: the frames wait in the slots' pinned staging buffers and are sent `chunk_frames` at a
// time; after every chunk K1 runs on the new frame pairs and K2 on what is resident, and a camera whose search is
// decided receives no further frames -- the device-side form of the reference reading frames one by one and
// stopping at the trigger (AnalyzerUnit.cpp:191-213: GetImage inside the loop, early exit at :297-300).
extern "C" int abub_trigger_stream(abub_ctx* ctx, int slot_first, int nslots, const int* nframes, int chunk_frames,
                                   int* trig, int* status, int* loc_thres, int* frames_sent) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, slot_first >= 0 && nslots >= 1 && slot_first + nslots <= ctx->max_slots && nframes, "slot range");
    ABUB_REQUIRE(ctx, chunk_frames >= 1, "chunk_frames");
    int rc = require_backgrounds(ctx);
    if (rc) return rc;
    cudaSetDevice(ctx->device);
    const int nc = ctx->ncams;
    const size_t n = (size_t)nslots * nc;
    std::vector<int> nf(n), sent(n, 0), avail(n, 0), res_t(n, 0), res_s(n, -3), res_l(n, -1), start(n, 1);
    std::vector<char> done(n, 0);
    int nf_max = 0;
    for (int s = 0; s < nslots; s++) {
        abub_slot_state& st = ctx->slots[slot_first + s];
        ABUB_REQUIRE(ctx, st.h_pinned != nullptr, "slot has no pinned buffer (call abub_pinned_acquire)");
        for (int c = 0; c < nc; c++) {
            const int v = nframes[s * nc + c];
            ABUB_REQUIRE(ctx, v >= 0 && v <= ctx->max_frames, "nframes out of range");
            nf[s * nc + c] = st.nframes[c] = v;
            if (v < 5) { done[s * nc + c] = 1; res_s[s * nc + c] = -9; }          // AnalyzerUnit.cpp:122-126
            else if (v > nf_max) nf_max = v;
        }
        st.hist_valid = false;
    }
    int* d_nf = ctx->d_nframes + (size_t)slot_first * nc;
    int* d_av = ctx->d_avail + (size_t)slot_first * nc;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(d_nf, nf.data(), n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->d_start + (size_t)slot_first * nc, start.data(), n * sizeof(int),
                                         cudaMemcpyHostToDevice, ctx->stream));
    const size_t hist_n = n * ctx->max_frames * 256;
    ABUB_CHECK_CUDA(ctx, cudaMemsetAsync(ctx->d_hist + (size_t)slot_first * nc * ctx->max_frames * 256, 0,
                                         hist_n * sizeof(uint32_t), ctx->stream));
    // Two groups of slots take turns on two streams: while one group's K1/K2/read-back runs, the other group's
    // frames are crossing PCIe, so the copy engine never waits for a decision.
    const int ngroups = nslots >= 2 ? 2 : 1;
    if (ngroups == 2 && !ctx->stream2) ABUB_CHECK_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
    if (!ctx->ev_stream) ABUB_CHECK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_stream, cudaEventDisableTiming));
    cudaStream_t gstream[2] = {ctx->stream, ngroups == 2 ? ctx->stream2 : ctx->stream};
    const int gfirst[2] = {0, nslots / 2}, gcount[2] = {ngroups == 2 ? nslots / 2 : nslots, nslots - nslots / 2};
    if (ngroups == 2) {                                           // the tables above must be in place before group 1 starts
        ABUB_CHECK_CUDA(ctx, cudaEventRecord(ctx->ev_stream, ctx->stream));
        ABUB_CHECK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream2, ctx->ev_stream, 0));
    }
    // read-backs must be asynchronous for the two groups to overlap: pinned result table
    if (ctx->h_raw_cap < n * 5) {
        if (ctx->h_raw) cudaFreeHost(ctx->h_raw);
        ctx->h_raw = nullptr;
        ctx->h_raw_cap = 0;
        ABUB_CHECK_CUDA(ctx, cudaHostAlloc((void**)&ctx->h_raw, n * 5 * sizeof(int), cudaHostAllocDefault));
        ctx->h_raw_cap = n * 5;
    }
    int* const raw = ctx->h_raw;
    // the per-camera frame limits go up from pinned memory too: a copy from pageable memory would make the host wait
    // for the frames it has just queued on that stream
    int* const avail_pin = ctx->h_raw + n * 4;
    bool pending[2] = {false, false};                             // results of the group's last chunk not read yet
    bool active[2] = {true, ngroups == 2};
    cudaStream_t const main_stream = ctx->stream;
    auto harvest = [&](int g) -> int {
        if (!pending[g]) return ABUB_OK;
        ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(gstream[g]));
        pending[g] = false;
        for (size_t i = (size_t)gfirst[g] * nc; i < (size_t)(gfirst[g] + gcount[g]) * nc; i++) {
            if (done[i] || raw[i * 4 + 1] == ABUB_STATUS_PENDING) continue;
            done[i] = 1;
            res_t[i] = raw[i * 4 + 0]; res_s[i] = raw[i * 4 + 1]; res_l[i] = raw[i * 4 + 2];
        }
        return ABUB_OK;
    };
    for (int lo = 0; lo < nf_max && (active[0] || active[1]); lo += chunk_frames) {
        const int hi = lo + chunk_frames < nf_max ? lo + chunk_frames : nf_max;
        for (int g = 0; g < ngroups; g++) {
            if (!active[g]) continue;
            if ((rc = harvest(g))) { ctx->stream = main_stream; return rc; }
            cudaStream_t st_g = gstream[g];
            bool any = false;
            for (int s = gfirst[g]; s < gfirst[g] + gcount[g]; s++) {
                abub_slot_state& st = ctx->slots[slot_first + s];
                for (int c = 0; c < nc; c++) {
                    const size_t i = (size_t)s * nc + c;
                    avail[i] = 0;
                    if (done[i]) continue;
                    any = true;
                    const int upto = hi < nf[i] ? hi : nf[i];
                    if (upto > sent[i]) {
                        const size_t off = ((size_t)c * ctx->max_frames + sent[i]) * ctx->npix;
                        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(st.d_frames + off, st.h_pinned + off, (size_t)(upto - sent[i]) * ctx->npix,
                                                             cudaMemcpyHostToDevice, st_g));
                        sent[i] = upto;
                    }
                    avail[i] = sent[i];
                }
            }
            if (!any) { active[g] = false; continue; }
            const size_t o = (size_t)gfirst[g] * nc, cnt = (size_t)gcount[g] * nc;
            memcpy(avail_pin + o, avail.data() + o, cnt * sizeof(int));     // the group's previous copy was harvested above
            ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(d_av + o, avail_pin + o, cnt * sizeof(int), cudaMemcpyHostToDevice, st_g));
            ctx->stream = st_g;                                   // the launchers enqueue on ctx->stream
            rc = abub_k1_launch_trigger_range(ctx, slot_first + gfirst[g], gcount[g], lo, hi, ctx->d_avail, false);
            if (!rc) rc = abub_k2_launch_avail(ctx, slot_first + gfirst[g], gcount[g], ctx->d_avail);
            ctx->stream = main_stream;
            if (rc) return rc;
            ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(raw + o * 4, ctx->d_trig + ((size_t)slot_first * nc + o) * 4, cnt * 4 * sizeof(int),
                                                 cudaMemcpyDeviceToHost, st_g));
            pending[g] = true;
        }
    }
    for (int g = 0; g < ngroups; g++)
        if ((rc = harvest(g))) return rc;
    if (ngroups == 2) {                                           // later work on the main stream follows both groups
        ABUB_CHECK_CUDA(ctx, cudaEventRecord(ctx->ev_stream, ctx->stream2));
        ABUB_CHECK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_stream, 0));
    }
    // leave the result table as abub_trigger_results expects it
    for (size_t i = 0; i < n; i++) { raw[i * 4] = res_t[i]; raw[i * 4 + 1] = res_s[i]; raw[i * 4 + 2] = res_l[i]; raw[i * 4 + 3] = sent[i]; }
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->d_trig + (size_t)slot_first * nc * 4, raw, n * 4 * sizeof(int),
                                         cudaMemcpyHostToDevice, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t i = 0; i < n; i++) {
        if (trig) trig[i] = res_t[i];
        if (status) status[i] = res_s[i];
        if (loc_thres) loc_thres[i] = res_l[i];
        if (frames_sent) frames_sent[i] = sent[i];
    }
    return ABUB_OK;
}

// Resident form of the early exit: the frames are in HBM already, so nothing is gained on the link, but the reference
// stops computing ProcessFrame at the trigger (AnalyzerUnit.cpp:297-300) and so does this call: K1 runs on
// chunk_frames frame pairs at a time, K2 after every round, and a (slot,cam) whose search is decided drops out of the
// following rounds.  No host round trip between the rounds: the per-camera frame limits are updated on the device.
// One search of abub_trigger_early, resumable between rounds: rounds are launched until the frame limit `upto` is reached
// (or the search is through).  d.lo < 0 = not started.
static int trigger_early_rounds(abub_ctx* ctx, abub_deferred_trigger& d, int upto) {
    int rc;
    if (d.lo < 0) {
        d.nf_max = 0;
        for (int s = d.first; s < d.first + d.n; s++) {
            for (int c = 0; c < ctx->ncams; c++) d.nf_max = ctx->slots[s].nframes[c] > d.nf_max ? ctx->slots[s].nframes[c] : d.nf_max;
            ctx->slots[s].hist_valid = false;                         // the tables stop where the searches stopped
        }
        if (d.nf_max == 0) { d.done = true; return abub_k2_launch(ctx, d.first, d.n); }      // status -9 everywhere
        // nothing can be decided on fewer than 5 frames (candidate at i >= 2 plus two look-ahead frames)
        d.lo = 0;
        d.hi = d.chunk < 5 ? 5 : d.chunk;
    }
    while (!d.done && d.lo < upto) {
        if (d.hi > d.nf_max) d.hi = d.nf_max;
        const bool first = d.lo == 0;
        if ((rc = abub_k2_next_avail(ctx, d.first, d.n, d.hi, first))) return rc;
        if ((rc = abub_k1_launch_trigger_range(ctx, d.first, d.n, d.lo, d.hi, ctx->d_avail, first))) return rc;
        if ((rc = abub_k2_launch_avail(ctx, d.first, d.n, ctx->d_avail, true, !first))) return rc;
        d.lo = d.hi;
        d.hi += d.chunk;
        if (d.lo >= d.nf_max) d.done = true;
    }
    return ABUB_OK;
}

static int trigger_early_check(abub_ctx* ctx, int slot_first, int nslots, int chunk_frames) {
    ABUB_REQUIRE(ctx, slot_first >= 0 && nslots >= 1 && slot_first + nslots <= ctx->max_slots, "slot range");
    ABUB_REQUIRE(ctx, chunk_frames >= 1, "chunk_frames");
    return require_backgrounds(ctx);
}

extern "C" int abub_trigger_early(abub_ctx* ctx, int slot_first, int nslots, int chunk_frames) {
    if (!ctx) return ABUB_E_INVALID;
    int rc = trigger_early_check(ctx, slot_first, nslots, chunk_frames);
    if (rc) return rc;
    cudaSetDevice(ctx->device);
    abub_deferred_trigger d{slot_first, nslots, chunk_frames};
    return trigger_early_rounds(ctx, d, 1 << 30);
}

// The same search, registered now and LAUNCHED by the next abub_localize_batch between its own stages (half of the rounds
// once its front-end kernels are queued, the rest behind its last kernel): the host round trips of that localizer call and
// whatever the caller does before it asks for these results are then covered by device work instead of an idle GPU.
// abub_trigger_deferred_flush launches what is still pending (a caller that ends up not calling the localizer).
extern "C" int abub_trigger_early_deferred(abub_ctx* ctx, int slot_first, int nslots, int chunk_frames) {
    if (!ctx) return ABUB_E_INVALID;
    int rc = trigger_early_check(ctx, slot_first, nslots, chunk_frames);
    if (rc) return rc;
    ctx->deferred.push_back(abub_deferred_trigger{slot_first, nslots, chunk_frames});
    return ABUB_OK;
}

// phase -1: the first round only (abub_trigger_deferred_start); phase 0: rounds up to the middle of the longest event;
// phase 1: everything that is left
int abub_run_deferred(abub_ctx* ctx, int phase) {
    int rc = ABUB_OK;
    for (abub_deferred_trigger& d : ctx->deferred) {
        if (d.done) continue;
        int upto = 1 << 30;
        if (phase < 0) {
            upto = 1;                                                    // rounds run while lo < upto: exactly the first
        } else if (phase == 0) {
            int nf = 0;
            for (int s = d.first; s < d.first + d.n; s++)
                for (int c = 0; c < ctx->ncams; c++) nf = ctx->slots[s].nframes[c] > nf ? ctx->slots[s].nframes[c] : nf;
            upto = nf / 2;
        }
        const int r = trigger_early_rounds(ctx, d, upto);
        if (r && !rc) rc = r;
    }
    if (phase > 0) ctx->deferred.clear();
    return rc;
}

// Launches the first round of every registered search at once: device work for the time the caller spends preparing its
// localizer call (item and job tables).
extern "C" int abub_trigger_deferred_start(abub_ctx* ctx) {
    if (!ctx) return ABUB_E_INVALID;
    cudaSetDevice(ctx->device);
    return abub_run_deferred(ctx, -1);
}

extern "C" int abub_trigger_deferred_flush(abub_ctx* ctx) {
    if (!ctx) return ABUB_E_INVALID;
    cudaSetDevice(ctx->device);
    return abub_run_deferred(ctx, 1);
}

// ---- single-frame parity tap --------------------------------------------------------
extern "C" int abub_process_frame(abub_ctx* ctx, int cam, const uint8_t* work, const uint8_t* prev, uint8_t* diff,
                                  uint32_t* hist) {
    if (!ctx) return ABUB_E_INVALID;
    ABUB_REQUIRE(ctx, cam >= 0 && cam < ctx->ncams && work && prev && diff, "abub_process_frame");
    ABUB_REQUIRE(ctx, ctx->cam[cam].has_background, "background not installed");
    cudaSetDevice(ctx->device);
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->d_tmp8[0], work, ctx->npix, cudaMemcpyHostToDevice, ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->d_tmp8[1], prev, ctx->npix, cudaMemcpyHostToDevice, ctx->stream));
    int rc = abub_k1_launch_single(ctx, ctx->d_tmp8[0], ctx->d_tmp8[1], ctx->cam[cam].d_k32, ctx->d_tmp8[2],
                                   ctx->d_hist_tmp, -1);
    if (rc) return rc;
    ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(diff, ctx->d_tmp8[2], ctx->npix, cudaMemcpyDeviceToHost, ctx->stream));
    if (hist)
        ABUB_CHECK_CUDA(ctx, cudaMemcpyAsync(hist, ctx->d_hist_tmp, 256 * sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                             ctx->stream));
    ABUB_CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ABUB_OK;
}
