// Internal declarations shared by the kernels of libabub_cuda.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>

#include "../../include/abub_cuda.h"

#define ABUB_MAX_CAMS 8
#define ABUB_K1_TILE_W 120          // output columns per warp (30 lanes x 4 px; lanes 0,31 = halo)
#define ABUB_MAX_POST (ABUB_MAX_TRACK + 1)

struct abub_cam_state {
    uint8_t*  d_mean = nullptr;     // TrainedAvgImage
    uint8_t*  d_sigma = nullptr;    // TrainedSigmaImage
    uint32_t* d_k32 = nullptr;      // per pixel (255-t)|(765-t)<<16, t=min(6*sigma,255)  (K1 stage A)
    uint8_t*  d_t8 = nullptr;       // t = min(6*sigma,255)
    int16_t*  d_k16 = nullptr;      // per pixel -(255+t) as s16 (TMA trigger kernel stage A)
    int16_t*  d_nt16 = nullptr;     // per pixel -t as s16 (localizer post-trigger kernel: relu(|a - avg| - t) in one VIADDMNMX)
    int training_set_size = 0;
    bool has_background = false;
    // training: kept frames live in device chunks until abub_train_finish
    std::vector<uint8_t*> chunks;   // each holds TRAIN_CHUNK frames
    std::vector<const uint8_t*> kept;   // device pointers in push order (Trainer.cpp:291-292)
    const uint8_t** d_table = nullptr;  // device copy of `kept`
    int table_cap = 0;
    int n_stored = 0;               // frames written into chunks (kept only)
    uint32_t* d_sum = nullptr;      // exact integer partials (multi-GPU)
    uint64_t* d_sumsq = nullptr;
    int*      d_count = nullptr;    // [0]=kept frames (this rank, then reduced), [1]=flag count
    uint8_t*  d_flag = nullptr;
    bool training = false;
    // bellows veto (k8_bellows.cu): cam<c>_bellows_template.png, rows padded to a 16-byte pitch
    uint8_t* d_templ = nullptr;
    int templ_rows = 0, templ_cols = 0, templ_pitch = 0;
    double templ_sum2 = 0;
};

#define ABUB_TRAIN_CHUNK 32

struct abub_slot_state {
    uint8_t* d_frames = nullptr;    // [cam][frame][rows*cols]
    uint8_t* h_pinned = nullptr;    // lazily allocated staging
    int nframes[ABUB_MAX_CAMS] = {0};
    bool hist_valid = false;
};

// Per-kernel-group device timing (abub_profile / abub_profile_read): CUDA events recorded on the stream a group is
// launched on, resolved when the table is read.  Off by default (no events are recorded then).
enum {
    ABUB_PROF_K1 = 0,   // k1_trigger_tma_kernel + k1_trigger_kernel (border rows) + k1_fill_bin0_kernel
    ABUB_PROF_K2,       // k2_trigger_scan_kernel (+ k2_next_avail_kernel)
    ABUB_PROF_K5A,      // localizer genesis frames: k5_genesis_tma_kernel + k1_genesis_border_kernel (or k1_genesis_jobs_kernel)
    ABUB_PROF_K5B,      // localizer post-trigger frames: k5_post_tma_kernel (or k5_post_stream_kernel / k5_post_kernel)
    ABUB_PROF_K5C,      // k5_otsu_kernel + flag fill
    ABUB_PROF_K6,       // k6_candidates_*_kernel
    ABUB_PROF_SORT,     // cub radix sort / scans (library kernels)
    ABUB_PROF_K7A,      // k7_trace1_kernel
    ABUB_PROF_K7B,      // k7_trace2_kernel
    ABUB_PROF_K7C,      // k7_external_kernel + k7_emit_kernel
    ABUB_PROF_K3,       // k3_gate_*_kernel
    ABUB_PROF_K4,       // k4_welford*_kernel / partials / finish / fixup
    ABUB_PROF_N
};
struct abub_prof_state {
    bool on = false;
    std::vector<cudaEvent_t> pool;
    size_t used = 0;
    struct Rec { int id; size_t e0, e1; };
    std::vector<Rec> recs;
    double ms[ABUB_PROF_N] = {0};
    long long calls[ABUB_PROF_N] = {0};
};

// a trigger search registered with abub_trigger_early_deferred (ctx.cu)
struct abub_deferred_trigger { int first, n, chunk; int lo = -1, hi = 0, nf_max = 0; bool done = false; };

struct abub_ctx {
    long long n_launches = 0;           // kernels of this library launched so far (ABUB_L at every launch site)
    abub_prof_state prof;
    int device = 0, rows = 0, cols = 0, ncams = 0, max_frames = 0, max_slots = 0;
    size_t npix = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;     // second stream of abub_trigger_stream (copy/compute overlap between slot groups)
    cudaEvent_t ev_stream = nullptr;
    int* h_raw = nullptr;               // pinned result table of abub_trigger_stream
    size_t h_raw_cap = 0;
    std::string err;
    abub_cam_state cam[ABUB_MAX_CAMS];
    std::vector<abub_slot_state> slots;
    // trigger tables
    uint32_t* d_hist = nullptr;     // [slot][cam][frame][256]
    int*      d_nframes = nullptr;  // [slot][cam]
    int*      d_start = nullptr;    // [slot][cam]
    int*      d_avail = nullptr;    // [slot][cam] frames resident so far (abub_trigger_stream)
    int*      d_trig = nullptr;     // [slot][cam] x {trig,status,loc_thres,scanned}
    double*   d_sig = nullptr;      // [slot][cam][frame]
    void*     d_k2_state = nullptr; // [slot][cam] K2State: searches of abub_trigger_early that wait for the next round (k2_trigger_scan.cu)
    int       trig_metric = 0;      // ABUB_METRIC_* used by K2 (abub_set_trigger_metric)
    float     trig_thr = 3.5f;      // entropyThreshold before the TrainingSetSize < 6 scaling (AnalyzerUnit.cpp:140-142)
    int*      d_tss = nullptr;      // [cam] TrainingSetSize
    uint32_t** d_k32_ptrs = nullptr;  // [cam] device array of K32 plane pointers
    uint8_t**  d_slot_ptrs = nullptr; // [slot] device array of slot base pointers
    // scratch for single-frame ops / training
    uint8_t*  d_tmp8[4] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t* d_hist_tmp = nullptr; // 16 x 256
    const uint8_t** d_gate_tab = nullptr;   // batched entropy gate: [n][2] frame pointers ...
    uint32_t* d_gate_hist = nullptr;        // ... and [n][16] histograms
    int gate_cap = 0;
    void*     loc = nullptr;        // localizer arena (k5_localize.cu)
    void*     bel = nullptr;        // bellows-veto arena (k8_bellows.cu)
    void*     png = nullptr;        // on-device PNG decode arena (k9_png.cu)
    int*      d_post_list = nullptr; // indices of the post-trigger jobs of a localizer chunk (k5_post_tma.cu)
    int*      h_post_list = nullptr; // its pinned staging copy; post_ev = the last upload from it
    cudaEvent_t post_ev = nullptr;
    std::vector<abub_deferred_trigger> deferred;   // pending deferred trigger searches
    void*     fix_buf[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // grow-only scratch of the multi-GPU fix-up (nccl_glue.cu)
    size_t    fix_cap[6] = {0, 0, 0, 0, 0, 0};
    int*      d_gen_list = nullptr;  // the same for the genesis jobs (k5_genesis_tma.cu)
    int*      h_gen_list = nullptr;
    cudaEvent_t gen_ev = nullptr;
    int       gen_cap = 0;
    int       post_cap = 0;
    void*     d_tmaps = nullptr;    // CUtensorMap[max_slots + 3*ncams]: frames of a slot, then per camera K16, TrainedAvgImage, -t
                                    // (k1_trigger_tma.cu builds them; null = TMA paths off)
};

#define ABUB_L(ctx) ((void)++(ctx)->n_launches)

// RAII scope around the launches of one kernel group; a no-op unless abub_profile(ctx, 1) was called
struct abub_prof_scope {
    abub_ctx* c; int id; size_t e0 = 0; bool live = false;
    static size_t ev(abub_ctx* c) {
        abub_prof_state& p = c->prof;
        if (p.used == p.pool.size()) { cudaEvent_t e; if (cudaEventCreate(&e) != cudaSuccess) return (size_t)-1; p.pool.push_back(e); }
        return p.used++;
    }
    abub_prof_scope(abub_ctx* ctx, int id_) : c(ctx), id(id_) {
        if (!c->prof.on) return;
        e0 = ev(c);
        if (e0 == (size_t)-1) return;
        cudaEventRecord(c->prof.pool[e0], c->stream);
        live = true;
    }
    ~abub_prof_scope() {
        if (!live) return;
        const size_t e1 = ev(c);
        if (e1 == (size_t)-1) return;
        cudaEventRecord(c->prof.pool[e1], c->stream);
        c->prof.recs.push_back({id, e0, e1});
    }
};

#define ABUB_CHECK_CUDA(ctx, call)                                                        \
    do {                                                                                  \
        cudaError_t e__ = (call);                                                         \
        if (e__ != cudaSuccess) {                                                         \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);             \
            return ABUB_E_CUDA;                                                           \
        }                                                                                 \
    } while (0)

#define ABUB_REQUIRE(ctx, cond, msg)                                                      \
    do {                                                                                  \
        if (!(cond)) {                                                                    \
            (ctx)->err = std::string("invalid argument: ") + (msg);                       \
            return ABUB_E_INVALID;                                                        \
        }                                                                                 \
    } while (0)

// ---- kernel launchers (one per .cu) ------------------------------------------------
// K1 (k1_processframe.cu)
int abub_k1_build_k32(abub_ctx* ctx, int cam);
int abub_k1_launch_trigger(abub_ctx* ctx, int slot_first, int nslots);
int abub_k1_launch_trigger_range(abub_ctx* ctx, int slot_first, int nslots, int j_first, int j_last, const int* d_limit,
                                 bool zero_hist);
int abub_k1_launch_single(abub_ctx* ctx, const uint8_t* d_a, const uint8_t* d_b, const uint32_t* d_k32,
                          uint8_t* d_out, uint32_t* d_hist, int tozero_thres);
// K1 TMA trigger path (k1_trigger_tma.cu)
bool abub_k1_tma_supported(const abub_ctx* ctx);
int abub_k1_tma_init(abub_ctx* ctx);
int abub_k1_build_k16(abub_ctx* ctx, int cam);
int abub_k1_launch_trigger_tma(abub_ctx* ctx, int slot_first, int nslots, int j_first, int j_last, const int* d_limit);
// K2 (k2_trigger_scan.cu)
int abub_k2_launch(abub_ctx* ctx, int slot_first, int nslots);
// streamed form: d_avail[slot*ncams+cam] frames are resident; a search that needs a frame beyond them reports
// ABUB_STATUS_PENDING
int abub_k2_launch_avail(abub_ctx* ctx, int slot_first, int nslots, const int* d_avail, bool keep_state = false, bool resume = false);
#define ABUB_STATUS_PENDING (-100)
// abub_trigger_early: d_avail (and d_start on the first round) of the range for the next round of frames
int abub_k2_next_avail(abub_ctx* ctx, int slot_first, int nslots, int next, bool first_round);
// K3/K4 (k34_trainer.cu)
int abub_k3_gate_hist16(abub_ctx* ctx, const uint8_t* d_f0, const uint8_t* d_f1, uint32_t* d_hist16);
int abub_k3_gate_batch(abub_ctx* ctx, const uint8_t* const* d_pairs, int n, uint32_t* d_hist16);
int abub_k4_welford(abub_ctx* ctx, const uint8_t* const* d_frame_table, int n, uint8_t* d_mean, uint8_t* d_sigma);
int abub_k4_partials(abub_ctx* ctx, const uint8_t* const* d_frame_table, int n, uint32_t* d_sum, uint64_t* d_sumsq, bool aligned16);
int abub_k4_finish_from_partials(abub_ctx* ctx, const uint32_t* d_sum, const uint64_t* d_sumsq, int n, uint8_t* d_mean,
                                 uint8_t* d_sigma, int* d_nflag, uint8_t* d_flag);
int abub_k4_fixup(abub_ctx* ctx, const uint8_t* const* d_frame_table, int n, const uint8_t* d_flag, uint8_t* d_mean,
                  uint8_t* d_sigma);
// one binarisation job of the localizer (a frame of an item)
struct abub_loc_job {
    const uint8_t* a;        // genesis: trigger frame ; post: frame[trig+k]
    const uint8_t* b;        // genesis: pre-trigger frame ; post: TrainedAvgImage
    const uint32_t* k32;     // genesis: K32 plane of the camera
    const uint8_t* t8;       // post: min(6 sigma, 255)
    const int16_t* nt16;     // post: -min(6 sigma, 255) as s16
    uint8_t* tz;             // out: TOZERO'd image (rows*cols)
    uint8_t* segflags;       // [rows][ceil(cols/240)] per 240-px row segment of tz: bit0 = holds a non-zero pixel now,
                             // bit1 = memory not known to be all zero (k5_localize.cu, sparse TOZERO frames)
    uint32_t* hist;          // out: 256-bin histogram of tz
    int tozero;              // loc_thres (genesis) or 3 (post)
    int genesis;
    int k;                   // 0 = genesis frame, 1.. = post-trigger frame
    int item;                // index of the abub_loc_item this job belongs to
    int slot, cam, zframe;   // jobs on the TMA paths: event slot, camera, cam * max_frames + frame of `a`
    int zb;                  // genesis jobs on the TMA path: cam * max_frames + the comparison (pre-trigger) frame
};
int abub_k1_launch_genesis_jobs(abub_ctx* ctx, const abub_loc_job* d_jobs, int njobs);
// K5a TMA path (k5_genesis_tma.cu): rows [2, rows-2) of the chunk's genesis jobs, sparse store + segment flags; builds
// ctx->d_gen_list (indices of the genesis jobs) and returns their number
bool abub_k5a_tma_supported(const abub_ctx* ctx);
int abub_k5a_launch_tma(abub_ctx* ctx, const abub_loc_job* d_jobs, int njobs, int* n_genesis);
// K5b TMA path (k5_post_tma.cu): post-trigger jobs of a chunk; false = shape not supported, use the streaming kernel
bool abub_k5b_tma_supported(const abub_ctx* ctx);
int abub_k5b_launch_tma(abub_ctx* ctx, const abub_loc_job* d_jobs, int njobs);
const abub_loc_job* abub_loc_host_jobs(abub_ctx* ctx);      // host copy of the chunk's job table (k5_localize.cu)

// K5-K7 (k5_localize.cu)
int abub_run_deferred(abub_ctx* ctx, int phase);           // ctx.cu: launches pending deferred trigger searches (phase 0 / 1)
int abub_loc_init(abub_ctx* ctx);
void abub_loc_free(abub_ctx* ctx);
int abub_loc_run_prepared(abub_ctx* ctx, const uint8_t* d_over, int tozero, abub_blob* out, int cap, int* n_out);
// K8 (k8_bellows.cu)
void abub_bel_free(abub_ctx* ctx);
// K9 (k9_png.cu)
void abub_png_free(abub_ctx* ctx);
