/* abub_cuda.h -- C ABI of libabub_cuda.so, the B200 (sm_100a) implementation of the
 * AutoBub3hs per-event image hot path.
 *
 * The reference (picoexperiment/AutoBub3hs) has no plugin / FFI layer: its seam is the
 * C++ class surface driven by AnyCamAnalysis()/main() (AutoBubStart3.cpp:67-125,
 * :295-388).  Every entry point below names the reference call it replaces.  Plain
 * pointers and sizes only; int return (0 = ok, negative = ABUB_E_*); no exceptions
 * cross the boundary; buffers are caller-owned unless stated.
 *
 * Threading: one abub_ctx (CUDA stream + scratch arena) per host worker thread
 * (the reference runs one OpenMP thread per event, AutoBubStart3.cpp:342); a ctx is
 * not internally locked.
 *
 * Frame layout everywhere: 8-bit gray, row-major, rows*cols bytes, no padding
 * (cv::Mat CV_8U continuous, as produced by Parser::GetImage, RawParser.cpp:34-47).
 */
#ifndef ABUB_CUDA_H
#define ABUB_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ABUB_OK            0
#define ABUB_E_INVALID    -101   /* bad argument                                   */
#define ABUB_E_CUDA       -102   /* CUDA runtime error (see abub_last_error)       */
#define ABUB_E_NOMEM      -103
#define ABUB_E_CAPACITY   -104   /* caller-provided output buffer too small        */
#define ABUB_E_NCCL       -105

#define ABUB_MAX_TRACK    10     /* NumFramesBubbleTrack, common/CommonParameters.h:5 */

typedef struct abub_ctx abub_ctx;

/* One external contour of one binarised frame, in cv::findContours order
 * (RETR_EXTERNAL, CHAIN_APPROX_TC89_L1; L3Localizer.cpp:264,793).  Mirrors what
 * BubbleImageFrame (bubble/bubble.hpp:22-31) is filled from. */
typedef struct abub_blob {
    int32_t k;          /* 0 = trigger (genesis) frame, 1..10 = post-trigger frame    */
    int32_t x, y, w, h; /* cv::boundingRect of the approximated polygon               */
    int32_t npts;       /* number of polygon vertices                                 */
    int32_t first_px;   /* raster index (y*cols+x) of the component's first pixel     */
    int32_t reserved;
    int64_t sum_x;      /* sum of polygon vertex x / y (centroid fallback,            */
    int64_t sum_y;      /*   L3Localizer.cpp:409-419)                                 */
    double  area;       /* cv::contourArea                                            */
    double  m00, m10, m01; /* cv::moments (spatial, polygon)                          */
} abub_blob;

/* ---- context ----------------------------------------------------------------- */
/* Replaces nothing in the reference (no device there).  max_slots = events resident
 * in HBM at once; each slot holds ncams*max_frames frames. */
int  abub_ctx_create(int device, int rows, int cols, int ncams, int max_frames,
                     int max_slots, abub_ctx** out);
void abub_ctx_destroy(abub_ctx* ctx);
const char* abub_last_error(const abub_ctx* ctx);   /* never NULL */
int  abub_sync(abub_ctx* ctx);                      /* cudaStreamSynchronize */
/* cudaStream_t of the ctx as an opaque pointer (for event timing by the harness). */
void* abub_stream(abub_ctx* ctx);

/* Harness counters (nothing in the reference): kernels of this library launched so far by the ctx, and device time
 * per kernel group.  abub_profile(ctx, 1) resets the table and records CUDA events around every group from then on
 * (on the stream the group is launched on); abub_profile_read fills up to `cap` entries (names are static strings, ms =
 * summed device time, calls = timed groups) and returns the number of groups.  Reading synchronises on the events. */
long long abub_launch_count(const abub_ctx* ctx);
int abub_profile(abub_ctx* ctx, int enable);
int abub_profile_read(abub_ctx* ctx, int cap, const char** names, double* ms, long long* calls);

/* ---- trained background (Trainer public fields, Trainer.hpp:35-54) ------------- */
/* Installs TrainedAvgImage / TrainedSigmaImage / TrainingSetSize for one camera;
 * replaces the per-analyzer deep copy `new Trainer(**TrainedData)`
 * (AnalyzerUnit.cpp:27).  Host pointers, rows*cols bytes each. */
int abub_set_background(abub_ctx* ctx, int cam, const uint8_t* mean, const uint8_t* sigma,
                        int training_set_size);
/* Same, from planes already in device memory (after abub_train_finish). */
int abub_set_background_from_training(abub_ctx* ctx, int cam);

/* ---- event frames -------------------------------------------------------------- */
/* Pinned staging buffer of a slot: ncams*max_frames*rows*cols bytes, camera-major
 * ([cam][frame][rows*cols]).  The host decodes Parser::GetImage output straight into
 * it (RawParser.cpp:39 / ZipParser.cpp:222). */
int abub_pinned_acquire(abub_ctx* ctx, int slot, uint8_t** host_ptr);
/* Async H2D of the slot's staging buffer (cudaMemcpyAsync on the ctx stream);
 * nframes[cam] = CameraFrames.size() of that camera (AnalyzerUnit.cpp:33). */
int abub_event_submit(abub_ctx* ctx, int slot, const int* nframes);
/* Copy one camera's frames from arbitrary host memory (pageable ok) into a slot. */
int abub_event_upload(abub_ctx* ctx, int slot, int cam, const uint8_t* frames, int nframes);
/* One camera of a slot whose frames were decoded straight into the pinned staging buffer (abub_pinned_acquire):
 * asynchronous H2D of its first `nframes` frames on the ctx stream.  The staging region must stay untouched until the
 * next synchronising call (abub_sync, abub_trigger_results, ...). */
int abub_event_submit_cam(abub_ctx* ctx, int slot, int cam, int nframes);
/* Same for a single frame (its region of the staging buffer); the frame count of the camera is not changed. */
int abub_event_upload_frame(abub_ctx* ctx, int slot, int cam, int frame);
/* Frames [frame_first, frame_first + nframes) of one camera from the staging buffer (asynchronous; frame counts unchanged):
 * what the localizer still needs after abub_trigger_stream stopped sending at the trigger -- the reference reads frames
 * trig+1 .. trig+10 in CalculatePostTriggerFrameParams (L3Localizer.cpp:770). */
int abub_event_submit_range(abub_ctx* ctx, int slot, int cam, int frame_first, int nframes);
/* Device address of the slot's [cam][frame][rows*cols] stack (for producers that
 * already live on the device) and the matching frame counts. */
int abub_event_device_ptr(abub_ctx* ctx, int slot, uint8_t** dev_ptr);
int abub_event_set_nframes(abub_ctx* ctx, int slot, const int* nframes);
/* The same for slots [slot_first, slot_first+nslots): nframes[nslots*ncams], one launch per 64 slots' worth of counts. */
int abub_event_set_nframes_range(abub_ctx* ctx, int slot_first, int nslots, const int* nframes);

/* ---- trigger: AnalyzerUnit::FindTriggerFrame (AnalyzerUnit.cpp:119-324) --------- */
/* K1: for every (slot, cam, j>=1) computes the 256-bin histogram of
 * ProcessFrame(frame[j], frame[j-d]) (AnalyzerUnit.cpp:346-377 + calcHist :441-456),
 * d = 2 (1 when TrainingSetSize < 6; frame 0 for j == 1).  Async. */
int abub_trigger_hist(abub_ctx* ctx, int slot_first, int nslots);
/* D2H of one (slot,cam) histogram table: hist[max_frames][256] (row 0 unused = 0). */
int abub_trigger_hist_read(abub_ctx* ctx, int slot, int cam, uint32_t* hist);
/* Parity tap: H2D of one (slot,cam) histogram table (same layout), so that K2 can be driven
 * with hand-built tables (int32-wrap / NaN / inf corner cases of AnalyzerUnit.cpp:478,529). */
int abub_trigger_hist_write(abub_ctx* ctx, int slot, int cam, const uint32_t* hist, int nframes);
/* K2: the significance state machine (calculateSignificanceFrame :435-504 +
 * FindTriggerFrame :119-324) for every (slot,cam) of the range, from
 * start_frames[(slot-slot_first)*ncams+cam] (NULL = 1 everywhere).  A start frame > 1
 * resumes with the history of frames 1..start-1 intact (AnalyzerUnit.cpp:165).  Async. */
int abub_trigger_scan(abub_ctx* ctx, int slot_first, int nslots, const int* start_frames);
/* Which per-frame quantity the state machine of abub_trigger_scan / abub_trigger / abub_trigger_stream thresholds.  The
 * reference compiles ONE of three lines (AnalyzerUnit.cpp:226-228 and :290-292); all three read the same K1 table:
 *   ABUB_METRIC_SIGNIFICANCE          calculateSignificanceFrame (:435-504), entropyThreshold 3.5 (:142) -- the default;
 *   ABUB_METRIC_ENTROPY               calculateEntropyFrame (:386-423): 128-bin Shannon entropy in float, threshold 8e-5 (:140);
 *   ABUB_METRIC_ENTROPY_SIGNIFICANCE  calculateEntropySignificance (:425-433): (E - mean)/sigma over the stored entropies, 1.9 (:141).
 * threshold <= 0 selects the reference's value for the metric.  On the two entropy paths loc_thres stays 3 (:29) and the
 * `sig` array of abub_trigger_results holds the metric.  Takes effect at the next scan. */
#define ABUB_METRIC_SIGNIFICANCE 0
#define ABUB_METRIC_ENTROPY 1
#define ABUB_METRIC_ENTROPY_SIGNIFICANCE 2
int abub_set_trigger_metric(abub_ctx* ctx, int metric, float threshold);
/* Results of the last scan, D2H (synchronises the stream).  Arrays of nslots*ncams:
 * trig = MatTrigFrame (valid when status 0), status = TriggerFrameIdentificationStatus
 * (0, -3, -9), loc_thres = AnalyzerUnit::loc_thres.  sig (may be NULL) =
 * nslots*ncams*max_frames doubles: stored significance of frame j (NaN-safe bit copy;
 * untouched frames hold 0). */
int abub_trigger_results(abub_ctx* ctx, int slot_first, int nslots, int* trig, int* status,
                         int* loc_thres, double* sig);
/* Convenience: FindTriggerFrame for one (slot,cam) with host buffers in the slot
 * (runs K1 for the slot if its table is stale, then K2, then reads back). */
int abub_trigger(abub_ctx* ctx, int slot, int cam, int start_frame, int* trig, int* status,
                 int* loc_thres, double* sig);

/* Streamed form of FindTriggerFrame (start frame 1) for every (slot,cam) of the range.  The frames wait in the
 * slots' pinned staging buffers (abub_pinned_acquire); nframes[(slot-slot_first)*ncams+cam] of them exist.  They are
 * sent chunk_frames at a time and a camera whose search is decided receives no further frames: the device-side
 * form of the reference reading frames one by one and returning at the trigger (AnalyzerUnit.cpp:191-213,
 * :297-300).  Results as abub_trigger_results; frames_sent (may be NULL) = frames transferred per (slot,cam).
 * The frames that were not sent are not resident: upload the slot before calling abub_localize on it. */
int abub_trigger_stream(abub_ctx* ctx, int slot_first, int nslots, const int* nframes, int chunk_frames,
                        int* trig, int* status, int* loc_thres, int* frames_sent);
/* Resident form of the same early exit (frames already in HBM, frame counts set by abub_event_submit / _upload /
 * _set_nframes): K1 over chunk_frames frame pairs at a time, K2 after every round, and a (slot,cam) whose search is
 * decided takes no part in the following rounds -- the reference stops computing ProcessFrame at the trigger
 * (AnalyzerUnit.cpp:297-300).  Start frame 1.  Async; read the results with abub_trigger_results (sig holds the
 * frames that were evaluated).  The histogram tables of the range are left partial: abub_trigger / abub_trigger_hist
 * rebuild them when a later call needs them (a re-trigger from start_frame > 1). */
int abub_trigger_early(abub_ctx* ctx, int slot_first, int nslots, int chunk_frames);
/* The same search for ANOTHER slot range, registered now and launched by the next abub_localize_batch between its own
 * stages (half of the rounds once its front-end kernels are queued, the rest behind its last kernel), on the same
 * stream: the localizer's host round trips, and whatever the caller does before it asks for these results, are covered by
 * device work instead of an idle GPU.  The frame counts of the range must be set before this call and its frames must be
 * resident before the localizer call.  abub_trigger_deferred_flush launches whatever is still pending (for a caller that
 * ends up not calling the localizer); abub_trigger_results then reads the results as after abub_trigger_early.
 * (The reference has one thread per event and no such overlap; AutoBubStart3.cpp:342.) */
int abub_trigger_early_deferred(abub_ctx* ctx, int slot_first, int nslots, int chunk_frames);
int abub_trigger_deferred_flush(abub_ctx* ctx);
/* Launches the FIRST round of every registered search right away (device work while the caller prepares its localizer call). */
int abub_trigger_deferred_start(abub_ctx* ctx);

/* ---- training: Trainer::MakeAvgSigmaImage (Trainer.cpp:221-331) ----------------- */
int abub_train_begin(abub_ctx* ctx, int cam);
/* Gate (Trainer.cpp:279-287): 16-bin histogram of frame1-frame0 on the device, entropy
 * in float on the host; when entropy <= 0.0005 both frames join the statistics.
 * `frames01` = 2*rows*cols host bytes (frame 0 then frame 1).  *entropy / *kept out. */
int abub_train_add_event(abub_ctx* ctx, int cam, const uint8_t* frames01, float* entropy, int* kept);
/* Same with the two frames taken from a resident slot (frames 0 and 1 of `cam`). */
int abub_train_add_slot(abub_ctx* ctx, int cam, int slot, float* entropy, int* kept);
/* Batched forms of the two calls above, for the event loop of Trainer.cpp:262-296 run over many events at
 * once: n events in the loop's order, one gate launch and one device->host read for all of them;
 * entropy[n] / kept[n] out (either may be NULL).  abub_train_add_events takes n consecutive
 * frame pairs (n * 2*rows*cols host bytes, pinned memory makes the upload asynchronous).
 * abub_train_add_slots keeps no copy: the statistics read the kept frames where they lie in the slots, so the
 * slots named here must stay untouched until abub_train_finish / abub_train_partials of this camera has run
 * (a slot may be named more than once, e.g. a ring of resident events). */
int abub_train_add_slots(abub_ctx* ctx, int cam, const int* slots, int n, float* entropy, int* kept);
int abub_train_add_events(abub_ctx* ctx, int cam, const uint8_t* frames01, int n, float* entropy, int* kept);
/* CalculateMeanSigmaImageVector (Trainer.cpp:144-216): exact sequential float32
 * Welford in push order.  mean/sigma host out (may be NULL), *set_size =
 * TrainingSetSize, *status = 0 or -7 (nothing kept, Trainer.cpp:308-313). */
int abub_train_finish(abub_ctx* ctx, int cam, uint8_t* mean, uint8_t* sigma, int* set_size, int* status);
/* Device self-test of the division shortcut inside the Welford kernel (Trainer.cpp:186 divides by the frame count):
 * the `steps`-step reciprocal-and-correct quotient is compared with the IEEE quotient for a = +0 and every float a,
 * 2^-50 <= |a| <= 256 (all that a difference sample - mean can be),
 * and every divisor k_first..k_last; *n_mismatch must come back 0.  example[2] (may be NULL) = {bits of a, k} of
 * one mismatch.  abub_k4_fdiv1_max() = the largest frame count for which the kernel uses steps = 1. */
int abub_selftest_fdiv(abub_ctx* ctx, int steps, int k_first, int k_last, unsigned long long* n_mismatch,
                       unsigned long long* example);
int abub_k4_fdiv1_max(void);
/* Multi-GPU: exact integer partial sums (u32 sum x, u64 sum x^2, kept-frame count) of
 * the frames this rank added; device buffers of rows*cols elements owned by the ctx,
 * exposed so that the host can all-reduce them (NCCL sum) in place. */
int abub_train_partials(abub_ctx* ctx, int cam, uint32_t** d_sum, uint64_t** d_sumsq, int** d_count);
/* ncclAllReduce(sum) of the partials over `nccl_comm` (ncclComm_t passed opaquely;
 * libnccl is dlopen'ed).  SURVEY.md 8(e): the only collective of the path. */
int abub_train_allreduce(abub_ctx* ctx, int cam, void* nccl_comm);
/* mean/sigma from reduced partials in fp64 + flags pixels within eps of a rounding
 * boundary.  n_flagged out. */
int abub_train_finish_partials(abub_ctx* ctx, int cam, uint8_t* mean, uint8_t* sigma,
                               int* set_size, int* status, int* n_flagged);

/* Exact multi-rank finish (after abub_train_partials + abub_train_allreduce on every rank): planes from the
 * reduced partials, then the flagged pixels are recomputed with the reference's sequential float32
 * recurrence from an ncclAllGather of their samples, ranks in order (= push order when ranks own
 * contiguous event ranges).  The result is bit-identical to abub_train_finish on one GPU and is
 * replicated on every rank. */
int abub_train_finish_allreduced(abub_ctx* ctx, int cam, void* nccl_comm, uint8_t* mean, uint8_t* sigma,
                                 int* set_size, int* status, int* n_flagged);

/* ---- on-device PNG decode (replaces cv::imread / cv::imdecode of RawParser.cpp:39, ZipParser.cpp:222
 * for the DAQ's 8-bit greyscale, non-interlaced PNG frames) ----
 * The caller parses the PNG container (signature, IHDR, IDAT chunks) and places the zlib stream -- the concatenated
 * IDAT payloads -- at the START of the frame's own region of the slot's pinned staging buffer (abub_pinned_acquire:
 * region of frame f of camera c begins at ((c * max_frames + f) * rows * cols)).  The library sends the zlen
 * compressed bytes to the device, inflates them (RFC 1951, one warp per frame), undoes the PNG row filters and
 * leaves the pixels in the frame.  ok[i] = 1 when job i decoded to exactly rows*(cols+1) filtered bytes with valid
 * filter types; otherwise the frame is zero-filled and ok[i] = 0 (decode it on the host instead).
 * Frame counts are set separately (abub_event_set_nframes).  Synchronises the stream. */
typedef struct abub_png_job {
    int32_t slot, cam, frame;
    uint32_t zlen;
} abub_png_job;
int abub_event_decode_png(abub_ctx* ctx, const abub_png_job* jobs, int n, int* ok);

/* 8-bit BMP frames keep their rows bottom-up in the file.  The caller reads the pixel array of such a file straight
 * into the frame's pinned region (one positional read, no row copies on the host; what cv::imread does for these
 * files behind Parser::GetImage, RawParser.cpp:34-47, is a row-by-row copy), sends it with the usual upload calls and
 * then has the rows of those frames turned over in place in HBM.  Async on the ctx stream: enqueue it after the
 * uploads of the frames it names.  Requires cols % 4 == 0 (the BMP row pitch equals cols then). */
typedef struct abub_frame_ref {
    int32_t slot, cam, frame;
} abub_frame_ref;
int abub_event_flip_rows(abub_ctx* ctx, const abub_frame_ref* frames, int n);

/* ---- bellows veto: L3Localizer.cpp:292-390 (taken when every genesis contour lies in the bellows mask) ---- */
/* The camera's cam<c>_bellows_template.png as rows*cols grey bytes (L3Localizer.cpp:294-295); NULL clears it
 * ("template not loadable", :297-301). */
int abub_set_bellows_template(abub_ctx* ctx, int cam, const uint8_t* templ, int trows, int tcols);
/* L3Localizer::TrackAFeature (L3Localizer.cpp:473-543) on frame `frame` of (slot, cam): matchTemplate(CCORR_NORMED)
 * over the whole frame, min-max normalisation, first maximum, 3x3 value-weighted centroid in float.
 * x, y = BestMatchLoc; peak_xy[2] (may be NULL) = the integer maximum.  The numerators are exact integers here,
 * OpenCV's come from a float32 DFT: the map agrees to ~1e-6 and the position to ~1e-3 px.
 * ABUB_E_INVALID when the maximum lies on the edge of the map (the reference reads out of bounds there). */
int abub_bellows_track(abub_ctx* ctx, int slot, int cam, int frame, float* x, float* y, int* peak_xy);
/* The CCORR_NORMED map of the last abub_bellows_track ((rows-th+1) x (cols-tw+1) floats, before normalisation);
 * out may be NULL to query the size. */
int abub_bellows_map(abub_ctx* ctx, float* out, int cap, int* map_rows, int* map_cols);
/* Block :303-390 given the two TrackAFeature results (pos_trig / pos_pre = {x, y} as returned, before the
 * one-pixel nudge of :316-323): the template is pasted at both positions, ProcessFrame on their overlap ROI
 * (GetDiffROI, :462-470) gives the bellows-motion model, which is subtracted from overTheSigma of (trig, pre_trig);
 * then TOZERO(loc_thres), Otsu and findContours as before.  out / *n_out = every external contour (k = 0), in
 * findContours order, with no bellows-mask filtering (:374-389).  *status = 0, or -6 where the reference throws
 * (a Rect outside the frame or an empty ROI).  roi_out[4] (may be NULL) = the ROI {x, y, w, h}. */
int abub_localize_bellows(abub_ctx* ctx, int slot, int cam, int trig, int pre_trig, int loc_thres,
                          const float* pos_trig, const float* pos_pre, abub_blob* out, int cap, int* n_out,
                          int* status, int* roi_out);

/* ---- NCCL plumbing for a process-per-GPU host (nothing in the reference: it is one OpenMP process) ----- */
/* ncclCommInitRank with the unique id published by rank 0 through `id_file` (written atomically, removed by
 * rank 0 once the communicator exists).  *comm receives the ncclComm_t. */
int abub_nccl_comm_create(abub_ctx* ctx, const char* id_file, int rank, int world, void** comm);
int abub_nccl_comm_destroy(void* comm);
/* all ranks wait for each other (one-int all-reduce on the ctx stream + stream synchronise) */
int abub_nccl_barrier(abub_ctx* ctx, void* comm);

/* ---- localizer: L3Localizer::LocalizeOMatic (L3Localizer.cpp:881-968) ----------- */
/* For the trigger frame (k=0: CalculateInitialBubbleParams :215-290, frames `trig` and
 * `pre_trig`, TOZERO at loc_thres, Otsu, external contours) and the post-trigger
 * frames k=1..n_post (CalculatePostTriggerFrameParams :764-836) of one (slot,cam):
 * writes the contour records in findContours order, frame after frame, into `out`
 * (capacity `cap`), and n_per_frame[k] for k=0..n_post.  Mask tests, BoxArea filters
 * and tracking stay on the host (L3Localizer.cpp:277-290,392-444,842-863). */
int abub_localize(abub_ctx* ctx, int slot, int cam, int trig, int pre_trig, int loc_thres,
                  int n_post, abub_blob* out, int cap, int* n_per_frame);
/* Batched form (one launch set for many (slot,cam) analyzers -- the shape the GPU wants):
 * items[i] describes one LocalizeOMatic call; blobs of item i, frame k are contiguous in
 * `out`, items in order; n_per_frame[i*(ABUB_MAX_TRACK+1)+k]; *n_total = blobs written. */
typedef struct abub_loc_item {
    int32_t slot, cam, trig, pre_trig, loc_thres, n_post;
} abub_loc_item;
int abub_localize_batch(abub_ctx* ctx, const abub_loc_item* items, int n_items, abub_blob* out,
                        int cap, int* n_per_frame, int* n_total);
/* Debug/parity taps: binary image of frame k of the last abub_localize (rows*cols). */
int abub_localize_binary(abub_ctx* ctx, int k, uint8_t* out);

/* ---- single-frame ops (parity taps; each is one kernel of the path) ------------- */
/* ProcessFrame (AnalyzerUnit.cpp:346-377) of two host frames with the installed
 * sigma of `cam`; diff out (rows*cols), hist out (256, may be NULL). */
int abub_process_frame(abub_ctx* ctx, int cam, const uint8_t* work, const uint8_t* prev,
                       uint8_t* diff, uint32_t* hist);

/* ---- allocation-free gate (no counterpart in the reference) ---------------------- */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The Trainer's entropy
 * gate (the arithmetic of abub_train_add_events: 16-bin histogram of frame1 - frame0, entropy in float, kept when
 * entropy <= 0.0005) for n_pairs frame pairs, with the per-pair bin-count scratch in a file-scope static __device__
 * array of the library (csrc/gate_static.cu) instead of a context arena: the call needs no abub_ctx and calls no
 * cudaMalloc / cudaFree / allocating routine, and it neither waits for the GPU nor copies to the host.
 *   d_pairs    DEVICE in : n_pairs x 2 x rows*cols bytes, frame 0 then frame 1 of each pair, pairs back to back
 *   d_entropy  DEVICE out: n_pairs floats          d_kept DEVICE out: n_pairs ints (may be NULL)
 *   d_hist16   DEVICE out: n_pairs x 16 bin counts (may be NULL)
 *   stream     cudaStream_t the work is queued on (NULL = the default stream); the device is the caller's current one
 *   err        receives a message when the return value is not ABUB_OK (may be NULL)
 * Limit: n_pairs <= ABUB_GATE_STATIC_MAX_PAIRS per call (the scratch is 16 counters per pair, 64 KiB per GPU, held from
 * library load to process exit) -- more returns ABUB_E_CAPACITY; rows*cols < 2^31.
 * All calls on a device share the one scratch array; each call clears the slice it uses on `stream` first.  Calls
 * whose GPU work can overlap (two streams without an ordering, two host threads) race: serialise them. */
#define ABUB_GATE_STATIC_MAX_PAIRS 1024
int abub_gate_entropy_static(const uint8_t* d_pairs, int n_pairs, int rows, int cols, float* d_entropy, int* d_kept,
                             uint32_t* d_hist16, void* stream, char* err, int err_cap);

#ifdef __cplusplus
}
#endif
#endif /* ABUB_CUDA_H */
