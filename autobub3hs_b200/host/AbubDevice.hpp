// AbubDevice -- the one piece of the host side that has no counterpart in the reference: it owns
// the abub_ctx (libabub_cuda.so, include/abub_cuda.h) that Trainer / AnalyzerUnit / L3Localizer
// call instead of OpenCV, and the event slots in HBM.
//
//   * Configure() is called once by main (or a test) with the device ordinal and capacities; the
//     context itself is created when the first frame tells the image size.
//   * An event occupies one slot ([cam][frame][rows*cols] in HBM).  Analyzers of the same event share
//     the slot (AcquireSlot is reference counted by event key); each camera's frames are decoded
//     straight into the slot's pinned staging buffer and sent with cudaMemcpyAsync.
//   * PrefetchEvents() lets the driver loop stage a whole batch of events (decode on every host core,
//     one K1 launch for the batch) before the per-camera objects are built; analyzers then find their
//     frames and histograms resident.
//   * The ctx is not internally locked (one CUDA stream): every C-ABI call goes through lock().  The slot table and
//     the mask cache have their own short lock, and an analyzer reads the results of its batch from the slot caches
//     without any lock (they are complete before the batch is handed to the analyzers), so that the per-camera objects
//     of one batch run on the host cores while the device works on the next batch.
#ifndef ABUB_DEVICE_HPP
#define ABUB_DEVICE_HPP

#include <map>
#include <mutex>
#include <atomic>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/abub_cuda.h"
#include "Parser.hpp"

struct AbubError : public std::runtime_error {
    explicit AbubError(const std::string& what) : std::runtime_error(what) {}
};

struct AbubSlot {
    std::string key;                 // "<run folder>|<event id>"
    int refs = 0;
    bool loaded[8] = {false};        // frames of this camera are in HBM
    int nframes[8] = {0};
    std::vector<int> bad[8];         // frame indices that failed to load/decode
    int resident_upto[8] = {0};      // frames [0, resident_upto) are in HBM (the streamed search stops sending at the decision)
    // results computed for the whole batch by PrefetchEvents (one K2 launch, one localizer call); an analyzer that asks
    // for exactly this work takes it from here instead of launching its own
    bool trig_valid[8] = {false};
    int trig[8] = {0}, trig_status[8] = {0}, trig_loc_thres[8] = {0};
    std::vector<double> sig[8];      // stored significance per frame
    bool loc_valid[8] = {false};
    int loc_trig[8] = {0}, loc_pre[8] = {0}, loc_thres[8] = {0}, loc_npost[8] = {0};
    int loc_npf[8][ABUB_MAX_TRACK + 1] = {{0}};
    std::vector<abub_blob> loc_blobs[8];
    // the trigger search of this slot's CURRENT frames was started ahead of time (AbubDevice::PrefetchTrigger) for these counts
    bool trig_prelaunched = false;
    int pre_nframes[8] = {0};
};

class AbubDevice {
public:
    static AbubDevice& Get();

    void Configure(int device, int ncams, int max_frames, int max_slots);
    // Which of the reference's three trigger quantities K2 thresholds (AnalyzerUnit.cpp:226-228; ABUB_METRIC_*); call before
    // the first frame is staged.  threshold <= 0: the reference's value for the metric (:140-142).
    void SetTriggerMetric(int metric, float threshold) { trigger_metric_ = metric; trigger_threshold_ = threshold; }
    bool configured() const { return ncams_ > 0; }
    int ncams() const { return ncams_; }
    int max_frames() const { return max_frames_; }
    int max_slots() const { return max_slots_; }
    int rows() const { return rows_; }
    int cols() const { return cols_; }

    // Creates the context on first use (rows x cols from the first decoded frame); throws AbubError
    // if a later frame has another size or the device cannot be opened.
    abub_ctx* Context(int rows, int cols);
    abub_ctx* ctx() const { return ctx_; }
    std::recursive_mutex& lock() { return mu_; }
    void Check(int rc, const char* what);       // throws AbubError(abub_last_error) when rc != 0

    // Slot of an event (allocated on first request, freed when the last holder releases it).
    int AcquireSlot(const std::string& key);
    void ReleaseSlot(const std::string& key);
    // Binds an event key to a given slot (whose frames a producer has filled); the slot must be free.
    int BindSlot(const std::string& key, int slot);
    bool SlotInUse(int slot);
    AbubSlot& slot(int s) { return slots_[s]; }

    // Decodes (in parallel over frames) and uploads the frames of one camera of an event into its slot;
    // no-op when already resident.  Frame names as listed by Parser::ParseAndSortFramesInFolder.
    void LoadCamera(int slot, int cam, Parser* parser, const std::string& event_id,
                    const std::vector<std::string>& frames);
    // Stages a batch: for every event all cameras are loaded, then one K1 launch covers the batch.
    // keys[i] is the slot key of events[i] (the one its analyzers use).  Returns the slots; every key holds one
    // reference that the caller releases with ReleaseSlot when the event is done.
    // training_set_sizes ([ncams], may be NULL): with it the trigger scan and the localizer also run for the whole
    // batch (the pre-trigger frame offset depends on TrainingSetSize, L3Localizer.cpp:919-934).
    std::vector<int> PrefetchEvents(Parser* parser, const std::vector<std::string>& keys, const std::vector<std::string>& events,
                                    const int* training_set_sizes = nullptr);

    // The device work of a batch whose frames are resident: FindTriggerFrame(1) of every (event, camera), LocalizeOMatic
    // of every triggered one (when training_set_sizes is given); results land in the slots' caches.
    // next_first / next_n / next_nframes (optional): the slots of the batch that will be analysed after this one, frames already
    // in HBM: their trigger search is queued inside this batch's localizer call (abub_trigger_early_deferred), so the GPU has
    // work while the host goes through the localizer's round trips and the bookkeeping between two batches
    void AnalyzeBatch(const std::vector<int>& slots, const int* training_set_sizes, int next_first = -1, int next_n = 0,
                      const int* next_nframes = nullptr);
    bool PrefetchTrigger(int slot_first, int nslots, const int* nframes);
    // Same for slots whose frames wait in the pinned staging buffers: frames cross PCIe chunk_frames at a time and only as
    // far as the search and the localizer need them.  Returns the number of frames sent.
    long long StreamBatch(const std::vector<int>& slots, const int* training_set_sizes, int chunk_frames);
    // Declares the frames of a slot present: in_hbm = a producer wrote them to abub_event_device_ptr; otherwise they are
    // in the pinned staging buffer (abub_pinned_acquire) and StreamBatch / EnsureAllResident send them.
    void MarkResident(int slot, const int* nframes, bool in_hbm, bool push = true);
    void PushFrameCounts(int slot_first, int nslots, const int* nframes);
    void EnsureAllResident(int slot);
    // frame pairs per round of the resident early exit (abub_trigger_early); 0 = full histogram table
    void SetEarlyChunk(int chunk) { early_chunk_ = chunk; }
    static int NPost(int trig, int nframes);                 // L3Localizer.cpp:944-954

    // cam<c>_bellows_template.png of `mask_dir` (L3Localizer.cpp:294-295), decoded and handed to the device on first
    // use; false when the file cannot be read or decoded ("template not loadable", :297-301).
    bool EnsureBellowsTemplate(int cam, const std::string& mask_dir);
    // cam<c>_mask.bmp / cam<c>_bellows_mask.bmp of `mask_dir`, read once per run (the reference re-reads the file in
    // every isInMask call, L3Localizer.cpp:983-984); nullptr when unreadable.
    const Image8* Mask(int cam, bool bellows, const std::string& mask_dir);

    void Shutdown();
    // Frames decoded on the device so far (PNG, K9) / on the host cores.
    long n_gpu_decoded = 0, n_cpu_decoded = 0;

    // wall-clock seconds spent in the ingest stages (printed by abub3hs when ABUB_TIMING is set)
    double t_read = 0, t_decode = 0, t_upload = 0, t_hist = 0, t_scan = 0, t_loc = 0;
    double t_submit = 0, t_flip = 0, t_tail = 0;     // parts of t_upload: upload calls, row-flip calls, end of StageBatch

private:
    AbubDevice() {}
    int device_ = 0, ncams_ = 0, max_frames_ = 0, max_slots_ = 0, rows_ = 0, cols_ = 0;
    abub_ctx* ctx_ = nullptr;
    std::recursive_mutex mu_;         // the device: every C-ABI call on ctx_ (held for the whole of a batch's device work)
    std::mutex tab_mu_;               // the slot table and the mask cache (short; may be taken while mu_ is held, never the other way round)
    std::vector<AbubSlot> slots_;
    // stages every camera of the given (slot, event) pairs: all frames are read on all cores; 8-bit grey PNG frames are
    // inflated and unfiltered on the device (abub_event_decode_png), anything else is decoded on the host
    void StageBatch(Parser* parser, const std::vector<int>& slots, const std::vector<std::string>& events);
    void LocalizeBatch(const std::vector<int>& slots, const int* training_set_sizes);
    std::vector<abub_blob> loc_blobs_;
    int early_chunk_ = 5;
    int trigger_metric_ = 0;
    float trigger_threshold_ = 0.f;
    std::map<std::string, int> by_key_;
    std::map<std::string, bool> templates_;     // "<mask dir>|<cam>" -> loadable
    std::map<std::string, Image8> masks_;       // "<mask dir>|<cam>|<bellows>" -> image (empty = unreadable)
};

#endif
