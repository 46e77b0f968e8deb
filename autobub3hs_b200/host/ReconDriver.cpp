#include "ReconDriver.hpp"

#include <omp.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <iostream>

#include "AbubDevice.hpp"
#include "L3Localizer.hpp"

static inline double wall() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

static void advance_cursor() {                     // common/UtilityFunctions.cpp:109-115
    static int pos = 0;
    static const char cursor[4] = {'/', '-', '\\', '|'};
    printf("%c\b", cursor[pos]);
    fflush(stdout);
    pos = (pos + 1) % 4;
}

static bool g_progress = true;

int AnyCamAnalysis(std::string EventID, std::string ImgDir, int camera, bool nonStopPref, Trainer** TrainingData,
                   OutputWriter** P60Output, std::string out_dir, int actualEventNumber, AnalyzerUnit** AGeneric) {
    (void)EventID; (void)ImgDir; (void)TrainingData;
    OutputWriter* Pico60Writer = *P60Output;
    AnalyzerUnit* AnalyzerCGeneric = *AGeneric;
    try {
        do {
            AnalyzerCGeneric->FindTriggerFrame(nonStopPref, AnalyzerCGeneric->MatTrigFrame + 1);
            if (AnalyzerCGeneric->okToProceed) {
                AnalyzerCGeneric->LocalizeOMatic(out_dir);
                if (AnalyzerCGeneric->okToProceed) {
                    Pico60Writer->stageCameraOutput(AnalyzerCGeneric->BubbleList, camera, AnalyzerCGeneric->MatTrigFrame,
                                                    actualEventNumber);
                } else {
                    Pico60Writer->stageCameraOutputError(camera, -8, actualEventNumber);
                    break;
                }
            } else {
                Pico60Writer->stageCameraOutputError(camera, AnalyzerCGeneric->TriggerFrameIdentificationStatus,
                                                     actualEventNumber);
                break;
            }
        } while (AnalyzerCGeneric->BubbleList.size() == 0);
    } catch (std::exception& e) {
        /* per-camera failure: -6 (a CUDA error inside an event lands here as well) */
        std::cout << e.what() << '\n';
        Pico60Writer->stageCameraOutputError(camera, -6, actualEventNumber);
    }
    if (g_progress) advance_cursor();
    return 0;
}

void DetectStagedBatch(DetectRun& run, const std::vector<std::string>& chunk, const std::vector<std::string>& keys, std::ofstream& out) {
    g_progress = run.progress;
    const int n = (int)chunk.size();
    const double t0 = wall();
    double t_write = 0;
    std::vector<Trainer*>& Trainers = *run.Trainers;
    /* events in parallel, cameras in series, rows committed in event order (AutoBubStart3.cpp:342, :360-365, :379-383) */
    const int nthreads = run.threads > 0 ? run.threads : omp_get_max_threads();
#pragma omp parallel for ordered schedule(dynamic, 1) num_threads(nthreads)
    for (int ci = 0; ci < n; ci++) {
        const std::string& ev = chunk[ci];
        OutputWriter* EventOutput = new OutputWriter(run.out_dir, run.run_number, run.frameOffset, run.numCams);
        std::string imageDir = run.eventDir + ev + run.imageFolder;
        int actualEventNumber = atoi(ev.c_str());
        if (run.progress) {
#pragma omp critical(abub_progress)
            {
                printf("\rProcessing event: %s / %d  ... ", ev.c_str(), run.nEvents - 1);
                advance_cursor();
            }
        }
        std::vector<AnalyzerUnit*> Analyzers;
        for (int icam = 0; icam < run.numCams; icam++) {
            Parser* tp = run.FileParser->clone();
            Analyzers.push_back(new L3Localizer(ev, imageDir, icam, run.debug_mode / 100 ? false : true, &Trainers[icam], run.mask_dir, tp));
            AnyCamAnalysis(ev, imageDir, icam, run.debug_mode % 10 ? false : true, &Trainers[icam], &EventOutput, run.out_dir,
                           actualEventNumber, &Analyzers[icam]);
        }
        std::string rows;
        EventOutput->appendCameraOutput(rows);
#pragma omp ordered
        {
            const double tw = wall();
            out.write(rows.data(), (std::streamsize)rows.size());
            out.flush();                              /* committed per event, as :379-383 */
            t_write += wall() - tw;
        }
        delete EventOutput;
        for (int icam = 0; icam < run.numCams; icam++) delete Analyzers[icam];
        AbubDevice::Get().ReleaseSlot(keys[ci]);
    }
    run.T_write += t_write;
    run.T_analyze += wall() - t0 - t_write;
}

void DetectPipeline::start(DetectRun* run, std::ofstream* out) {
    stop();
    run_ = run; out_ = out;
    stop_ = false; inflight_ = 0; err_ = nullptr;
    q_.clear();
    th_ = std::thread(&DetectPipeline::loop, this);
    started_ = true;
}

void DetectPipeline::loop() {
    for (;;) {
        Job job;
        {
            std::unique_lock<std::mutex> lk(m_);
            cv_.wait(lk, [&] { return stop_ || !q_.empty(); });
            if (q_.empty()) return;
            job = std::move(q_.front());
            q_.pop_front();
        }
        try {
            DetectStagedBatch(*run_, job.chunk, job.keys, *out_);
        } catch (...) {
            std::lock_guard<std::mutex> lk(m_);
            if (!err_) err_ = std::current_exception();
        }
        {
            std::lock_guard<std::mutex> lk(m_);
            inflight_--;
        }
        cv_.notify_all();
    }
}

void DetectPipeline::submit(std::vector<std::string> chunk, std::vector<std::string> keys) {
    std::unique_lock<std::mutex> lk(m_);
    cv_.wait(lk, [&] { return inflight_ == 0; });
    if (err_) { std::exception_ptr e = err_; err_ = nullptr; std::rethrow_exception(e); }
    q_.push_back(Job{std::move(chunk), std::move(keys)});
    inflight_++;
    lk.unlock();
    cv_.notify_all();
}

void DetectPipeline::drain() {
    std::unique_lock<std::mutex> lk(m_);
    cv_.wait(lk, [&] { return inflight_ == 0; });
    if (err_) { std::exception_ptr e = err_; err_ = nullptr; std::rethrow_exception(e); }
}

void DetectPipeline::stop() {
    if (!started_) return;
    {
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [&] { return inflight_ == 0; });
        stop_ = true;
    }
    cv_.notify_all();
    if (th_.joinable()) th_.join();
    started_ = false;
}

void MemParser::ParseAndSortFramesInFolder(std::string EventID, int Camera, std::vector<std::string>& Contents) {
    Contents.clear();
    int n = 0;
    {
        std::lock_guard<std::mutex> g(tab_->mu);
        auto it = tab_->nframes.find(EventID);
        if (it == tab_->nframes.end() || Camera < 0 || Camera >= (int)it->second.size()) return;
        n = it->second[Camera];
    }
    char name[64];
    for (int f = 0; f < n; f++) {
        snprintf(name, sizeof(name), "cam%d_image%d.png", Camera, frame_offset_ + f);
        Contents.push_back(name);
    }
}

void MemParser::GetEventDirLists(std::vector<std::string>& EventList) {
    std::lock_guard<std::mutex> g(tab_->mu);
    EventList.assign(tab_->order.begin(), tab_->order.end());
}

void MemParser::SetEvent(const std::string& event, const int* nframes, int ncams, size_t keep) {
    std::lock_guard<std::mutex> g(tab_->mu);
    auto it = tab_->nframes.find(event);
    if (it != tab_->nframes.end()) { it->second.assign(nframes, nframes + ncams); return; }
    tab_->nframes.emplace(event, std::vector<int>(nframes, nframes + ncams));
    tab_->order.push_back(event);
    while (tab_->order.size() > keep) {
        tab_->nframes.erase(tab_->order.front());
        tab_->order.pop_front();
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Harness entry points (libabubhost.so, loaded with ctypes by bench.py and the tests): the full recon of events whose
// frames are produced in memory -- in HBM slots (abub_event_device_ptr) or in the slots' pinned staging buffers
// (abub_pinned_acquire) -- through the same host classes and the same detect loop as the abub3hs binary.
// ---------------------------------------------------------------------------------------------------------------------
namespace {
struct HostRun {
    DetectRun run;
    std::vector<Trainer*> trainers;
    MemParser* parser = nullptr;
    std::ofstream out;
    std::string err;
    DetectPipeline pipe;
    int sub_batch = 32;
    double base[5] = {0, 0, 0, 0, 0};           // counters at the last abubhost_stats call
    int hint_first = -1, hint_n = 0;            // abubhost_recon_hint: the slots of the next call ...
    std::vector<int> hint_nframes;              // ... and their frame counts
} g;
int g_threads = 0;
}  // namespace

extern "C" {

const char* abubhost_last_error() { return g.err.c_str(); }

/* Configures the device singleton and creates the context; *ctx receives the abub_ctx the C ABI of include/abub_cuda.h
 * takes (training, slot pointers, pinned buffers). */
int abubhost_open(int device, int rows, int cols, int ncams, int max_frames, int max_slots, int early_chunk, abub_ctx** ctx) {
    try {
        AbubDevice& dev = AbubDevice::Get();
        dev.Configure(device, ncams, max_frames, max_slots);
        dev.SetEarlyChunk(early_chunk);
        *ctx = dev.Context(rows, cols);
        return 0;
    } catch (std::exception& e) {
        g.err = e.what();
        return -1;
    }
}

/* Begins a run: header + output file abub3hs_<run>.txt in out_dir (truncated), Trainers that adopt the planes already
 * installed on the device (abub_set_background / abub_set_background_from_training), training_set_sizes[ncams]. */
int abubhost_begin_run(const char* run_id, const char* out_dir, const char* mask_dir, int frame_offset, const int* training_set_sizes) {
    try {
        AbubDevice& dev = AbubDevice::Get();
        g.pipe.stop();
        for (Trainer* t : g.trainers) delete t;
        g.trainers.clear();
        delete g.parser;
        DetectRun& r = g.run;
        const double keepA = r.T_analyze, keepW = r.T_write;
        r = DetectRun();
        r.T_analyze = keepA; r.T_write = keepW;
        r.run_number = run_id;
        r.out_dir = out_dir;
        if (!r.out_dir.empty() && r.out_dir.back() != '/') r.out_dir += "/";
        r.mask_dir = mask_dir ? mask_dir : "";
        r.eventDir = std::string("mem://") + run_id + "/";
        r.imageFolder = "/Images/";
        r.frameOffset = frame_offset;
        r.numCams = dev.ncams();
        r.progress = false;
        g.parser = new MemParser(r.eventDir, r.imageFolder, "cam%d_image%u.png", frame_offset);
        r.FileParser = g.parser;
        for (int c = 0; c < r.numCams; c++) {
            Trainer* t = new Trainer(c, std::vector<std::string>(), r.eventDir, "cam%d_image%u.png", r.imageFolder, g.parser->clone());
            t->AdoptDevicePlanes(training_set_sizes[c]);
            g.trainers.push_back(t);
        }
        r.Trainers = &g.trainers;
        OutputWriter w(r.out_dir, r.run_number, r.frameOffset, r.numCams);
        w.writeHeader();
        g.pipe.stop();
        if (g.out.is_open()) g.out.close();
        g.out.open(w.fileName(), std::fstream::out | std::fstream::app);
        if (!g.out) { g.err = "cannot open " + w.fileName(); return -1; }
        r.threads = g_threads;
        g.pipe.start(&g.run, &g.out);
        return 0;
    } catch (std::exception& e) {
        g.err = e.what();
        return -1;
    }
}

/* One batch: slots [slot_first, slot_first + nslots) hold the events event_numbers[nslots] with nframes[nslots*ncams]
 * frames per camera.  mode 0: the frames are in HBM; mode 1: in the pinned staging buffers, sent chunk_frames at a
 * time only as far as the trigger search and the localizer need them.  Trigger, localizer, per-camera bookkeeping and
 * output rows (appended to the run's file per event).  stats[8] (may be NULL) += {device trigger s, result read-back s,
 * localizer s, host analysis s, file write s, frames sent, 0, 0}. */
int abubhost_recon(int slot_first, int nslots, const int* event_numbers, const int* nframes, int mode, int chunk_frames, double* stats) {
    try {
        AbubDevice& dev = AbubDevice::Get();
        DetectRun& r = g.run;
        const int nc = r.numCams;
        std::vector<int> tss;
        for (int c = 0; c < nc; c++) tss.push_back(g.trainers[c]->TrainingSetSize);
        long long sent = 0;
        if (mode != 0) g.hint_first = -1;          /* an announcement only holds for a resident-form call that follows it directly */
        /* sub-batches: while the per-camera objects of one run on the host cores (DetectPipeline), the device works on the next */
        const int sub = g.sub_batch > 0 ? g.sub_batch : nslots;
        for (int s0 = 0; s0 < nslots; s0 += sub) {
            const int s1 = s0 + sub < nslots ? s0 + sub : nslots;
            std::vector<std::string> chunk, keys;
            std::vector<int> slots;
            /* a slot whose previous event is still with the host stage: wait for it (a caller that alternates between
             * two sets of slots never waits here) */
            for (int s = s0; s < s1; s++)
                if (dev.SlotInUse(slot_first + s)) { g.pipe.drain(); break; }
            for (int s = s0; s < s1; s++) {
                const std::string ev = std::to_string(event_numbers[s]);
                chunk.push_back(ev);
                keys.push_back(r.eventDir + ev + r.imageFolder + "|" + ev);
                g.parser->SetEvent(ev, nframes + (size_t)s * nc, nc);
                dev.BindSlot(keys.back(), slot_first + s);
                dev.MarkResident(slot_first + s, nframes + (size_t)s * nc, mode == 0, false);
                slots.push_back(slot_first + s);
            }
            if (mode == 0) dev.PushFrameCounts(slot_first + s0, s1 - s0, nframes + (size_t)s0 * nc);   /* one launch, not one per slot */
            if (mode == 0) {
                /* what the device will be asked for next -- the following sub-batch of this call, or the batch announced
                 * with abubhost_recon_hint -- has its trigger search queued inside this sub-batch's localizer call */
                if (s1 < nslots) {
                    const int n1 = (s1 + sub < nslots ? s1 + sub : nslots) - s1;
                    dev.AnalyzeBatch(slots, tss.data(), slot_first + s1, n1, nframes + (size_t)s1 * nc);
                } else if (g.hint_first >= 0 && (int)g.hint_nframes.size() == g.hint_n * nc) {
                    const int hf = g.hint_first, hn = g.hint_n;
                    g.hint_first = -1;
                    dev.AnalyzeBatch(slots, tss.data(), hf, hn, g.hint_nframes.data());
                } else {
                    dev.AnalyzeBatch(slots, tss.data());
                }
            } else {
                sent += dev.StreamBatch(slots, tss.data(), chunk_frames);
            }
            g.pipe.submit(std::move(chunk), std::move(keys));
        }
        if (stats) stats[5] += (double)sent;
        (void)stats;
        return 0;
    } catch (std::exception& e) {
        g.err = e.what();
        return -1;
    }
}

/* Optional: the NEXT abubhost_recon call (mode 0) will be for exactly these slots with these frame counts, and their frames
 * are already in HBM.  The current call then queues that batch's trigger search inside its own localizer call
 * (abub_trigger_early_deferred), and the next call finds it running or done.  A hint that does not come true costs one
 * trigger search; a next call that differs from the hint simply starts its own. */
int abubhost_recon_hint(int slot_first, int nslots, const int* nframes) {
    const int nc = g.run.numCams;
    if (slot_first < 0 || nslots < 1 || !nframes || nc < 1) { g.hint_first = -1; return -1; }
    g.hint_first = slot_first;
    g.hint_n = nslots;
    g.hint_nframes.assign(nframes, nframes + (size_t)nslots * nc);
    return 0;
}

/* Waits until every submitted event has been analysed and its rows are in the file. */
int abubhost_drain() {
    try {
        g.pipe.drain();
        return 0;
    } catch (std::exception& e) {
        g.err = e.what();
        return -1;
    }
}

/* Host threads of the event loop (0 = OpenMP's default) and events per sub-batch of abubhost_recon (the unit that is
 * handed from the device stage to the host stage; 0 = the whole call).  Take effect at the next abubhost_begin_run. */
int abubhost_set_threads(int threads, int sub_batch) {
    g_threads = threads < 0 ? 0 : threads;
    g.sub_batch = sub_batch < 0 ? 0 : sub_batch;
    return 0;
}

/* Seconds spent since the previous call of this function (after draining the pipeline): out[5] = {device trigger
 * launches, trigger read-back (waits for K1/K2), localizer, host analysis, file writes}. */
int abubhost_stats(double* out) {
    if (abubhost_drain()) return -1;
    AbubDevice& dev = AbubDevice::Get();
    const double now[5] = {dev.t_hist, dev.t_scan, dev.t_loc, g.run.T_analyze, g.run.T_write};
    for (int i = 0; i < 5; i++) { out[i] = now[i] - g.base[i]; g.base[i] = now[i]; }
    return 0;
}

int abubhost_end_run() {
    int rc = 0;
    try {
        g.pipe.stop();
    } catch (std::exception& e) {
        g.err = e.what();
        rc = -1;
    }
    if (g.out.is_open()) g.out.close();
    g.hint_first = -1;
    return rc;
}

void abubhost_close() {
    abubhost_end_run();
    for (Trainer* t : g.trainers) delete t;
    g.trainers.clear();
    delete g.parser;
    g.parser = nullptr;
    AbubDevice::Get().Shutdown();
}

}  // extern "C"
