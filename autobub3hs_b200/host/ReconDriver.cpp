#include "ReconDriver.hpp"

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
#pragma omp parallel for ordered schedule(static, 1)
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

void MemParser::ParseAndSortFramesInFolder(std::string EventID, int Camera, std::vector<std::string>& Contents) {
    Contents.clear();
    for (size_t e = 0; e < events_.size(); e++) {
        if (events_[e] != EventID) continue;
        if (Camera < 0 || Camera >= (int)nframes_[e].size()) return;
        char name[64];
        for (int f = 0; f < nframes_[e][Camera]; f++) {
            snprintf(name, sizeof(name), "cam%d_image%d.png", Camera, frame_offset_ + f);
            Contents.push_back(name);
        }
        return;
    }
}

void MemParser::SetEvent(const std::string& event, const int* nframes, int ncams) {
    for (size_t e = 0; e < events_.size(); e++)
        if (events_[e] == event) { nframes_[e].assign(nframes, nframes + ncams); return; }
    events_.push_back(event);
    nframes_.emplace_back(nframes, nframes + ncams);
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
} g;
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
        for (Trainer* t : g.trainers) delete t;
        g.trainers.clear();
        delete g.parser;
        DetectRun& r = g.run;
        r = DetectRun();
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
        if (g.out.is_open()) g.out.close();
        g.out.open(w.fileName(), std::fstream::out | std::fstream::app);
        if (!g.out) { g.err = "cannot open " + w.fileName(); return -1; }
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
        std::vector<std::string> chunk, keys;
        std::vector<int> slots, tss;
        g.parser->Clear();
        for (int s = 0; s < nslots; s++) {
            const std::string ev = std::to_string(event_numbers[s]);
            chunk.push_back(ev);
            keys.push_back(r.eventDir + ev + r.imageFolder + "|" + ev);
            g.parser->SetEvent(ev, nframes + (size_t)s * nc, nc);
            dev.BindSlot(keys.back(), slot_first + s);
            dev.MarkResident(slot_first + s, nframes + (size_t)s * nc, mode == 0);
            slots.push_back(slot_first + s);
        }
        for (int c = 0; c < nc; c++) tss.push_back(g.trainers[c]->TrainingSetSize);
        const double a0 = dev.t_hist, a1 = dev.t_scan, a2 = dev.t_loc, a3 = r.T_analyze, a4 = r.T_write;
        long long sent = 0;
        if (mode == 0) dev.AnalyzeBatch(slots, tss.data());
        else sent = dev.StreamBatch(slots, tss.data(), chunk_frames);
        DetectStagedBatch(r, chunk, keys, g.out);
        if (stats) {
            stats[0] += dev.t_hist - a0; stats[1] += dev.t_scan - a1; stats[2] += dev.t_loc - a2;
            stats[3] += r.T_analyze - a3; stats[4] += r.T_write - a4; stats[5] += (double)sent;
        }
        return 0;
    } catch (std::exception& e) {
        g.err = e.what();
        return -1;
    }
}

int abubhost_end_run() {
    if (g.out.is_open()) g.out.close();
    return 0;
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
