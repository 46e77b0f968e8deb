// ReconDriver -- the detect-mode loop of main (reference: AutoBubStart3.cpp:342-388) over a batch of events that are
// staged on the device, shared by the abub3hs binary and by the harness entry points below.
//
// The reference runs `#pragma omp parallel for ordered schedule(static,1)` over events, the cameras of an event in series
// (AnyCamAnalysis, :67-125), and appends the rows of an event to the output file inside `#pragma omp ordered`
// (:379-383: "so in the event of a crash, its not lost").  DetectStagedBatch keeps exactly that shape: the per-camera
// objects of the events of a batch are built and run on the host cores in parallel (they find their trigger and
// localizer results in the slot caches filled by AbubDevice::AnalyzeBatch / StreamBatch), and every event's rows are
// committed to the file, in event order, as soon as the event is done.
#ifndef ABUB_RECONDRIVER_HPP
#define ABUB_RECONDRIVER_HPP

#include <fstream>
#include <string>
#include <vector>

#include "AnalyzerUnit.hpp"
#include "PICOFormatWriterV4.hpp"
#include "Parser.hpp"
#include "Trainer.hpp"

/* the same routine for every camera (AutoBubStart3.cpp:67-125) */
int AnyCamAnalysis(std::string EventID, std::string ImgDir, int camera, bool nonStopPref, Trainer** TrainingData,
                   OutputWriter** P60Output, std::string out_dir, int actualEventNumber, AnalyzerUnit** AGeneric);

struct DetectRun {
    std::string run_number, out_dir, mask_dir, eventDir, imageFolder;
    int frameOffset = 30, numCams = 4, debug_mode = 0, nEvents = 0;
    bool progress = true;                       // the "\rProcessing event" line and spinner of the reference
    std::vector<Trainer*>* Trainers = nullptr;
    Parser* FileParser = nullptr;               // cloned per analyzer, as the reference does (:362)
    double T_analyze = 0, T_write = 0;          // wall-clock seconds, accumulated
};

// Analyzes the events `chunk` (ids) staged under `keys` (one slot reference each, released here) and appends their rows to
// `out` in chunk order, flushing after every event.
void DetectStagedBatch(DetectRun& run, const std::vector<std::string>& chunk, const std::vector<std::string>& keys, std::ofstream& out);

// Frames that are already in device slots or pinned staging buffers (a device-side producer, a DAQ ring, the benchmark):
// lists the 40l-19 frame names of an event without touching a file system.  GetImage has nothing to decode and fails.
class MemParser : public Parser {
public:
    MemParser(std::string run_folder, std::string image_folder, std::string image_format, int frame_offset)
        : Parser(run_folder, image_folder, image_format), frame_offset_(frame_offset) {}
    Parser* clone() override { MemParser* p = new MemParser(RunFolder, ImageFolder, ImageFormat, frame_offset_); p->nframes_ = nframes_; p->events_ = events_; return p; }
    int GetImage(std::string, std::string, Image8&) override { return 0; }
    int GetImageBytes(std::string, std::string, std::vector<uint8_t>&) override { return 0; }
    int FailureCode() const override { return 0; }
    void GetEventDirLists(std::vector<std::string>& EventList) override { EventList = events_; }
    void GetFileLists(const char*, std::vector<std::string>&, const char*) override {}
    void ParseAndSortFramesInFolder(std::string EventID, int Camera, std::vector<std::string>& Contents) override;
    void SetEvent(const std::string& event, const int* nframes, int ncams);
    void Clear() { events_.clear(); nframes_.clear(); }

private:
    int frame_offset_;
    std::vector<std::string> events_;
    std::vector<std::vector<int>> nframes_;
};

#endif
