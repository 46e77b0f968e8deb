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

#include <condition_variable>
#include <deque>
#include <exception>
#include <map>
#include <memory>
#include <fstream>
#include <mutex>
#include <string>
#include <thread>
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
    int threads = 0;                            // host threads of the event loop (0 = OpenMP's default, i.e. OMP_NUM_THREADS / all cores)
    std::vector<Trainer*>* Trainers = nullptr;
    Parser* FileParser = nullptr;               // cloned per analyzer, as the reference does (:362)
    double T_analyze = 0, T_write = 0;          // wall-clock seconds, accumulated
};

// Analyzes the events `chunk` (ids) staged under `keys` (one slot reference each, released here) and appends their rows to
// `out` in chunk order, flushing after every event.
void DetectStagedBatch(DetectRun& run, const std::vector<std::string>& chunk, const std::vector<std::string>& keys, std::ofstream& out);

// One batch behind the device: DetectStagedBatch of batch k runs on a worker thread (with its OpenMP team) while the caller
// stages and launches batch k+1.  Batches are analysed in submission order, so the rows reach the file in event order.
class DetectPipeline {
public:
    DetectPipeline() {}
    ~DetectPipeline() { stop(); }
    void start(DetectRun* run, std::ofstream* out);
    // hands a staged batch to the worker; waits first until the batch before it has been analysed (so at most one batch
    // is being analysed while one is being staged) and rethrows what the worker threw
    void submit(std::vector<std::string> chunk, std::vector<std::string> keys);
    void drain();                                // returns when nothing is queued or running
    void stop();

private:
    struct Job { std::vector<std::string> chunk, keys; };
    void loop();
    std::thread th_;
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<Job> q_;
    int inflight_ = 0;
    bool stop_ = false, started_ = false;
    std::exception_ptr err_;
    DetectRun* run_ = nullptr;
    std::ofstream* out_ = nullptr;
};

// Frames that are already in device slots or pinned staging buffers (a device-side producer, a DAQ ring, the benchmark):
// lists the 40l-19 frame names of an event without touching a file system.  GetImage has nothing to decode and fails.
class MemParser : public Parser {
public:
    MemParser(std::string run_folder, std::string image_folder, std::string image_format, int frame_offset)
        : Parser(run_folder, image_folder, image_format), frame_offset_(frame_offset), tab_(std::make_shared<Table>()) {}
    // clones share the event table (the analyzers' parsers are clones made on other threads while the driver adds events)
    Parser* clone() override { MemParser* p = new MemParser(RunFolder, ImageFolder, ImageFormat, frame_offset_); p->tab_ = tab_; return p; }
    int GetImage(std::string, std::string, Image8&) override { return 0; }
    int GetImageBytes(std::string, std::string, std::vector<uint8_t>&) override { return 0; }
    int FailureCode() const override { return 0; }
    void GetEventDirLists(std::vector<std::string>& EventList) override;
    void GetFileLists(const char*, std::vector<std::string>&, const char*) override {}
    void ParseAndSortFramesInFolder(std::string EventID, int Camera, std::vector<std::string>& Contents) override;
    // declares an event; the table keeps the last `keep` events (older ones have been analysed long ago)
    void SetEvent(const std::string& event, const int* nframes, int ncams, size_t keep = 4096);

private:
    struct Table {
        std::mutex mu;
        std::map<std::string, std::vector<int>> nframes;
        std::deque<std::string> order;
    };
    int frame_offset_;
    std::shared_ptr<Table> tab_;
};

#endif
