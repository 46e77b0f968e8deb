// abub3hs -- drop-in driver (reference: AutoBubStart3.cpp:67-125 AnyCamAnalysis, :148-405 main).
// Same flags, data-series table, error codes (-5 unreadable run, -7 training failed, -6 per-camera
// exception, -8/-9/-3 from the analyzers) and output file.  What differs is the schedule: instead of
// one OpenMP thread per event calling OpenCV, events are staged in batches on the GPU (decode on all
// host cores -> pinned slot -> cudaMemcpyAsync -> one K1 launch per batch) and the per-camera objects
// then run against resident data.  Extra flags (not in the reference): --device, --batch, --max_frames,
// and --rank/--world/--nccl_id for one process per GPU on a multi-GPU box.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <exception>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "AbubDevice.hpp"
#include "AnalyzerUnit.hpp"
#include "L3Localizer.hpp"
#include "PICOFormatWriterV4.hpp"
#include "RawParser.hpp"
#include "RankSync.hpp"
#include "ReconDriver.hpp"
#include <chrono>
#include <time.h>

#include "Trainer.hpp"
#include "ZipParser.hpp"

static bool eventNameOrderSort(std::string a, std::string b) { return std::stoi(a) < std::stoi(b); }

static std::string usage() {
    return "Usage: abub3hs [-hzme] [-D data_series] [-c cam_mask_dir] [--debug code] -d data_dir -r run_ID -o out_dir\n"
           "Run the AutoBub3hs bubble finding algorithm on a PICO run (B200 build)\n\n"
           "Required arguments:\n"
           "  -d, --data_dir = Dir\t\tpath to the directory in which the run folder/file is stored\n"
           "  -r, --run_id = Str\t\trun ID, formatted as YYYYMMDD_*\n"
           "  -o, --out_dir = Dir\t\tdirectory to write the output file to\n\n"
           "Optional arguments:\n"
           "  -h, --help\t\t\tgive this help message\n"
           "  -z, --zip\t\t\tindicate the run is stored as a zip file; otherwise assumed to be in a directory\n"
           "  -c, --cam_mask_dir = Dir\tdirectory containing the camera mask images\n"
           "  -m, --mask_check\t\tuse camera masks in default directory. Not needed if directory specified\n"
           "  -D, --data_series = Str\tname of the data series, e.g. 40l-19, 30l-16, etc.\n"
           "  -e, --event = Int\t\tspecify a single event to process. Mostly just useful for debugging and testing\n"
           "  --debug = Int\t\t\t3 digit int; eg: 101: first digit = localizer debug; second digit = multithread off; third digit = analyzer debug\n"
           "GPU arguments (not in the reference):\n"
           "  --device = Int\t\tCUDA device ordinal (default 0, or the rank with --world)\n"
           "  --batch = Int\t\t\tevents staged on the GPU at once (default 8; 282 MB of HBM each at 4x40 frames of 1680x1050)\n"
           "  --early_chunk = Int\t\tframe pairs per round of the trigger search's early exit (default 5; 0 = all pairs at once)\n"
           "  --max_frames = Int\t\tframes per camera a slot can hold (default: what the first events hold, at least 40)\n"
           "  --trigger_metric = Str	significance (default) | entropy | entropy_significance: which of the three lines of\n"
           "\t\t\t\tAnalyzerUnit.cpp:226-228 is the live one; --trigger_threshold = Float overrides its entropyThreshold\n"
           "  --rank/--world = Int\t\tone process per GPU: this process handles a contiguous range of the events\n"
           "  --nccl_id = File\t\tfile through which rank 0 publishes the NCCL id (default <out_dir>/.abub3hs_<run>.ncclid)\n";
}

struct Options {
    std::string dataLoc, run_number, out_dir, mask_dir, data_series, nccl_id, trigger_metric = "significance", trigger_threshold = "0";
    int event_user = -1, debug_mode = 0, device = -1, batch = 8, early_chunk = 5, max_frames = 0, rank = 0, world = 1;
    bool zipped = false, mask_check = false, help = false;
};

/* boost::program_options replacement: "-x value", "--long value" and "--long=value" */
static bool parse_args(int argc, char** argv, Options& o) {
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i], val;
        bool has_val = false;
        if (a.rfind("--", 0) == 0) {
            const size_t eq = a.find('=');
            if (eq != std::string::npos) { val = a.substr(eq + 1); a = a.substr(0, eq); has_val = true; }
        }
        auto need = [&](std::string& dst) {
            if (!has_val) { if (i + 1 >= argc) return false; val = argv[++i]; }
            dst = val;
            return true;
        };
        auto need_int = [&](int& dst) { std::string s; if (!need(s)) return false; dst = atoi(s.c_str()); return true; };
        bool ok = true;
        if (a == "-h" || a == "--help") o.help = true;
        else if (a == "-z" || a == "--zip") o.zipped = true;
        else if (a == "-m" || a == "--mask_check") o.mask_check = true;
        else if (a == "-D" || a == "--data_series") ok = need(o.data_series);
        else if (a == "-d" || a == "--data_dir") ok = need(o.dataLoc);
        else if (a == "-r" || a == "--run_num" || a == "--run_id") ok = need(o.run_number);
        else if (a == "-o" || a == "--out_dir") ok = need(o.out_dir);
        else if (a == "-c" || a == "--cam_mask_dir") ok = need(o.mask_dir);
        else if (a == "-e" || a == "--event") ok = need_int(o.event_user);
        else if (a == "--debug") ok = need_int(o.debug_mode);
        else if (a == "--device") ok = need_int(o.device);
        else if (a == "--batch") ok = need_int(o.batch);
        else if (a == "--early_chunk") ok = need_int(o.early_chunk);
        else if (a == "--max_frames") ok = need_int(o.max_frames);
        else if (a == "--trigger_metric") ok = need(o.trigger_metric);
        else if (a == "--trigger_threshold") ok = need(o.trigger_threshold);
        else if (a == "--rank") ok = need_int(o.rank);
        else if (a == "--world") ok = need_int(o.world);
        else if (a == "--nccl_id") ok = need(o.nccl_id);
        else { std::cerr << "unrecognised option '" << a << "'" << std::endl; return false; }
        if (!ok) { std::cerr << "the required argument for option '" << a << "' is missing" << std::endl; return false; }
    }
    return true;
}

int main(int argc, char** argv) {
    Options opt;
    if (!parse_args(argc, argv, opt)) return -1;
    if (opt.help || argc == 1) {
        std::cout << usage() << std::endl;
        return 1;
    }
    if (opt.dataLoc == "" || opt.run_number == "" || opt.out_dir == "") {
        std::cerr << "Insufficient required arguments; use \"autobub3hs -h\" to view required arguments" << std::endl;
        return -1;
    }
    std::string dataLoc = opt.dataLoc, run_number = opt.run_number, out_dir = opt.out_dir, mask_dir = opt.mask_dir,
                data_series = opt.data_series;
    const int debug_mode = opt.debug_mode, event_user = opt.event_user;
    const bool lead = opt.rank == 0;

    printf("This is AutoBub v3, the automatic unified bubble finder code for all chambers\n");

    std::string this_path = argv[0];
    std::string abub_dir = this_path.substr(0, this_path.find_last_of("/") + 1);
    if (opt.mask_check && mask_dir == "") mask_dir = abub_dir + "cam_masks/" + data_series;
    else if (!opt.mask_check && mask_dir == "") std::cout << "Not performing mask check on this run." << std::endl;

    std::string eventDir = dataLoc + "/" + run_number + "/";
    if (out_dir[out_dir.length() - 1] != '/') out_dir += "/";

    /* how the different experiments stored their images (AutoBubStart3.cpp:216-245) */
    std::string imageFormat, imageFolder;
    int frameOffset, numCams;
    if (data_series == "01l-21" || data_series == "2l-16") {
        imageFormat = "cam%dimage %u.bmp"; imageFolder = "/"; frameOffset = 0; numCams = 2;
    } else if (data_series == "40l-19") {
        imageFormat = "cam%d_image%u.png"; imageFolder = "/Images/"; frameOffset = 30;
        numCams = (run_number >= "20200713_7") ? 4 : 2;
        if (run_number < "20200501_1") frameOffset = 20;
    } else {                                       /* 40l-22 and the default */
        imageFormat = "cam%d_image%u.png"; imageFolder = "/Images/"; frameOffset = 30; numCams = 4;
    }

    OutputWriter* PICO60Output = new OutputWriter(out_dir, run_number, frameOffset, numCams);
    if (lead) PICO60Output->writeHeader();

    std::vector<std::string> EventList;
    Parser* FileParser = nullptr;
    try {
        if (opt.zipped) FileParser = new ZipParser(eventDir, imageFolder, imageFormat);
        else FileParser = new RawParser(eventDir, imageFolder, imageFormat);
        FileParser->GetEventDirLists(EventList);
        std::sort(EventList.begin(), EventList.end(), eventNameOrderSort);
    } catch (...) {
        /* the run cannot be read: -5 (AutoBubStart3.cpp:273-281) */
        std::cout << "Failed to read the images from run " << run_number << ". Autobub cannot continue.\n";
        if (lead) {
            for (int icam = 0; icam < numCams; icam++) PICO60Output->stageCameraOutputError(icam, -5, -1);
            PICO60Output->writeCameraOutput();
        }
        return -5;
    }
    const int nEvents = (int)EventList.size();

    /* device capacities */
    int max_frames = opt.max_frames;
    if (max_frames <= 0) {
        max_frames = 40;
        for (int evi = 0; evi < nEvents && evi < 4; evi++)
            for (int icam = 0; icam < numCams; icam++) {
                std::vector<std::string> fr;
                FileParser->ParseAndSortFramesInFolder(EventList[evi], icam, fr);
                max_frames = std::max(max_frames, (int)fr.size());
            }
    }
    const int batch = std::max(1, opt.batch);
    void* nccl = nullptr;
    try {
        /* two batches of slots: one being analysed on the host cores while the next is staged and run on the device */
        AbubDevice::Get().Configure(opt.device >= 0 ? opt.device : opt.rank, numCams, max_frames, 2 * batch);
        const int metric = opt.trigger_metric == "significance" ? ABUB_METRIC_SIGNIFICANCE
                         : opt.trigger_metric == "entropy" ? ABUB_METRIC_ENTROPY
                         : opt.trigger_metric == "entropy_significance" ? ABUB_METRIC_ENTROPY_SIGNIFICANCE : -1;
        if (metric < 0) throw std::runtime_error("--trigger_metric: significance | entropy | entropy_significance");
        AbubDevice::Get().SetEarlyChunk(opt.early_chunk);
        AbubDevice::Get().SetTriggerMetric(metric, (float)atof(opt.trigger_threshold.c_str()));
    } catch (std::exception& e) {
        std::cerr << e.what() << std::endl;
        return -5;
    }

    /* event range of this process (contiguous ranges, rank order = file order) */
    const int ev_lo = (int)((long long)nEvents * opt.rank / opt.world), ev_hi = (int)((long long)nEvents * (opt.rank + 1) / opt.world);

    printf("**Starting training. AutoBub is in learn mode**\n");
    auto wall = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double T_train0 = wall();
    double T_stage = 0, T_analyze = 0, T_write = 0;
    std::vector<Trainer*> Trainers;
    for (int icam = 0; icam < numCams; icam++)
        Trainers.push_back(new Trainer(icam, EventList, eventDir, imageFormat, imageFolder, FileParser->clone(), debug_mode / 100));
    bool succeeded = true;
    RankSync sync;
    sync.base = out_dir + ".abub3hs_" + run_number + ".status";
    sync.rank = opt.rank; sync.world = opt.world; sync.t_start = (long long)time(nullptr);
    const std::string part_path = out_dir + "abub3hs_" + run_number + ".txt" + (opt.world > 1 ? ".rank" + std::to_string(opt.rank) : "");
    try {
        if (opt.world > 1) {
            /* ranks train on their own events and all-reduce exact integer partials; needs a context first */
            const std::string idfile = opt.nccl_id != "" ? opt.nccl_id : out_dir + ".abub3hs_" + run_number + ".ncclid";
            for (int icam = 0; icam < numCams; icam++) { Trainers[icam]->ShardBegin = ev_lo; Trainers[icam]->ShardEnd = ev_hi; }
            /* size the context from the first frame this rank can read */
            for (int evi = ev_lo; evi < ev_hi && !AbubDevice::Get().ctx(); evi++) {
                std::vector<std::string> fr;
                FileParser->ParseAndSortFramesInFolder(EventList[evi], 0, fr);
                Image8 im;
                if (!fr.empty() && FileParser->GetImage(EventList[evi], fr[0], im) != FileParser->FailureCode() && !im.empty())
                    AbubDevice::Get().Context(im.rows, im.cols);
            }
            /* every rank must have a context before any of them enters ncclCommInitRank */
            const bool have_ctx = AbubDevice::Get().ctx() != nullptr;
            if (!have_ctx) std::cout << "rank " << opt.rank << ": no readable frame in this rank's event range\n";
            if (sync.agree("context", have_ctx ? 0 : -1) < 0) throw AbubError("a rank cannot take part in the training: no communicator is created");
            AbubDevice::Get().Check(abub_nccl_comm_create(AbubDevice::Get().ctx(), idfile.c_str(), opt.rank, opt.world, &nccl),
                                    "abub_nccl_comm_create");
            for (int icam = 0; icam < numCams; icam++) Trainers[icam]->NcclComm = nccl;
        }
        for (int icam = 0; icam < numCams; icam++) Trainers[icam]->MakeAvgSigmaImage(false);
    } catch (std::exception& e) {
        std::cout << e.what() << std::endl;
        succeeded = false;
    }
    for (int icam = 0; icam < numCams; icam++)
        if (Trainers[icam]->StatusCode) succeeded = false;
    /* the ranks take the -7 path together or not at all */
    if (opt.world > 1 && sync.agree("trained", succeeded ? 0 : -1) < 0) succeeded = false;

    if (!succeeded) {
        /* -7 for every camera of every event (AutoBubStart3.cpp:315-325) */
        std::cout << "Failed to train on images from run " << run_number << ". Autobub cannot continue.\n";
        if (lead)
            for (int evi = 0; evi < nEvents; evi++) {
                int actualEventNumber = atoi(EventList[evi].c_str());
                for (int icam = 0; icam < numCams; icam++) PICO60Output->stageCameraOutputError(icam, -7, actualEventNumber);
                PICO60Output->writeCameraOutput();
            }
        if (opt.world > 1) {
            /* nothing of a failed multi-rank run stays behind: part file, status files, the NCCL id file */
            unlink(part_path.c_str());
            usleep(300000);                       /* the peers have read this rank's status by now */
            sync.forget(0);
            if (lead) unlink((opt.nccl_id != "" ? opt.nccl_id : out_dir + ".abub3hs_" + run_number + ".ncclid").c_str());
            if (nccl) abub_nccl_comm_destroy(nccl);
        }
        return -7;
    }
    printf("***Training complete. AutoBub is now in detect mode***\n");
    delete PICO60Output;
    const double T_train1 = wall();

    /* Detect mode: batches of events staged on the device, rows appended in event order */
    const std::string part_name = out_dir + "abub3hs_" + run_number + ".txt" + (opt.world > 1 ? ".rank" + std::to_string(opt.rank) : "");
    if (opt.world > 1) { std::ofstream trunc(part_name); }
    /* rows are committed event by event (AutoBubStart3.cpp:379-383); the file stays open between events */
    std::ofstream part_out(part_name, std::fstream::out | std::fstream::app);
    DetectRun drun;
    drun.run_number = run_number; drun.out_dir = out_dir; drun.mask_dir = mask_dir; drun.eventDir = eventDir;
    drun.imageFolder = "/Images/"; drun.frameOffset = frameOffset; drun.numCams = numCams; drun.debug_mode = debug_mode;
    drun.nEvents = nEvents; drun.Trainers = &Trainers; drun.FileParser = FileParser;
    DetectPipeline pipe;
    pipe.start(&drun, &part_out);
    for (int b0 = ev_lo; b0 < ev_hi; b0 += batch) {
        const int b1 = std::min(b0 + batch, ev_hi);
        std::vector<std::string> chunk;
        for (int evi = b0; evi < b1; evi++)
            if (!(event_user >= 0 && evi != event_user)) chunk.push_back(EventList[evi]);
        if (chunk.empty()) continue;
        /* stage the batch under the keys the analyzers will ask for (AnalyzerUnit: ImageDir|EventID) */
        std::vector<std::string> keys;
        for (const std::string& ev : chunk) keys.push_back(eventDir + ev + "/Images/" + "|" + ev);
        const double ts0 = wall();
        try {
            Parser* stage_parser = FileParser->clone();
            std::vector<int> tss;
            for (int icam = 0; icam < numCams; icam++) tss.push_back(Trainers[icam]->TrainingSetSize);
            AbubDevice::Get().PrefetchEvents(stage_parser, keys, chunk, tss.data());
            delete stage_parser;
        } catch (std::exception& e) {
            std::cout << e.what() << std::endl;       /* the analyzers will meet the same problem and report -6 */
        }
        T_stage += wall() - ts0;
        /* the per-camera objects of this batch run on a worker thread while the next batch is staged */
        pipe.submit(chunk, keys);
    }
    pipe.stop();
    T_analyze = drun.T_analyze; T_write = drun.T_write;
    part_out.close();
    printf("run complete.\n");
    if (getenv("ABUB_TIMING")) {
        AbubDevice& d = AbubDevice::Get();
        fprintf(stderr, "ABUB_TIMING train %.3f s | detect: stage %.3f s (setup %.3f read+decode %.3f upload %.3f K1 %.3f K2 %.3f localizer %.3f) analyze %.3f s write %.3f s\n",
                T_train1 - T_train0, T_stage, d.t_read, d.t_decode, d.t_upload, d.t_hist, d.t_scan, d.t_loc, T_analyze, T_write);
        fprintf(stderr, "ABUB_UPLOAD upload = submit calls %.3f + flip calls %.3f + end of staging %.3f s\n", d.t_submit, d.t_flip, d.t_tail);
        fprintf(stderr, "ABUB_DECODE frames decoded on the device %ld, on the host %ld\n", d.n_gpu_decoded, d.n_cpu_decoded);
    }

    if (opt.world > 1) {
        /* every rank has flushed its part; rank 0 concatenates in rank order = event order (omp ordered, :380).  The
         * ranks meet through the status files first: a rank that died in detect mode must not hang the barrier. */
        const int all_done = sync.agree("detected", 0);
        if (all_done < 0) {
            std::cout << "A rank did not finish its events: the part files are left in place, no merged output.\n";
            sync.forget(0);
            return -6;
        }
        AbubDevice::Get().Check(abub_nccl_barrier(AbubDevice::Get().ctx(), nccl), "abub_nccl_barrier");
        sync.forget(0);
        if (lead) {
            std::ofstream out(out_dir + "abub3hs_" + run_number + ".txt", std::fstream::out | std::fstream::app);
            for (int r = 0; r < opt.world; r++) {
                const std::string pn = out_dir + "abub3hs_" + run_number + ".txt.rank" + std::to_string(r);
                std::ifstream in(pn);
                out << in.rdbuf();
                in.close();
                unlink(pn.c_str());
            }
        }
        abub_nccl_comm_destroy(nccl);
    }

    for (int icam = 0; icam < numCams; icam++) delete Trainers[icam];
    delete FileParser;
    AbubDevice::Get().Shutdown();
    printf("AutoBub done analyzing this run. Thank you.\n");
    return 0;
}
