#include "Trainer.hpp"

#include <stdio.h>
#include <string.h>

#include <iostream>

#include "AbubDevice.hpp"
#include "ImageCodec.hpp"

Trainer::Trainer(int camera_, std::vector<std::string> EventList_, std::string EventDir_, std::string ImageFormat_,
                 std::string ImageFolder_, Parser* parser, bool debug_) {
    camera = camera_;
    EventList = EventList_;
    EventDir = EventDir_;
    ImageFormat = ImageFormat_;
    ImageFolder = ImageFolder_;
    FileParser = parser;
    debug = debug_;
    StatusCode = 0;
    TrainingSetSize = 0;
    SearchPattern = ImageFormat.substr(0, ImageFormat.find("%u"));     // Trainer.cpp: frame name up to the index
}

// Trainer.cpp:45-73: copies the trained data; the copy does not own a parser
Trainer::Trainer(const Trainer& o) {
    camera = o.camera; EventDir = o.EventDir; ImageFolder = o.ImageFolder; debug = o.debug;
    FileParser = nullptr;
    installed = o.installed;
    ImageFormat = o.ImageFormat; SearchPattern = o.SearchPattern;
    TrainedAvgImage = o.TrainedAvgImage; TrainedSigmaImage = o.TrainedSigmaImage;
    isLBPApplied = o.isLBPApplied; StatusCode = o.StatusCode; TrainingSetSize = o.TrainingSetSize;
    EventList = o.EventList; TrainingSequence = o.TrainingSequence;
}

Trainer::~Trainer(void) {
    if (FileParser) delete FileParser;
}

void Trainer::InstallOnDevice() {
    if (installed) return;
    AbubDevice& dev = AbubDevice::Get();
    if (TrainedAvgImage.empty() || TrainedSigmaImage.empty()) throw AbubError("Trainer has no trained planes to install");
    std::lock_guard<std::recursive_mutex> g(dev.lock());
    abub_ctx* ctx = dev.Context(TrainedAvgImage.rows, TrainedAvgImage.cols);
    dev.Check(abub_set_background(ctx, camera, TrainedAvgImage.ptr(), TrainedSigmaImage.ptr(), TrainingSetSize),
              "abub_set_background");
    installed = true;
}

void Trainer::MakeAvgSigmaImage(bool PerformLBPOnImages) {
    if (PerformLBPOnImages) throw AbubError("LBP training (MakeAvgSigmaImage(true)) is not part of the abub3hs path");
    isLBPApplied = false;
    AbubDevice& dev = AbubDevice::Get();
    printf("Camera %d training ... ", camera);
    const int ev_begin = ShardBegin, ev_end = ShardEnd < 0 ? (int)EventList.size() : ShardEnd;
    const int failure = FileParser->FailureCode();
    bool began = false;
    int kept_frames = 0;
    GateEntropies.clear();

    // events are staged in chunks: the encoded frames 0 and 1 are read in order, decoded on every core, then
    // gated and pushed in event order (push order defines the float32 Welford result, Trainer.cpp:185-187)
    const int CHUNK = 32;
    for (int c0 = ev_begin; c0 < ev_end; c0 += CHUNK) {
        const int c1 = c0 + CHUNK < ev_end ? c0 + CHUNK : ev_end;
        const int nev = c1 - c0;
        std::vector<std::vector<uint8_t>> enc(2 * (size_t)nev);
        std::vector<char> good(nev, 0);
        std::vector<std::string> names(2 * (size_t)nev);
        for (int e = 0; e < nev; e++) {
            CameraFrames.clear();
            FileParser->ParseAndSortFramesInFolder(EventList[c0 + e], camera, CameraFrames);
            if (CameraFrames.size() >= 2) {                       // Trainer.cpp:254: frames {0,1} of the event
                good[e] = 1;
                for (int k = 0; k < 2; k++) names[2 * e + k] = CameraFrames[TrainingSequence[k]];
            } else {
                std::cout << "Event " << EventList[c0 + e] << " is nonexistant on the disk. Skipping training on this event\n";
            }
            CameraFrames.clear();
        }
        // the two files of every event are read on all cores (positional reads in the zip case)
#pragma omp parallel for schedule(dynamic, 1)
        for (int i = 0; i < 2 * nev; i++) {
            const int e = i / 2;
            if (names[i].empty()) continue;
            int err = failure;
            try { err = FileParser->GetImageBytes(EventList[c0 + e], names[i], enc[i]); } catch (...) { err = failure; enc[i].clear(); }
            if (err == failure || enc[i].empty()) {
#pragma omp critical
                { std::cout << "Skipping corrupted image for training.\n"; good[e] = 0; }
            }
        }
        // size the context from the first readable frame
        if (!dev.ctx()) {
            for (int e = 0; e < nev && !dev.ctx(); e++) {
                int r, c;
                if (good[e] && abubcodec::peek_size(enc[2 * e].data(), enc[2 * e].size(), &r, &c)) dev.Context(r, c);
            }
            if (!dev.ctx()) continue;
        }
        const size_t npix = (size_t)dev.rows() * dev.cols();
        std::vector<uint8_t> px(2 * (size_t)nev * npix);
#pragma omp parallel for schedule(dynamic, 1)
        for (int i = 0; i < 2 * nev; i++) {
            const int e = i / 2;
            if (!good[e]) continue;
            bool dec = false;
            try { dec = abubcodec::decode_into(enc[i].data(), enc[i].size(), px.data() + (size_t)i * npix, dev.rows(), dev.cols()); } catch (...) {}
            if (!dec) {
#pragma omp critical
                { std::cout << "Skipping corrupted image for training.\n"; good[e] = 0; }
            }
        }
        std::lock_guard<std::recursive_mutex> g(dev.lock());
        if (!began) { dev.Check(abub_train_begin(dev.ctx(), camera), "abub_train_begin"); began = true; }
        // runs of readable events are contiguous in px: one gate launch and one read-back per run
        std::vector<float> entropy(nev);
        std::vector<int> kept(nev);
        for (int e = 0; e < nev;) {
            if (!good[e]) { e++; continue; }
            int e1 = e;
            while (e1 < nev && good[e1]) e1++;
            dev.Check(abub_train_add_events(dev.ctx(), camera, px.data() + 2 * (size_t)e * npix, e1 - e, entropy.data() + e,
                                            kept.data() + e), "abub_train_add_events");
            for (int i = e; i < e1; i++) {
                GateEntropies.push_back(entropy[i]);
                if (kept[i]) kept_frames += 2;
            }
            e = e1;
        }
    }

    if (!NcclComm && kept_frames == 0) {                          // Trainer.cpp:308-313
        std::cout << "Training image set for camera " << camera
                  << " has 0 frames. This means that the event is malformed." << std::endl;
        StatusCode = -7;
        return;
    }
    std::lock_guard<std::recursive_mutex> g(dev.lock());
    if (!dev.ctx()) { StatusCode = -7; return; }
    if (!began) dev.Check(abub_train_begin(dev.ctx(), camera), "abub_train_begin");
    const size_t npix = (size_t)dev.rows() * dev.cols();
    TrainedAvgImage.create(dev.rows(), dev.cols());
    TrainedSigmaImage.create(dev.rows(), dev.cols());
    int status = 0;
    if (NcclComm) {
        // exact integer partials, one NCCL all-reduce, fp64 finish + boundary fix-up
        uint32_t* d_sum; uint64_t* d_sumsq; int* d_count;
        dev.Check(abub_train_partials(dev.ctx(), camera, &d_sum, &d_sumsq, &d_count), "abub_train_partials");
        dev.Check(abub_train_allreduce(dev.ctx(), camera, NcclComm), "abub_train_allreduce");
        dev.Check(abub_train_finish_allreduced(dev.ctx(), camera, NcclComm, TrainedAvgImage.ptr(), TrainedSigmaImage.ptr(),
                                               &TrainingSetSize, &status, &FlaggedPixels), "abub_train_finish_allreduced");
    } else {
        dev.Check(abub_train_finish(dev.ctx(), camera, TrainedAvgImage.ptr(), TrainedSigmaImage.ptr(), &TrainingSetSize,
                                    &status), "abub_train_finish");
    }
    (void)npix;
    if (status != 0) {
        std::cout << "Training image set for camera " << camera
                  << " has 0 frames. This means that the event is malformed." << std::endl;
        StatusCode = status;
        TrainedAvgImage.release(); TrainedSigmaImage.release();
        return;
    }
    dev.Check(abub_set_background_from_training(dev.ctx(), camera), "abub_set_background_from_training");
    installed = true;
    StatusCode = 0;
    printf("complete.\n");
}
