// Trainer -- per-camera background statistics (reference: AlgorithmTraining/Trainer.hpp:18-75,
// Trainer.cpp:221-331).  Same constructor, MakeAvgSigmaImage() and public fields; the gate histogram
// (K3) and the float32 Welford pass (K4) run on the device through abub_train_* and the trained
// planes stay resident there for the analyzers (the reference's per-analyzer deep copy of the
// Trainer, AnalyzerUnit.cpp:27, is not reproduced: analyzers keep a pointer).
#ifndef ABUB_TRAINER_HPP
#define ABUB_TRAINER_HPP

#include <string>
#include <vector>

#include "Image8.hpp"
#include "Parser.hpp"

class Trainer {
private:
    std::vector<std::string> CameraFrames;
    int camera;
    std::string EventDir;
    std::string ImageFolder;
    bool debug;
    Parser* FileParser;
    bool installed = false;          // planes of this object are the ones on the device

public:
    std::string ImageFormat;
    std::string SearchPattern;

    /*direct access to the trained data*/
    Image8 TrainedAvgImage;
    Image8 TrainedSigmaImage;
    bool isLBPApplied = false;       // the LBP branch (Trainer.cpp:295-298) is never taken by abub3hs: unsupported here
    int StatusCode;
    int TrainingSetSize;
    std::vector<std::string> EventList;
    std::vector<int> TrainingSequence{0, 1};

    /* multi-GPU: this rank trains on EventList[ShardBegin, ShardEnd) and the partial sums are all-reduced
     * over NcclComm (ncclComm_t) before the planes are derived; defaults = the whole list, no collective */
    int ShardBegin = 0, ShardEnd = -1;
    void* NcclComm = nullptr;
    int FlaggedPixels = 0;           // pixels re-evaluated by the boundary fix-up
    std::vector<float> GateEntropies;   // entropy of every gated event (diagnostics / tests)

    Trainer(int camera, std::vector<std::string> EventList, std::string EventDir, std::string ImageFormat,
            std::string ImageFolder, Parser* parser, bool debug = false);
    Trainer(const Trainer& other_trainer);
    ~Trainer(void);

    void MakeAvgSigmaImage(bool PerformLBPOnImages);
    /* installs TrainedAvgImage/TrainedSigmaImage/TrainingSetSize of this object on the device (no-op when they
     * already are): lets a caller that filled the public fields by hand use the analyzers */
    void InstallOnDevice();
    /* the planes of this camera are on the device already (abub_set_background / abub_set_background_from_training by
     * the caller): only the bookkeeping the analyzers read is set */
    void AdoptDevicePlanes(int training_set_size) { TrainingSetSize = training_set_size; StatusCode = 0; installed = true; }
    int Camera() const { return camera; }
};

#endif
