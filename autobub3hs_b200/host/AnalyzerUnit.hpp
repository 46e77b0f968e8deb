// AnalyzerUnit -- per-(event, camera) base class: frame list and trigger search (reference:
// AnalyzerUnit.hpp:20-100, AnalyzerUnit.cpp:17-324).  Same constructor, FindTriggerFrame() and public
// state; ProcessFrame + calcHist (K1) and the significance state machine (K2) run on the device.
// A retrigger (startframe > 1) resumes with the history of frames 1..startframe-1, exactly like the
// reference keeps pix_counts (AnalyzerUnit.cpp:165): the history is the device-resident histogram table.
#ifndef ABUB_ANALYZERUNIT_HPP
#define ABUB_ANALYZERUNIT_HPP

#include <string>
#include <vector>

#include "Image8.hpp"
#include "Parser.hpp"
#include "Trainer.hpp"
#include "bubble.hpp"

class AnalyzerUnit {
private:
    int StatusCode = 0;
    int minEvalFrameNumber = 2;
    int loc_thres_max = 3;

protected:
    std::string ImageDir;
    std::string EventID;
    std::string MaskDir;
    std::vector<std::string> CameraFrames;
    int CameraNumber;
    Parser* FileParser;

    /* device residency of this analyzer's frames */
    std::string SlotKey;
    int Slot = -1;
    void EnsureFramesOnDevice();
    bool FrameIsBad(int first, int last) const;     // any undecodable frame in [first, last]

public:
    AnalyzerUnit(std::string EventID, std::string ImageDir, int CameraNumber, Trainer** TrainedData, std::string MaskDir,
                 Parser* parser);
    virtual ~AnalyzerUnit(void);

    int MatTrigFrame;
    int loc_thres;
    std::vector<double> entropies;                  // significance of every frame the last search stored

    void FindTriggerFrame(bool nonStopMode, int startframe);
    virtual void LocalizeOMatic(std::string) = 0;

    Trainer* TrainedData;                            // borrowed (the reference holds a deep copy)
    std::vector<bubble*> BubbleList;

    bool okToProceed = true;
    int TriggerFrameIdentificationStatus = 0;       // 0 ok | -3 no trigger | -9 malformed frame set

    int NumFrames() const { return (int)CameraFrames.size(); }
};

#endif
