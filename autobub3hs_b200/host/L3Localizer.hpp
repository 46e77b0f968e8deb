// L3Localizer -- blob extraction on the trigger frame and the post-trigger frames (reference:
// BubbleLocalizer/L3Localizer.hpp:27-108, L3Localizer.cpp:215-460, :764-869, :881-1012).
// The pixel work (ProcessFrame, blur, TOZERO, Otsu, external TC89_L1 contours, boundingRect,
// contourArea, moments) is one abub_localize call (K5-K7); the mask tests, the BoxArea filters and
// the bubble tracking below are the reference's host logic.
#ifndef ABUB_L3LOCALIZER_HPP
#define ABUB_L3LOCALIZER_HPP

#include "../../include/abub_cuda.h"
#include "AnalyzerUnit.hpp"

class L3Localizer : public AnalyzerUnit {
private:
    bool nonStopMode;

    static BubbleImageFrame FrameFromBlob(const abub_blob& b, bool genesis);
    const Image8* cam_mask_ = nullptr;
    const Image8* bellows_mask_ = nullptr;
    bool cam_mask_looked_ = false, bellows_mask_looked_ = false;
    int preTrigNumber = 0;                                           // frame used as preTrigFrame (:932-934)
    void GenesisFromBlobs(const abub_blob* blobs, int n);            // CalculateInitialBubbleParams bookkeeping
    bool BellowsVeto(std::vector<abub_blob>& again);                 // block :292-372; false = template not loadable
    void TrackFromBlobs(const abub_blob* blobs, int n);              // CalculatePostTriggerFrameParams bookkeeping

public:
    L3Localizer(std::string EventID, std::string ImageDir, int CameraNumber, bool nonStopPref, Trainer** TrainedData,
                std::string MaskDir, Parser* parser);
    ~L3Localizer();

    int numBubbleMultiplicity = 0;
    bool Level1SuspicionFlag;

    void LocalizeOMatic(std::string) override;
    bool isInMask(RectI* genesis_coords, bool bellows = false);
    void printBubbleList(void);
};

#endif
