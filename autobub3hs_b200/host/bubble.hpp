// bubble -- track record of one bubble (reference: bubble/bubble.hpp:22-82, bubble.cpp:34-118).
// Host glue: consumes the abub_blob records the localizer kernels emit.
#ifndef ABUB_BUBBLE_HPP
#define ABUB_BUBBLE_HPP

#include <utility>
#include <vector>

#include "Image8.hpp"

struct MomentsD {            // the three cv::Moments fields the path reads (L3Localizer.cpp:403-407)
    double m00 = 0, m10 = 0, m01 = 0;
};

struct BubbleImageFrame {    // bubble.hpp:22-31
    RectI newPosition;
    double ContArea = 0;
    double ContRadius = 0;
    MomentsD moments;
    Point2F MassCentres;
};

class bubble {
public:
    bubble(BubbleImageFrame bubbleGenesisStructure);
    ~bubble();

    std::vector<BubbleImageFrame> KnownDescriptors;
    float last_x;
    float last_y;
    RectI GenesisPosition;
    Point2F GenesisPositionCentroid;
    std::vector<float> dz;
    bool lockThisIteration;

    float dRdT(void);
    float dZdT(void);
    bool isNewPositionProbable(int& x, int& y);
    void printAllXY(void);
    void operator<<(BubbleImageFrame newPosImgFrameStructure);
};

#endif
