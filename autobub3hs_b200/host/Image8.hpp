// Image8 -- the 8-bit single-channel image that stands where the reference passes cv::Mat
// (CV_8U, continuous; Parser::GetImage output, RawParser.cpp:34-47).  OpenCV is not part
// of this build: every pixel operation of the path runs in libabub_cuda.so.
#ifndef ABUB_IMAGE8_HPP
#define ABUB_IMAGE8_HPP

#include <stdint.h>

#include <vector>

struct Image8 {
    int rows = 0, cols = 0;
    std::vector<uint8_t> data;          // rows*cols, row-major, no padding

    bool empty() const { return data.empty(); }
    void create(int r, int c) { rows = r; cols = c; data.assign((size_t)r * c, 0); }
    void release() { rows = cols = 0; data.clear(); }
    uint8_t at(int y, int x) const { return data[(size_t)y * cols + x]; }
    uint8_t* ptr(int y = 0) { return data.data() + (size_t)y * cols; }
    const uint8_t* ptr(int y = 0) const { return data.data() + (size_t)y * cols; }
};

// cv::Rect stand-in (bubble.hpp:24, L3Localizer.cpp:278)
struct RectI {
    int x = 0, y = 0, width = 0, height = 0;
};
struct Point2F {
    float x = 0.f, y = 0.f;
};

#endif
