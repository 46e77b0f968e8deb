/* oracle/refshim/opencv2/opencv.hpp -- TEST INFRASTRUCTURE (never part of the product).
 *
 * A minimal stand-in for the OpenCV C++ API, just large enough to compile the UNMODIFIED first-party sources of the
 * reference (picoexperiment/AutoBub3hs: AutoBubStart3.cpp, AnalyzerUnit.cpp, AlgorithmTraining/Trainer.cpp,
 * BubbleLocalizer/L3Localizer.cpp, bubble/bubble.cpp, PICOFormatWriter/PICOFormatWriterV4.cpp, ParseFolder/Parser.cpp,
 * ParseFolder/RawParser.cpp, FrameSorter.cpp, common/UtilityFunctions.cpp) where they lie under /root/reference into
 * oracle/_ref/abub3hs_ref (oracle/Makefile, target _ref).  OpenCV C++ headers and Boost are not in this image; the
 * reference's own arithmetic (significance, trigger state machine, Welford, entropy gate, tracking, dZdT/dRdT, the
 * writer's formatting, the event loop and its error codes) is thereby executed from the reference's own code.
 *
 * What stands behind the cv:: calls here is NOT OpenCV: it is oracle/abub_ref.c (the cv2-free restatement of OpenCV
 * 4.13.0's observable behaviour for GaussianBlur 5x5, blur 3x3, threshold TOZERO / BINARY|OTSU, findContours
 * RETR_EXTERNAL + CHAIN_APPROX_TC89_L1, boundingRect, contourArea, moments -- fuzzed against cv2 in
 * tests/test_oracle_cpu.py) plus the element-wise saturating arithmetic written out in cvshim.cpp.  OpenCV-internal
 * arithmetic therefore stays pinned to cv2 4.13.0, not to the reference (which does not pin an OpenCV version).
 * Functions reached only from the reference's dead or debug code (imshow, ellipse, minAreaRect, cvtColor, LUT ...) are
 * declared and abort when called.
 */
#ifndef ABUB_REFSHIM_OPENCV_HPP
#define ABUB_REFSHIM_OPENCV_HPP

#include <assert.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>
#include <memory>
#include <string>
#include <vector>

#define CV_VERSION_MAJOR 4
#define CV_8U 0
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_32F 5
#define CV_32FC1 5
#define CV_TM_SQDIFF 0
#define CV_TM_SQDIFF_NORMED 1
#define CV_TM_CCORR 2
#define CV_TM_CCORR_NORMED 3
#define CV_TM_CCOEFF 4
#define CV_TM_CCOEFF_NORMED 5

namespace cv {

typedef unsigned char uchar;

[[noreturn]] void shim_unsupported(const char* what);

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <typename U> Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
    Point_& operator+=(const Point_& o) { x += o.x; y += o.y; return *this; }
    Point_& operator-=(const Point_& o) { x -= o.x; y -= o.y; return *this; }
    template <typename S> Point_& operator*=(S s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
template <typename T> Point_<T> operator+(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(a.x + b.x, a.y + b.y); }
template <typename T> Point_<T> operator-(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(a.x - b.x, a.y - b.y); }
/* OpenCV: Point_<T> * float -> saturate_cast<T>(a.x * b); for T = float that is the float product */
inline Point_<float> operator*(const Point_<float>& a, float b) { return Point_<float>(a.x * b, a.y * b); }
inline Point_<float> operator*(float b, const Point_<float>& a) { return Point_<float>(a.x * b, a.y * b); }
inline Point_<float> operator*(const Point_<float>& a, double b) { return Point_<float>((float)(a.x * b), (float)(a.y * b)); }
template <typename T> std::ostream& operator<<(std::ostream& o, const Point_<T>& p) { return o << "[" << p.x << ", " << p.y << "]"; }
typedef Point_<int> Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;

template <typename T> struct Size_ {
    T width, height;
    Size_() : width(0), height(0) {}
    Size_(T w, T h) : width(w), height(h) {}
    T area() const { return width * height; }
};
typedef Size_<int> Size;
typedef Size_<float> Size2f;

template <typename T> struct Rect_ {
    T x, y, width, height;
    Rect_() : x(0), y(0), width(0), height(0) {}
    Rect_(T x_, T y_, T w, T h) : x(x_), y(y_), width(w), height(h) {}
    T area() const { return width * height; }
    Point_<T> tl() const { return Point_<T>(x, y); }
    Point_<T> br() const { return Point_<T>(x + width, y + height); }
};
template <typename T> Rect_<T> operator&(const Rect_<T>& a, const Rect_<T>& b) {
    T x1 = std::max(a.x, b.x), y1 = std::max(a.y, b.y);
    T w = std::min(a.x + a.width, b.x + b.width) - x1, h = std::min(a.y + a.height, b.y + b.height) - y1;
    if (w <= 0 || h <= 0) return Rect_<T>();
    return Rect_<T>(x1, y1, w, h);
}
template <typename T> Rect_<T> operator|(const Rect_<T>& a, const Rect_<T>& b) {
    if (a.width <= 0 || a.height <= 0) return b;
    if (b.width <= 0 || b.height <= 0) return a;
    T x1 = std::min(a.x, b.x), y1 = std::min(a.y, b.y);
    return Rect_<T>(x1, y1, std::max(a.x + a.width, b.x + b.width) - x1, std::max(a.y + a.height, b.y + b.height) - y1);
}
typedef Rect_<int> Rect;

struct Scalar {
    double val[4];
    Scalar() { val[0] = val[1] = val[2] = val[3] = 0; }
    Scalar(double a, double b = 0, double c = 0, double d = 0) { val[0] = a; val[1] = b; val[2] = c; val[3] = d; }
    double operator[](int i) const { return val[i]; }
};

struct RotatedRect {
    Point2f center;
    Size2f size;
    float angle = 0;
    RotatedRect() {}
    RotatedRect(const Point2f& c, const Size2f& s, float a) : center(c), size(s), angle(a) {}
    Rect boundingRect() const { shim_unsupported("RotatedRect::boundingRect"); }
};

struct Moments {
    double m00 = 0, m10 = 0, m01 = 0, m20 = 0, m11 = 0, m02 = 0, m30 = 0, m21 = 0, m12 = 0, m03 = 0;
    double mu20 = 0, mu11 = 0, mu02 = 0, mu30 = 0, mu21 = 0, mu12 = 0, mu03 = 0;
    double nu20 = 0, nu11 = 0, nu02 = 0, nu30 = 0, nu21 = 0, nu12 = 0, nu03 = 0;
};

enum { THRESH_BINARY = 0, THRESH_BINARY_INV = 1, THRESH_TRUNC = 2, THRESH_TOZERO = 3, THRESH_TOZERO_INV = 4, THRESH_OTSU = 8 };
enum { RETR_EXTERNAL = 0, RETR_LIST = 1, RETR_CCOMP = 2, RETR_TREE = 3 };
enum { CHAIN_APPROX_NONE = 1, CHAIN_APPROX_SIMPLE = 2, CHAIN_APPROX_TC89_L1 = 3, CHAIN_APPROX_TC89_KCOS = 4 };
enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4, NORM_MINMAX = 32 };
enum { IMREAD_UNCHANGED = -1, IMREAD_GRAYSCALE = 0, IMREAD_COLOR = 1 };
enum { COLOR_BGR2GRAY = 6, COLOR_GRAY2BGR = 8, COLOR_RGB2GRAY = 7 };
enum { WINDOW_NORMAL = 0, WINDOW_AUTOSIZE = 1 };
enum { TM_SQDIFF = 0, TM_SQDIFF_NORMED = 1, TM_CCORR = 2, TM_CCORR_NORMED = 3, TM_CCOEFF = 4, TM_CCOEFF_NORMED = 5 };
enum { BORDER_DEFAULT = 4, BORDER_REFLECT_101 = 4 };
enum { LINE_8 = 8 };

}  // namespace cv

#include "shim_mat.hpp"
#include "shim_funcs.hpp"

#endif
