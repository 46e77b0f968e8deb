/* cv:: free functions of oracle/refshim (see opencv.hpp).  Live ones are implemented in cvshim.cpp on top of
 * oracle/abub_ref.c; the ones only dead or debug code of the reference reaches abort. */
#ifndef ABUB_REFSHIM_FUNCS_HPP
#define ABUB_REFSHIM_FUNCS_HPP

namespace cv {

typedef std::string String;

/* ---- imgcodecs: 8-bit BMP (any palette -> grey through BT.601 fixed point as cv::imread does) and 8-bit grey PNG ---- */
Mat imread(const std::string& path, int flags = IMREAD_COLOR);
bool imwrite(const std::string& path, const Mat& img);                 /* debug dumps of the reference: discarded */

/* ---- imgproc (live) ---- */
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT);
void blur(InputArray src, OutputArray dst, Size ksize, Point anchor = Point(-1, -1), int borderType = BORDER_DEFAULT);
void absdiff(InputArray a, InputArray b, OutputArray dst);
double threshold(InputArray src, OutputArray dst, double thresh, double maxval, int type);
void calcHist(const Mat* images, int nimages, const int* channels, InputArray mask, OutputArray hist, int dims, const int* histSize,
              const float** ranges, bool uniform = true, bool accumulate = false);
void findContours(InputArray image, std::vector<std::vector<Point> >& contours, int mode, int method, Point offset = Point());
Rect boundingRect(const std::vector<Point>& pts);
double contourArea(const std::vector<Point>& contour, bool oriented = false);
Moments moments(const std::vector<Point>& contour, bool binaryImage = false);
void matchTemplate(InputArray image, InputArray templ, OutputArray result, int method, InputArray mask = Mat());
void normalize(InputArray src, InputOutputArray dst, double alpha = 1, double beta = 0, int norm_type = NORM_L2, int dtype = -1,
               InputArray mask = Mat());
void minMaxLoc(InputArray src, double* minVal, double* maxVal = 0, Point* minLoc = 0, Point* maxLoc = 0, InputArray mask = Mat());
void rectangle(Mat& img, Rect rec, const Scalar& color, int thickness = 1, int lineType = LINE_8, int shift = 0);   /* drawing: no effect on results */
void rectangle(Mat& img, Point pt1, Point pt2, const Scalar& color, int thickness = 1, int lineType = LINE_8, int shift = 0);

/* ---- reached from dead or debug code only ---- */
inline void cvtColor(InputArray, OutputArray, int, int = 0) { shim_unsupported("cvtColor"); }
inline void LUT(InputArray, InputArray, OutputArray) { shim_unsupported("LUT"); }
inline RotatedRect minAreaRect(const std::vector<Point>&) { shim_unsupported("minAreaRect"); }
inline RotatedRect fitEllipse(const std::vector<Point>&) { shim_unsupported("fitEllipse"); }
inline void ellipse(Mat&, const RotatedRect&, const Scalar&, int = 1, int = LINE_8) {}
inline void ellipse(Mat&, Point, Size, double, double, double, const Scalar&, int = 1, int = LINE_8, int = 0) {}
inline void circle(Mat&, Point, int, const Scalar&, int = 1, int = LINE_8, int = 0) {}
inline void line(Mat&, Point, Point, const Scalar&, int = 1, int = LINE_8, int = 0) {}
inline void drawContours(Mat&, const std::vector<std::vector<Point> >&, int, const Scalar&, int = 1, int = LINE_8) {}
inline void namedWindow(const std::string&, int = WINDOW_AUTOSIZE) { shim_unsupported("namedWindow"); }
inline void imshow(const std::string&, InputArray) { shim_unsupported("imshow"); }
inline int waitKey(int = 0) { shim_unsupported("waitKey"); }
inline void bitwise_not(InputArray, OutputArray) { shim_unsupported("bitwise_not"); }
inline void equalizeHist(InputArray, OutputArray) { shim_unsupported("equalizeHist"); }
inline void medianBlur(InputArray, OutputArray, int) { shim_unsupported("medianBlur"); }
inline void Canny(InputArray, OutputArray, double, double, int = 3, bool = false) { shim_unsupported("Canny"); }
inline void dilate(InputArray, OutputArray, InputArray, Point = Point(-1, -1), int = 1) { shim_unsupported("dilate"); }
inline void erode(InputArray, OutputArray, InputArray, Point = Point(-1, -1), int = 1) { shim_unsupported("erode"); }
inline void addWeighted(InputArray, double, InputArray, double, double, OutputArray, int = -1) { shim_unsupported("addWeighted"); }
inline void subtract(InputArray a, InputArray b, OutputArray dst) { dst.ref() = a - b; }
inline Scalar mean(InputArray) { shim_unsupported("mean"); }
inline Scalar sum(InputArray) { shim_unsupported("sum"); }

}  // namespace cv

#endif
