/* cv::Mat stand-in of oracle/refshim (see opencv.hpp): reference-counted 2-D array of CV_8U or CV_32F, one channel,
 * with ROI views that share the buffer -- the subset the reference's live code uses. */
#ifndef ABUB_REFSHIM_MAT_HPP
#define ABUB_REFSHIM_MAT_HPP

namespace cv {

class Mat {
public:
    int rows = 0, cols = 0;
    size_t step = 0;                 /* bytes per row of the underlying buffer */
    uchar* data = nullptr;
    int flags = CV_8U;               /* depth | ((channels-1) << 3) */

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(Size s, int type) { create(s.height, s.width, type); }
    Mat(const Mat& m, const Rect& roi) { *this = m; adjust(roi); }
    Mat(const Mat&) = default;
    Mat& operator=(const Mat&) = default;

    void create(int r, int c, int type) {
        if (data && r == rows && c == cols && type == flags) return;   /* as cv::Mat::create: a view of the right shape is kept */
        flags = type;
        rows = r; cols = c;
        step = (size_t)c * elemSize();
        buf = std::shared_ptr<uchar>(new uchar[std::max<size_t>(step * (size_t)r, 1)], std::default_delete<uchar[]>());
        data = buf.get();
    }
    void create(Size s, int type) { create(s.height, s.width, type); }
    void release() { buf.reset(); data = nullptr; rows = cols = 0; step = 0; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t total() const { return (size_t)rows * cols; }
    int type() const { return flags; }
    int depth() const { return flags & 7; }
    int channels() const { return (flags >> 3) + 1; }
    size_t elemSize() const { return (size_t)((flags & 7) == CV_32F ? 4 : 1) * ((flags >> 3) + 1); }
    Size size() const { return Size(cols, rows); }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }

    static Mat zeros(int r, int c, int type) {
        Mat m(r, c, type);
        if (m.data) memset(m.data, 0, m.step * (size_t)r);
        return m;
    }
    static Mat zeros(Size s, int type) { return zeros(s.height, s.width, type); }

    Mat clone() const {
        Mat m;
        copyToImpl(m);
        return m;
    }
    void copyToImpl(Mat& dst) const {
        if (empty()) { dst.release(); return; }
        dst.create(rows, cols, flags);
        for (int y = 0; y < rows; y++) memmove(dst.data + (size_t)y * dst.step, data + (size_t)y * step, (size_t)cols * elemSize());
    }
    void copyTo(Mat& dst) const { copyToImpl(dst); }
    void copyTo(Mat&& dst) const { Mat d = dst; copyToImpl(d); }          /* a ROI temporary: written through the shared buffer */

    Mat operator()(const Rect& roi) const { return Mat(*this, roi); }

    template <typename T> T& at(int r, int c) { return *reinterpret_cast<T*>(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const { return *reinterpret_cast<const T*>(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    /* single index: element i of a single-row or single-column array */
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
    template <typename T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }

    void convertTo(Mat& dst, int rtype, double alpha = 1, double beta = 0) const;

private:
    std::shared_ptr<uchar> buf;
    /* cv::Mat(m, roi) asserts 0 <= roi.x, roi.x + roi.width <= m.cols, ... and throws cv::Exception otherwise */
    void adjust(const Rect& roi);
};

/* what cv::Exception is to the reference: something derived from std::exception (caught at AutoBubStart3.cpp:114) */
struct Exception : public std::exception {
    std::string msg;
    explicit Exception(const std::string& m) : msg(m) {}
    const char* what() const noexcept override { return msg.c_str(); }
};

inline void Mat::adjust(const Rect& roi) {
    if (!(0 <= roi.x && 0 <= roi.width && roi.x + roi.width <= cols && 0 <= roi.y && 0 <= roi.height && roi.y + roi.height <= rows))
        throw Exception("(-215:Assertion failed) 0 <= roi.x && 0 <= roi.width && roi.x + roi.width <= m.cols && 0 <= roi.y && "
                        "0 <= roi.height && roi.y + roi.height <= m.rows in function 'Mat'");
    data += (size_t)roi.y * step + (size_t)roi.x * elemSize();
    rows = roi.height;
    cols = roi.width;
}

/* Input/output argument wrappers: a named Mat is written in place (and may be reallocated), a temporary (a ROI such as
 * diff_frame(ROI)) is written through the buffer it shares. */
struct _OutputArray {
    mutable Mat tmp;
    Mat* p;
    _OutputArray(Mat& m) : p(&m) {}
    _OutputArray(Mat&& m) : tmp(m), p(&tmp) {}
    Mat& ref() const { return *p; }
};
typedef const _OutputArray& OutputArray;
typedef const _OutputArray& InputOutputArray;
typedef const Mat& InputArray;

/* ---- the MatExpr arithmetic the reference writes: all evaluated eagerly, element-wise, with OpenCV's saturation ---- */
Mat operator-(const Mat& a, const Mat& b);        /* CV_8U: saturate_cast<uchar>(a - b) */
Mat operator+(const Mat& a, const Mat& b);
Mat operator*(double s, const Mat& a);            /* CV_8U: saturate_cast<uchar>(cvRound(a * s)) */
Mat operator*(const Mat& a, double s);
Mat operator/(const Mat& a, double s);            /* CV_32F: a * (float)(1 / s), as convertTo applies a MatExpr scale */
Mat& operator-=(Mat& a, const Mat& b);
Mat& operator+=(Mat& a, const Mat& b);
Mat& operator*=(Mat& a, double s);

}  // namespace cv

#endif
