/* oracle/refshim/ref_harness.cpp -- TEST INFRASTRUCTURE: calls single functions of the reference's own object code
 * (AnalyzerUnit.cpp, AlgorithmTraining/Trainer.cpp compiled unmodified from /root/reference against oracle/refshim)
 * so that the golden vectors of tests/golden can be checked against them:
 *
 *   ref_harness significance <in> <out>   replays a recorded sequence of AnalyzerUnit::calculateSignificanceFrame calls
 *                                         (AnalyzerUnit.cpp:435-504, with CalcMean / CalcStdDev :514-532) on images that
 *                                         have exactly the given 256-bin histograms
 *   ref_harness welford <in> <out>        Trainer::CalculateMeanSigmaImageVector (Trainer.cpp:144-216) on a frame stack
 *   ref_harness gate <in> <out>           Trainer::calculateEntropyFrame (Trainer.cpp:341-376) on frame1 - frame0
 *
 * The two private member functions are reached by compiling THIS file with -Dprivate=public; the reference's own
 * translation units are compiled as they are.  Binary little-endian files, layouts below. */
#include <opencv2/opencv.hpp>      /* the standard headers and the shim first: only the reference's classes are opened up */
#include <dirent.h>
#include <functional>
#define private public
#define protected public
#include "AnalyzerUnit.hpp"
#include "AlgorithmTraining/Trainer.hpp"
#undef private
#undef protected

#include <stdio.h>

namespace {

class NullParser : public Parser {
public:
    NullParser() : Parser("", "", "cam%d_image%u.png") {}
    ~NullParser() override {}
    Parser* clone() override { return new NullParser(); }
    int GetImage(std::string, std::string, cv::Mat&) override { return -1; }
    void GetEventDirLists(std::vector<std::string>&) override {}
    void GetFileLists(const char*, std::vector<std::string>&, const char*) override {}
    void ParseAndSortFramesInFolder(std::string, int, std::vector<std::string>&) override {}
    void GetRunFileInfo(std::vector<std::string>&) override {}
};

class BareAnalyzer : public AnalyzerUnit {
public:
    BareAnalyzer(Trainer** t) : AnalyzerUnit("0", "", 0, t, "", new NullParser()) {}
    void LocalizeOMatic(std::string) override {}
};

template <typename T> bool rd(FILE* f, T* v, size_t n = 1) { return fread(v, sizeof(T), n, f) == n; }
template <typename T> void wr(FILE* f, const T* v, size_t n = 1) { fwrite(v, sizeof(T), n, f); }

/* in: int32 ncases; per case: int32 tss, int32 ncalls; per call: int32 store, int32 total, uint32 hist[256]
 * out: per call: double significance, int32 loc_thres (AnalyzerUnit::loc_thres after the call)
 * A new analyzer per case = the state FindTriggerFrame starts from at startframe 1 (AnalyzerUnit.cpp:165-169). */
int significance(FILE* in, FILE* out) {
    int32_t ncases;
    if (!rd(in, &ncases)) return 2;
    for (int c = 0; c < ncases; c++) {
        int32_t tss, ncalls;
        if (!rd(in, &tss) || !rd(in, &ncalls)) return 2;
        Trainer* tr = new Trainer(0, std::vector<std::string>(), "", "cam%d_image%u.png", "/Images/", new NullParser());
        tr->TrainingSetSize = tss;
        BareAnalyzer a(&tr);
        for (int k = 0; k < ncalls; k++) {
            int32_t store, total;
            uint32_t h[256];
            if (!rd(in, &store) || !rd(in, &total) || !rd(in, h, 256)) return 2;
            cv::Mat img(1, total, CV_8U);
            size_t o = 0;
            for (int v = 0; v < 256; v++)
                for (uint32_t i = 0; i < h[v] && o < (size_t)total; i++) img.data[o++] = (cv::uchar)v;
            if (o != (size_t)total) return 3;
            const double s = a.calculateSignificanceFrame(img, store != 0, false);
            const int32_t lt = a.loc_thres;
            wr(out, &s);
            wr(out, &lt);
        }
        delete tr;
    }
    return 0;
}

/* in: int32 n, rows, cols; n frames of rows*cols bytes.  out: mean plane, sigma plane, int32 TrainingSetSize */
int welford(FILE* in, FILE* out) {
    int32_t n, rows, cols;
    if (!rd(in, &n) || !rd(in, &rows) || !rd(in, &cols)) return 2;
    std::vector<cv::Mat> stack;
    for (int i = 0; i < n; i++) {
        cv::Mat m(rows, cols, CV_8U);
        if (!rd(in, m.data, (size_t)rows * cols)) return 2;
        stack.push_back(m);
    }
    Trainer tr(0, std::vector<std::string>(), "", "cam%d_image%u.png", "/Images/", new NullParser());
    cv::Mat mean, sigma;
    tr.CalculateMeanSigmaImageVector(stack, mean, sigma);
    wr(out, mean.data, (size_t)rows * cols);
    wr(out, sigma.data, (size_t)rows * cols);
    const int32_t tss = tr.TrainingSetSize;
    wr(out, &tss);
    return 0;
}

/* in: int32 n, rows, cols; n pairs (frame0, frame1).  out: n floats = calculateEntropyFrame(frame1 - frame0), Trainer.cpp:279-280 */
int gate(FILE* in, FILE* out) {
    int32_t n, rows, cols;
    if (!rd(in, &n) || !rd(in, &rows) || !rd(in, &cols)) return 2;
    Trainer tr(0, std::vector<std::string>(), "", "cam%d_image%u.png", "/Images/", new NullParser());
    for (int i = 0; i < n; i++) {
        cv::Mat f0(rows, cols, CV_8U), f1(rows, cols, CV_8U);
        if (!rd(in, f0.data, (size_t)rows * cols) || !rd(in, f1.data, (size_t)rows * cols)) return 2;
        cv::Mat d = f1 - f0;
        const float e = tr.calculateEntropyFrame(d);
        wr(out, &e);
    }
    return 0;
}

}  // namespace

int main(int argc, char** argv) {
    if (argc != 4) { fprintf(stderr, "usage: ref_harness significance|welford|gate <in> <out>\n"); return 1; }
    FILE* in = fopen(argv[2], "rb");
    FILE* out = fopen(argv[3], "wb");
    if (!in || !out) return 1;
    const std::string cmd = argv[1];
    int rc = 1;
    if (cmd == "significance") rc = significance(in, out);
    else if (cmd == "welford") rc = welford(in, out);
    else if (cmd == "gate") rc = gate(in, out);
    fclose(in);
    fclose(out);
    return rc;
}
