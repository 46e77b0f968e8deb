#include "PICOFormatWriterV4.hpp"

#include <charconv>
#include <cmath>
#include <cstdio>

#include <fstream>

#include "CommonParameters.h"

OutputWriter::OutputWriter(std::string OutDir, std::string run_number_, int frameOffset_, int NumCams_) {
    OutputDir = OutDir;
    run_number = run_number_;
    abubOutFilename = OutputDir + "abub3hs_" + run_number + ".txt";      // PICOFormatWriterV4.cpp:26
    frameOffset = frameOffset_;
    NumCams = NumCams_;
    for (int i = 0; i < NumCams; i++) AllBubbleData.push_back(new BubbleData());
}

OutputWriter::~OutputWriter(void) {
    for (BubbleData* d : AllBubbleData) delete d;
}

// column legend + printf template of one row (:54-88)
std::string OutputWriter::headerText(void) {
    std::ostringstream h;
    const int nt = NumFramesBubbleTrack;
    h << "Output of AutoBub v3 - the automatic unified bubble finder code by Pitam, using OpenCV.\n";
    h << "run  ev  ibubimage  TotalBub4CamImg  camera  frame0  hori  vert  GenesisW  GenesisH  dZdt  dRdt  ";
    h << "TrkFrame(" << nt << ")  TrkHori(" << nt << ")  TrkVert(" << nt << ")  ";
    h << "TrkBubW(" << nt << ")  TrkBubH(" << nt << ")  TrkBubRadius(" << nt << ")  FakeValue\n";
    h << "%12s  %5d  %d  %d  %d  %d  %.02f  %.02f  %d  %d  %.02f  %.02f  ";
    for (int j = 0; j < nt; j++) h << "%d  ";
    for (int block = 0; block < 5; block++)
        for (int j = 0; j < nt; j++) h << "%.02f  ";
    h << "%d";
    h << "\n8\n\n\n";
    return h.str();
}

void OutputWriter::writeHeader(void) {
    std::ofstream f(abubOutFilename);          // truncates: a run always starts a fresh file (:56)
    f << headerText();
}

// an empty bubble list stages the (transient) error -1 (:99-110)
void OutputWriter::stageCameraOutput(std::vector<bubble*> BubbleRectIn, int camera, int frame0, int event) {
    BubbleData* d = AllBubbleData[camera];
    d->StatusCode = BubbleRectIn.empty() ? -1 : 0;
    d->BubbleObjectData = BubbleRectIn;
    d->frame0 = frame0;
    d->event = event;
}

void OutputWriter::stageCameraOutputError(int camera, int error, int event) {
    AllBubbleData[camera]->StatusCode = error;
    AllBubbleData[camera]->event = event;
}

// Field sink of the batch formatter: what `stream << value` prints under precision(2) + fixed, produced with
// std::to_chars.  Integers: shortest decimal.  Floating point: the value as a double with two decimals, correctly rounded
// (what "%.2f" prints); non-finite values are left to snprintf so that "nan" / "-nan" / "inf" come out as the C library
// spells them.
namespace {
struct FastSink {
    std::string& s;
    explicit FastSink(std::string& dst) : s(dst) {}
    FastSink& operator<<(const char* t) { s.append(t); return *this; }
    FastSink& operator<<(const std::string& t) { s.append(t); return *this; }
    FastSink& operator<<(int v) {
        char buf[16];
        auto r = std::to_chars(buf, buf + sizeof(buf), v);
        s.append(buf, r.ptr);
        return *this;
    }
    FastSink& operator<<(double v) {
        char buf[64];
        if (std::isfinite(v)) {
            auto r = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::fixed, 2);
            if (r.ec == std::errc()) { s.append(buf, r.ptr); return *this; }
        }
        char big[400];
        const int n = snprintf(big, sizeof(big), "%.2f", v);
        s.append(big, (size_t)n);
        return *this;
    }
    FastSink& operator<<(float v) { return *this << (double)v; }
};
}  // namespace

template <class Sink>
void OutputWriter::formRows(Sink& o, int camera, int& ibubImageStart, int nBubTotal) const {
    const BubbleData* wd = AllBubbleData[camera];
    const int nt = NumFramesBubbleTrack;
    const char* sep = "  ";

    if (wd->StatusCode != 0) {                 // error row: zeros, the code in the frame0 column (:146-174)
        o << run_number << sep << wd->event << sep << 0 << sep << 0 << sep << camera << sep << wd->StatusCode << sep
          << 0.0 << sep << 0.0 << sep << 0 << sep << 0 << sep << 0.0 << sep << 0.0 << sep;
        for (int j = 0; j < nt; j++) o << 0 << sep;
        for (int j = 0; j < 5 * nt; j++) o << 0.0 << sep;
        o << "1  \n";
        return;
    }

    for (size_t i = 0; i < wd->BubbleObjectData.size(); i++) {
        bubble* b = wd->BubbleObjectData[i];
        const int f0 = wd->frame0 + frameOffset;
        o << run_number << sep << wd->event << sep << ibubImageStart + (int)i << sep << nBubTotal << sep << camera << sep;
        o << f0 << sep;
        const float width = (float)b->GenesisPosition.width, height = (float)b->GenesisPosition.height;
        const float x = b->GenesisPositionCentroid.x, y = b->GenesisPositionCentroid.y;
        const float dzdt = b->dZdT(), drdt = b->dRdT();
        o << x << sep << y << sep << (int)width << sep << (int)height << sep << dzdt << sep << drdt << sep;

        const int tracked = (int)b->KnownDescriptors.size() - 1;
        const int excess = nt > tracked ? nt - tracked : 0;         // columns padded with -1 (:228-272)
        for (int j = 1; j <= tracked; j++) o << f0 + j << sep;
        for (int j = 0; j < excess; j++) o << f0 + tracked + j << sep;
        for (int j = 1; j <= tracked; j++) o << b->KnownDescriptors[j].MassCentres.x << sep;
        for (int j = 0; j < excess; j++) o << -1 << sep;
        for (int j = 1; j <= tracked; j++) o << b->KnownDescriptors[j].MassCentres.y << sep;
        for (int j = 0; j < excess; j++) o << -1 << sep;
        for (int j = 1; j <= tracked; j++) o << b->KnownDescriptors[j].newPosition.width << sep;
        for (int j = 0; j < excess; j++) o << -1 << sep;
        for (int j = 1; j <= tracked; j++) o << b->KnownDescriptors[j].newPosition.height << sep;
        for (int j = 0; j < excess; j++) o << -1 << sep;
        for (int j = 1; j <= tracked; j++) o << b->KnownDescriptors[j].ContRadius << sep;
        for (int j = 0; j < excess; j++) o << -1 << sep;
        o << "1  \n";
    }
    ibubImageStart += (int)wd->BubbleObjectData.size();
}

void OutputWriter::formEachBubbleOutput(int camera, int& ibubImageStart, int nBubTotal) {
    std::stringstream& o = _StreamOutput;
    o.clear();
    o.precision(2);
    o.setf(std::ios::fixed, std::ios::floatfield);
    formRows(o, camera, ibubImageStart, nBubTotal);
}

void OutputWriter::appendCameraOutput(std::string& dst) {
    int ibubImageStart = 1;
    int nBubTotal = 0;
    for (int i = 0; i < NumCams; i++)
        nBubTotal += AllBubbleData[i]->StatusCode != 0 ? 0 : (int)AllBubbleData[i]->BubbleObjectData.size();
    FastSink sink(dst);
    for (int i = 0; i < NumCams; i++) formRows(sink, i, ibubImageStart, nBubTotal);
}

void OutputWriter::appendToFile(const std::string& file, const std::string& rows) {
    std::ofstream f(file, std::fstream::out | std::fstream::app);
    f.write(rows.data(), (std::streamsize)rows.size());
}

std::string OutputWriter::formatValueFast(double v) {
    std::string s;
    FastSink k(s);
    k << v;
    return s;
}

std::string OutputWriter::formatValueStream(double v) {
    std::stringstream o;
    o.precision(2);
    o.setf(std::ios::fixed, std::ios::floatfield);
    o << v;
    return o.str();
}

std::string OutputWriter::formatCameraOutput(void) {
    int ibubImageStart = 1;
    int nBubTotal = 0;
    for (int i = 0; i < NumCams; i++)
        nBubTotal += AllBubbleData[i]->StatusCode != 0 ? 0 : (int)AllBubbleData[i]->BubbleObjectData.size();
    _StreamOutput.str("");
    for (int i = 0; i < NumCams; i++) formEachBubbleOutput(i, ibubImageStart, nBubTotal);
    return _StreamOutput.str();
}

// rows of every camera of the event, appended (:287-302)
void OutputWriter::writeCameraOutput(void) {
    const std::string rows = formatCameraOutput();
    std::ofstream f(abubOutFilename, std::fstream::out | std::fstream::app);
    f << rows;
}
