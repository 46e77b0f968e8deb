// OutputWriter -- abub3hs_<run>.txt (reference: PICOFormatWriter/PICOFormatWriterV4.hpp:15-62,
// PICOFormatWriterV4.cpp:19-302).  Same interface; the row text is byte-identical to the reference's
// stream output (fixed, precision 2; libstdc++ prints NaN as nan/-nan).
#ifndef ABUB_PICOWRITER_HPP
#define ABUB_PICOWRITER_HPP

#include <sstream>
#include <string>
#include <vector>

#include "bubble.hpp"

class OutputWriter {
private:
    int frameOffset;
    std::string OutputDir;
    std::string run_number;
    std::string abubOutFilename;
    std::stringstream _StreamOutput;
    int NumCams;

public:
    struct BubbleData {
        std::vector<bubble*> BubbleObjectData;
        int StatusCode = 0;
        int frame0 = 0;
        int event = 0;
        BubbleData() {}
    };
    std::vector<BubbleData*> AllBubbleData;

    OutputWriter(std::string OutDir, std::string run_number, int frameOffset, int NumCams);
    ~OutputWriter(void);

    void writeHeader(void);
    void stageCameraOutput(std::vector<bubble*> BubbleRectIn, int camera, int frame0, int event);
    void stageCameraOutputError(int camera, int error, int event);
    void formEachBubbleOutput(int camera, int& ibubImageStart, int nBubTotal);
    template <class Sink> void formRows(Sink& o, int camera, int& ibubImageStart, int nBubTotal) const;
    void writeCameraOutput(void);

    /* the staged rows of this event as text, without touching the file (multi-GPU ranks buffer their
     * event range and rank 0 concatenates in event order = the omp ordered file order) */
    std::string formatCameraOutput(void);
    /* SURVEY 8(f)-3, batch form: the same rows appended to `dst` by a to_chars formatter (no iostream, no locale, no
     * file access).  A driver formats every event of a staged batch into one buffer and opens/appends/closes the
     * output once per batch instead of once per event (PICOFormatWriterV4.cpp:299-301). */
    void appendCameraOutput(std::string& dst);
    /* open / append / close, once for however many events `rows` holds */
    static void appendToFile(const std::string& file, const std::string& rows);
    /* the two formatters on one value, for the self-test */
    static std::string formatValueFast(double v);
    static std::string formatValueStream(double v);
    static std::string headerText(void);
    const std::string& fileName() const { return abubOutFilename; }
};

#endif
