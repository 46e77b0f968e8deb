// Parser -- image source abstraction of the reference (ParseFolder/Parser.hpp:10-34) with
// cv::Mat replaced by Image8.  Same method names, argument meaning and return conventions:
//   RawParser::GetImage returns 1 ok / 0 failed (RawParser.cpp:46: `!Image.empty()`),
//   ZipParser::GetImage returns -1 when the entry is missing or undecodable, else 1
//   (ZipParser.cpp:198,227,237) -- callers test `== -1` (AnalyzerUnit.cpp:208, Trainer.cpp:262).
// One parser instance per analyzer/trainer (clone()): file handles are not shared.
#ifndef ABUB_PARSER_HPP
#define ABUB_PARSER_HPP

#include <string>
#include <vector>

#include "Image8.hpp"

class Parser {
protected:
    std::string RunFolder;
    std::string ImageFolder;
    std::string ImageFormat;
    int StatusCode = 0;

public:
    Parser(std::string run_folder, std::string image_folder, std::string image_format)
        : RunFolder(run_folder), ImageFolder(image_folder), ImageFormat(image_format) {}
    virtual ~Parser() {}

    virtual Parser* clone() = 0;
    virtual int GetImage(std::string EventID, std::string FrameName, Image8& Image) = 0;
    virtual void GetEventDirLists(std::vector<std::string>& EventList) = 0;
    virtual void GetFileLists(const char* EventFolder, std::vector<std::string>& FileList, const char* camera_out_name) = 0;
    virtual void ParseAndSortFramesInFolder(std::string EventID, int Camera, std::vector<std::string>& Contents) = 0;

    // Not in the reference: the encoded bytes of a frame, so that the caller can decode straight into
    // the pinned staging buffer of a device slot (abub_pinned_acquire).  Same return convention as
    // GetImage of the concrete class.
    virtual int GetImageBytes(std::string EventID, std::string FrameName, std::vector<uint8_t>& Bytes) = 0;
    // Not in the reference: fast path for the DAQ's raw frames.  When the frame is an uncompressed 8-bit BMP of exactly
    // rows x cols with the identity grey palette (and cols % 4 == 0, so that the file's row pitch is cols), its pixel
    // array is read with one positional read into dst (rows*cols bytes, e.g. the pinned staging region) in FILE order and
    // *bottom_up says whether the rows still have to be turned over (abub_event_flip_rows does that on the device).
    // false = not such a file / not supported by this parser: use GetImageBytes + decode.
    virtual bool ReadGray8BmpPixels(const std::string& EventID, const std::string& FrameName, int rows, int cols, uint8_t* dst,
                                    bool* bottom_up) {
        (void)EventID; (void)FrameName; (void)rows; (void)cols; (void)dst; (void)bottom_up;
        return false;
    }
    // value GetImage returns for "could not produce this frame"
    virtual int FailureCode() const = 0;
};

#endif
