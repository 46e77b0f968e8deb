// RawParser -- run stored as a directory tree <run>/<event>/<ImageFolder>/<frame> (ParseFolder/RawParser.cpp).
#ifndef ABUB_RAWPARSER_HPP
#define ABUB_RAWPARSER_HPP

#include "Parser.hpp"

class RawParser : public Parser {
public:
    RawParser(std::string RunFolder, std::string ImageFolder, std::string ImageFormat);
    ~RawParser() override;
    RawParser* clone() override;
    int GetImage(std::string EventID, std::string FrameName, Image8& Image) override;
    int GetImageBytes(std::string EventID, std::string FrameName, std::vector<uint8_t>& Bytes) override;
    int FailureCode() const override { return 0; }
    void GetEventDirLists(std::vector<std::string>& EventList) override;
    void GetFileLists(const char* EventFolder, std::vector<std::string>& FileList, const char* camera_out_name) override;
    void ParseAndSortFramesInFolder(std::string EventID, int Camera, std::vector<std::string>& Contents) override;
    bool ReadGray8BmpPixels(const std::string& EventID, const std::string& FrameName, int rows, int cols, uint8_t* dst,
                            bool* bottom_up) override;
    // regex_match(path, "^.*cam<Camera>.*(png|bmp)") of the reference (RawParser.cpp:145-150), written out
    static bool FramePathMatches(const std::string& path, int Camera);

private:
    std::string FramePath(const std::string& EventID, const std::string& FrameName) const;
};

#endif
