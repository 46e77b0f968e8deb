// ZipParser -- run stored as <run>.zip (ParseFolder/ZipParser.cpp).  The reference reads the
// archive through minizip-ng; here a small central-directory reader (ZIP + ZIP64, stored and
// deflate entries, zlib inflate) does the same job.  throw -10 on an unreadable archive, as the
// reference (ZipParser.cpp:66-70).
#ifndef ABUB_ZIPPARSER_HPP
#define ABUB_ZIPPARSER_HPP

#include <stdio.h>

#include <map>
#include <memory>

#include "Parser.hpp"

struct ZipEntryLoc {
    uint64_t local_off = 0, comp_size = 0, uncomp_size = 0;
    int method = 0;
};

class ZipParser : public Parser {
public:
    ZipParser(std::string RunFolder, std::string ImageFolder, std::string ImageFormat);
    ~ZipParser() override;
    ZipParser* clone() override;
    int GetImage(std::string EventID, std::string FrameName, Image8& Image) override;
    int GetImageBytes(std::string EventID, std::string FrameName, std::vector<uint8_t>& Bytes) override;
    int FailureCode() const override { return -1; }
    void GetEventDirLists(std::vector<std::string>& EventList) override;
    void GetFileLists(const char* EventFolder, std::vector<std::string>& FileList, const char* camera_out_name) override;
    void ParseAndSortFramesInFolder(std::string EventID, int Camera, std::vector<std::string>& Contents) override;

private:
    void BuildFileList();
    bool ReadEntry(const ZipEntryLoc& loc, std::vector<uint8_t>& out);

    // The central-directory index is immutable once built and shared by the clones made after that (the reference
    // copies the lists into every clone, ZipParser.cpp: one copy per analyzer, i.e. per event and camera -- with 1.6 M
    // entries in a 10,000-event archive that copy alone would dominate the host side).
    struct Index {
        std::vector<std::string> FileContents;                                  // every entry name
        std::map<std::string, std::map<std::string, std::string>> ImageNames;   // [event][frame] -> entry name
        std::map<std::string, std::map<std::string, ZipEntryLoc>> ImageLocs;    // [event][frame] -> location
        uint64_t archive_size = 0;                                              // bytes of the archive file (entry sizes are checked against it)
    };
    FILE* fp = nullptr;
    std::shared_ptr<Index> idx = std::make_shared<Index>();
};

#endif
