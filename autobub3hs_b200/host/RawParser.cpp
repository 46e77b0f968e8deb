// RawParser (reference: ParseFolder/RawParser.cpp:28-157) on std::filesystem instead of Boost (the one regex is written out).
#include "RawParser.hpp"

#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <filesystem>

#include "ImageCodec.hpp"

namespace fs = std::filesystem;

RawParser::RawParser(std::string RunFolder, std::string ImageFolder, std::string ImageFormat)
    : Parser(RunFolder, ImageFolder, ImageFormat) {}
RawParser::~RawParser() {}

RawParser* RawParser::clone() { return new RawParser(RunFolder, ImageFolder, ImageFormat); }

std::string RawParser::FramePath(const std::string& EventID, const std::string& FrameName) const {
    // boost::filesystem operator/ : an absolute right-hand side ("/Images/") replaces the path in newer
    // Boost but appends in the Boost the reference was written against; the run layout needs "append".
    std::string p = RunFolder;
    auto join = [&p](const std::string& part) {
        size_t a = 0;
        while (a < part.size() && part[a] == '/') a++;
        if (!p.empty() && p.back() != '/') p += '/';
        p += part.substr(a);
    };
    join(EventID);
    join(ImageFolder);
    join(FrameName);
    return p;
}

int RawParser::GetImageBytes(std::string EventID, std::string FrameName, std::vector<uint8_t>& Bytes) {
    return abubcodec::read_file(FramePath(EventID, FrameName), Bytes) ? 1 : 0;
}

// One open, one 1078-byte read (file header, info header, palette) and one positional read of the pixel array straight
// into the caller's buffer: no intermediate copy of the file and no row copies.
bool RawParser::ReadGray8BmpPixels(const std::string& EventID, const std::string& FrameName, int rows, int cols, uint8_t* dst,
                                   bool* bottom_up) {
    if ((cols & 3) != 0) return false;
    const int fd = ::open(FramePath(EventID, FrameName).c_str(), O_RDONLY | O_CLOEXEC);
    if (fd < 0) return false;
    uint8_t h[14 + 40 + 1024];
    bool ok = ::pread(fd, h, sizeof(h), 0) == (ssize_t)sizeof(h) && h[0] == 'B' && h[1] == 'M';
    auto rd32 = [&h](int o) { return (uint32_t)h[o] | ((uint32_t)h[o + 1] << 8) | ((uint32_t)h[o + 2] << 16) | ((uint32_t)h[o + 3] << 24); };
    uint32_t off = 0;
    if (ok) {
        off = rd32(10);
        const int32_t hh = (int32_t)rd32(22);
        const uint32_t pal_n = rd32(46);
        ok = rd32(14) == 40 && (int32_t)rd32(18) == cols && (hh == rows || hh == -rows) && (h[28] | (h[29] << 8)) == 8 &&
             rd32(30) == 0 && (pal_n == 0 || pal_n == 256) && off >= sizeof(h);
        *bottom_up = hh > 0;
        // identity grey palette: entry i = (i, i, i, x)
        for (int i = 0; ok && i < 256; i++) ok = h[54 + 4 * i] == i && h[55 + 4 * i] == i && h[56 + 4 * i] == i;
    }
    if (ok) {
        const size_t n = (size_t)rows * cols;
        size_t got = 0;
        while (got < n) {
            const ssize_t r = ::pread(fd, dst + got, n - got, (off_t)off + (off_t)got);
            if (r <= 0) break;
            got += (size_t)r;
        }
        ok = got == n;
    }
    ::close(fd);
    return ok;
}

// RawParser.cpp:34-47: cv::imread(path, 0); returns !Image.empty()
int RawParser::GetImage(std::string EventID, std::string FrameName, Image8& Image) {
    std::vector<uint8_t> bytes;
    Image.release();
    if (!abubcodec::read_file(FramePath(EventID, FrameName), bytes)) return 0;
    abubcodec::decode(bytes.data(), bytes.size(), Image);
    return !Image.empty();
}

// RawParser.cpp:52-86: names in EventFolder containing camera_out_name, hidden files skipped
void RawParser::GetFileLists(const char* EventFolder, std::vector<std::string>& FileList, const char* camera_out_name) {
    std::error_code ec;
    fs::directory_iterator it(EventFolder, ec);
    if (ec) { StatusCode = 1; return; }
    for (const auto& e : it) {
        const std::string name = e.path().filename().string();
        if (name.empty() || name[0] == '.') continue;
        if (name.find(camera_out_name) != std::string::npos) FileList.push_back(name);
    }
}

// RawParser.cpp:89-137: every sub-directory of the run folder is an event
void RawParser::GetEventDirLists(std::vector<std::string>& EventList) {
    std::error_code ec;
    fs::directory_iterator it(RunFolder, ec);
    if (ec) { StatusCode = 1; return; }
    for (const auto& e : it) {
        const std::string name = e.path().filename().string();
        if (name.empty() || name[0] == '.') continue;
        std::error_code ec2;
        if (fs::is_directory(e.path(), ec2)) EventList.push_back(name);
    }
}

// regex_match of the whole path against "^.*cam<N>.*(png|bmp)": the path ends in png or bmp and "cam<N>" occurs before
// those three characters.  Written out: std::regex costs 5 us per path, 3 ms per event for each pass over its files.
bool RawParser::FramePathMatches(const std::string& p, int Camera) {
    const std::string needle = "cam" + std::to_string(Camera);
    const size_t n = p.size();
    if (n < 3 || (p.compare(n - 3, 3, "png") != 0 && p.compare(n - 3, 3, "bmp") != 0)) return false;
    if (p.find_first_of("\n\r") != std::string::npos) return false;       // '.' does not match a line break
    const size_t at = p.find(needle);
    return at != std::string::npos && at + needle.size() <= n - 3;
}

// RawParser.cpp:139-157: full-path regex "^.*cam<N>.*(png|bmp)", lexicographic sort of the file names
void RawParser::ParseAndSortFramesInFolder(std::string EventID, int Camera, std::vector<std::string>& Contents) {
    std::string dir = FramePath(EventID, "");
    std::error_code ec;
    if (fs::exists(dir, ec) && fs::is_directory(dir, ec)) {
        for (const auto& x : fs::directory_iterator(dir, ec)) {
            if (FramePathMatches(x.path().string(), Camera)) Contents.push_back(x.path().filename().string());
        }
        std::sort(Contents.begin(), Contents.end());
    }
}
