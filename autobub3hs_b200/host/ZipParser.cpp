// ZipParser (reference: ParseFolder/ZipParser.cpp:38-327).
#include "ZipParser.hpp"

#include <unistd.h>

#include <string.h>
#include <zlib.h>

#include <algorithm>
#include <iostream>
#include <regex>

#include "ImageCodec.hpp"

namespace {
inline uint16_t r16(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }
inline uint32_t r32(const uint8_t* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
inline uint64_t r64(const uint8_t* p) { return (uint64_t)r32(p) | ((uint64_t)r32(p + 4) << 32); }

// positional read on the descriptor: no shared file position, so entries can be read from several threads at once
bool read_at(FILE* f, uint64_t off, void* dst, size_t n) {
    const int fd = fileno(f);
    uint8_t* p = static_cast<uint8_t*>(dst);
    while (n) {
        const ssize_t got = pread(fd, p, n, (off_t)off);
        if (got <= 0) return false;
        p += got; off += (uint64_t)got; n -= (size_t)got;
    }
    return true;
}
}  // namespace

ZipParser::ZipParser(std::string RunFolder_, std::string ImageFolder_, std::string ImageFormat_)
    : Parser(RunFolder_, ImageFolder_, ImageFormat_) {
    // remove the trailing separator, add ".zip" when there is no extension (ZipParser.cpp:41-45)
    while (RunFolder.size() > 1 && RunFolder.back() == '/') RunFolder.pop_back();
    const size_t slash = RunFolder.find_last_of('/');
    const std::string leaf = slash == std::string::npos ? RunFolder : RunFolder.substr(slash + 1);
    if (leaf.find('.') == std::string::npos || leaf.find('.') == 0) RunFolder += ".zip";
    fp = fopen(RunFolder.c_str(), "rb");
    if (!fp) {
        std::cerr << "Error initializing zip file; cannot continue" << std::endl;
        throw -10;
    }
}

ZipParser::~ZipParser() {
    if (fp) fclose(fp);
}

ZipParser* ZipParser::clone() {
    ZipParser* other = new ZipParser(RunFolder, ImageFolder, ImageFormat);
    if (!idx->FileContents.empty()) other->idx = idx;        // built: shared, read-only from here on
    return other;
}

// Walks the central directory (ZipParser.cpp:106-169): records every entry name, and for image entries
// the event ([1]) and frame name ([2]) of "^.*/(\d+)/.*/?(cam\d.*image\s*\d+.*(png|bmp))".
void ZipParser::BuildFileList() {
    fseeko(fp, 0, SEEK_END);
    const uint64_t fsize = (uint64_t)ftello(fp);
    const uint64_t tail = fsize < 66000 ? fsize : 66000;
    std::vector<uint8_t> t(tail);
    if (tail < 22 || !read_at(fp, fsize - tail, t.data(), tail)) throw -10;
    int64_t e = -1;
    for (int64_t i = (int64_t)tail - 22; i >= 0; i--)
        if (r32(&t[i]) == 0x06054b50u) { e = i; break; }
    if (e < 0) { std::cout << "Error when creating file list; cannot continue." << std::endl; throw -10; }
    uint64_t n_entries = r16(&t[e + 10]), cd_size = r32(&t[e + 12]), cd_off = r32(&t[e + 16]);
    if (e >= 20 && r32(&t[e - 20]) == 0x07064b50u) {          // ZIP64 locator -> ZIP64 end of central directory
        const uint64_t e64 = r64(&t[e - 20 + 8]);
        uint8_t rec[56];
        if (!read_at(fp, e64, rec, 56) || r32(rec) != 0x06064b50u) throw -10;
        n_entries = r64(rec + 32); cd_size = r64(rec + 40); cd_off = r64(rec + 48);
    }
    // sizes come from the archive: they must lie inside the file before anything is allocated from them
    if (cd_off > fsize || cd_size > fsize - cd_off || n_entries > cd_size / 46 + 1) {
        std::cout << "Error when creating file list; cannot continue." << std::endl;
        throw -10;
    }
    idx->archive_size = fsize;
    std::vector<uint8_t> cd(cd_size);
    if (cd_size && !read_at(fp, cd_off, cd.data(), cd_size)) throw -10;
    const std::regex re("^.*/(\\d+)/.*/?(cam\\d.*image\\s*\\d+.*(png|bmp))");
    std::smatch m;
    size_t p = 0;
    for (uint64_t i = 0; i < n_entries; i++) {
        if (p + 46 > cd.size() || r32(&cd[p]) != 0x02014b50u) {
            std::cout << "Error when creating file list; cannot continue." << std::endl;
            throw -10;
        }
        ZipEntryLoc loc;
        loc.method = r16(&cd[p + 10]);
        loc.comp_size = r32(&cd[p + 20]);
        loc.uncomp_size = r32(&cd[p + 24]);
        const size_t nlen = r16(&cd[p + 28]), xlen = r16(&cd[p + 30]), clen = r16(&cd[p + 32]);
        loc.local_off = r32(&cd[p + 42]);
        if (p + 46 + nlen + xlen + clen > cd.size()) throw -10;
        const std::string name((const char*)&cd[p + 46], nlen);
        // ZIP64 extended information (header id 1): present fields follow in the fixed order
        size_t x = p + 46 + nlen;
        const size_t xend = x + xlen;
        while (x + 4 <= xend) {
            const uint16_t id = r16(&cd[x]), sz = r16(&cd[x + 2]);
            if (id == 1) {
                size_t q = x + 4;
                if (loc.uncomp_size == 0xFFFFFFFFu && q + 8 <= xend) { loc.uncomp_size = r64(&cd[q]); q += 8; }
                if (loc.comp_size == 0xFFFFFFFFu && q + 8 <= xend) { loc.comp_size = r64(&cd[q]); q += 8; }
                if (loc.local_off == 0xFFFFFFFFu && q + 8 <= xend) { loc.local_off = r64(&cd[q]); q += 8; }
            }
            x += 4 + sz;
        }
        idx->FileContents.push_back(name);
        if (std::regex_match(name, m, re)) {
            idx->ImageNames[m[1].str()][m[2].str()] = name;
            idx->ImageLocs[m[1].str()][m[2].str()] = loc;
        }
        p += 46 + nlen + xlen + clen;
    }
}

bool ZipParser::ReadEntry(const ZipEntryLoc& loc, std::vector<uint8_t>& out) {
    uint8_t lh[30];
    if (!read_at(fp, loc.local_off, lh, 30) || r32(lh) != 0x04034b50u) return false;
    const uint64_t data = loc.local_off + 30 + r16(lh + 26) + r16(lh + 28);
    // a corrupt central directory must end as "frame does not decode" (-9 / -6 for the camera), not as bad_alloc inside
    // a parallel region: the compressed bytes must lie inside the archive, and one frame never inflates past kMaxEntry
    constexpr uint64_t kMaxEntry = (uint64_t)256 << 20;
    if (data > idx->archive_size || loc.comp_size > idx->archive_size - data || loc.uncomp_size > kMaxEntry ||
        (loc.method == 0 && loc.comp_size > kMaxEntry))
        return false;
    std::vector<uint8_t> comp(loc.comp_size);
    if (loc.comp_size && !read_at(fp, data, comp.data(), loc.comp_size)) return false;
    if (loc.method == 0) { out.swap(comp); return true; }
    if (loc.method != 8) return false;
    out.resize(loc.uncomp_size);
    z_stream zs;
    memset(&zs, 0, sizeof(zs));
    if (inflateInit2(&zs, -MAX_WBITS) != Z_OK) return false;
    zs.next_in = comp.data(); zs.avail_in = (uInt)comp.size();
    zs.next_out = out.data(); zs.avail_out = (uInt)out.size();
    const int rc = inflate(&zs, Z_FINISH);
    inflateEnd(&zs);
    return rc == Z_STREAM_END && zs.total_out == loc.uncomp_size;
}

int ZipParser::GetImageBytes(std::string EventID, std::string FrameName, std::vector<uint8_t>& Bytes) {
    if (idx->FileContents.size() == 0) BuildFileList();
    auto ev = idx->ImageLocs.find(EventID);
    if (ev == idx->ImageLocs.end()) return -1;
    auto fr = ev->second.find(FrameName);
    if (fr == ev->second.end()) return -1;
    return ReadEntry(fr->second, Bytes) ? 1 : -1;
}

// ZipParser.cpp:186-239: -1 when the entry is not listed or does not decode
int ZipParser::GetImage(std::string EventID, std::string FrameName, Image8& Image) {
    std::vector<uint8_t> bytes;
    Image.release();
    if (GetImageBytes(EventID, FrameName, bytes) == -1) return -1;
    if (!abubcodec::decode(bytes.data(), bytes.size(), Image)) {
        std::cerr << "Failed to decode image " << FrameName << std::endl;
        return -1;
    }
    return 1;
}

void ZipParser::GetFileLists(const char*, std::vector<std::string>&, const char*) {}   // empty in the reference too (:242-244)

// ZipParser.cpp:248-283: directory entries "^.*/(\d+)/$"
void ZipParser::GetEventDirLists(std::vector<std::string>& EventList) {
    if (idx->FileContents.size() == 0) BuildFileList();
    const std::regex expression("^.*/(\\d+)/$");
    std::smatch what;
    for (const std::string& name : idx->FileContents)
        if (std::regex_match(name, what, expression)) EventList.push_back(what[1].str());
}

// ZipParser.cpp:286-315
void ZipParser::ParseAndSortFramesInFolder(std::string EventID, int camera, std::vector<std::string>& Contents) {
    if (idx->FileContents.size() == 0) BuildFileList();
    const std::regex re("cam" + std::to_string(camera) + ".*image.*(png|bmp)");
    auto ev = idx->ImageNames.find(EventID);
    if (ev != idx->ImageNames.end())
        for (const auto& entry : ev->second)
            if (std::regex_match(entry.first, re)) Contents.push_back(entry.first);
    std::sort(Contents.begin(), Contents.end());
}
