/* oracle/refshim/ref_stubs.cpp -- TEST INFRASTRUCTURE: link-time stand-ins for the two parts of the reference that are
 * not compiled into oracle/_ref/abub3hs_ref:
 *   * ZipParser (ParseFolder/ZipParser.cpp needs the minizip-ng submodule built with its CMake files): the class is
 *     declared by the reference's own header; every use ends the run the way an unreadable archive does;
 *   * the LBP operators (LBP/*.cpp): reachable only through MakeAvgSigmaImage(true), which main never calls
 *     (AutoBubStart3.cpp:306 passes false). */
#include <stdexcept>

#include "LBP/LBPUser.hpp"
#include "ParseFolder/ZipParser.hpp"

ZipParser::ZipParser(std::string RunFolder, std::string ImageFolder, std::string ImageFormat) : Parser(RunFolder, ImageFolder, ImageFormat) {
    throw std::runtime_error("oracle/_ref: zip archives are not supported by this build of the reference (minizip-ng is not compiled)");
}
ZipParser::~ZipParser() {}
ZipParser* ZipParser::clone() { return nullptr; }
int ZipParser::GetImage(std::string, std::string, cv::Mat&) { return -1; }
void ZipParser::GetEventDirLists(std::vector<std::string>&) {}
void ZipParser::GetFileLists(const char*, std::vector<std::string>&, const char*) {}
void ZipParser::ParseAndSortFramesInFolder(std::string, int, std::vector<std::string>&) {}
void ZipParser::GetRunFileInfo(std::vector<std::string>&) {}
void ZipParser::BuildFileList() {}
int ZipParser::SearchForFile(int, int, std::string) { return -1; }

cv::Mat lbpImage(cv::Mat&) { cv::shim_unsupported("lbpImage"); }
cv::Mat lbpImageSingleChan(cv::Mat&) { cv::shim_unsupported("lbpImageSingleChan"); }
cv::Mat lbpImageSingleChanExperimental(cv::Mat&) { cv::shim_unsupported("lbpImageSingleChanExperimental"); }
