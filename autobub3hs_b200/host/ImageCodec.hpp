// Decoders for the frame formats the reference reads through cv::imread(path, 0) /
// cv::imdecode(buf, 0) (RawParser.cpp:39, ZipParser.cpp:222): BMP and PNG -> 8-bit gray.
#ifndef ABUB_IMAGECODEC_HPP
#define ABUB_IMAGECODEC_HPP

#include <stddef.h>

#include <string>

#include "Image8.hpp"

namespace abubcodec {

// Decodes an in-memory BMP or PNG to 8-bit gray.  `dst` may be NULL-backed (Image8) or a
// caller buffer of rows*cols bytes (decode_into: used to decode straight into the pinned
// staging buffer of an event slot).  Returns false (and leaves `img` empty) on any error.
bool decode(const uint8_t* buf, size_t n, Image8& img);
// Peeks the dimensions without decoding the pixels.
bool peek_size(const uint8_t* buf, size_t n, int* rows, int* cols);
// Decodes into caller memory; fails if the image is not rows x cols.
bool decode_into(const uint8_t* buf, size_t n, uint8_t* dst, int rows, int cols);

bool read_file(const std::string& path, std::vector<uint8_t>& out);
// For the on-device decoder (abub_event_decode_png): when buf is an 8-bit greyscale, non-interlaced PNG of exactly
// rows x cols, copies its zlib stream (the concatenated IDAT payloads) to dst (capacity cap) and returns its length;
// 0 when the file is anything else or does not fit.
size_t png_gray8_stream(const uint8_t* buf, size_t n, int rows, int cols, uint8_t* dst, size_t cap);

// Minimal writers (8-bit gray): used by the tests and the synthetic-run generator.
bool write_bmp_gray(const std::string& path, const uint8_t* px, int rows, int cols);

}  // namespace abubcodec

#endif
