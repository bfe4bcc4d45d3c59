// Minimal PNG encoder/decoder on zlib for the slice files: 8-bit, non-interlaced.  Stands in for cv::imwrite /
// cv::imread (src/image_writer.cpp:7, src/reassemble.cpp:58), which this build does not have.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace icr_png {
// rgba: h*w*4 bytes.  Sub filter on every row and deflate at `level`: what OpenCV's writer does when
// IMWRITE_PNG_COMPRESSION is given (the reference passes COMPRESSION_LEVEL = 4, include/configuration.h:11).
std::vector<uint8_t> encode_rgba(const uint8_t* rgba, int w, int h, int level);
bool write_file(const std::string& path, const std::vector<uint8_t>& bytes);
// Decodes grey, grey+alpha, RGB and RGBA files of bit depth 8 to RGBA (ensure_rgba, src/reassemble.cpp:14-28).
// Returns false with `err` set for anything else.
bool decode_rgba(const std::vector<uint8_t>& file, std::vector<uint8_t>& rgba, int& w, int& h, std::string& err);
bool read_file(const std::string& path, std::vector<uint8_t>& bytes);
}  // namespace icr_png
