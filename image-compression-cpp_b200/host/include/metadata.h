// Drop-in header for the reference's include/metadata.h:6-32 (same types, same two functions, same bytes on disk:
// 16-byte "SLIC" header, then per record a packed 22-byte fixed part and the filename bytes).
#pragma once
#include <stdint.h>
#include <string>
#include <vector>

struct SliceMetadata {
    int label;
    std::string filename;
    int x, y, width, height;
};

#pragma pack(push, 1)
struct SliceRecordHeader {
    uint32_t magic;  // "SLIC" = 0x534C4943
    uint32_t count;
    uint32_t original_width;
    uint32_t original_height;
};

struct SliceRecordFixed {
    int32_t label;
    int32_t x;
    int32_t y;
    int32_t width;
    int32_t height;
    uint16_t filename_len;
};
#pragma pack(pop)

void write_metadata_binary(const std::vector<SliceMetadata>& metadata, const std::string& path, uint32_t image_width, uint32_t image_height);
// throws std::runtime_error("Cannot open metadata file") / ("Invalid metadata file (magic mismatch)") as src/metadata.cpp:39-49
std::vector<SliceMetadata> read_metadata_binary(const std::string& path, uint32_t& image_width, uint32_t& image_height);
