// Sidecar of a slice directory ("metadata.bin"), source-compatible with the reference's include/metadata.h:6-32 for its
// callers: the record type and the two functions keep their names, member names and signatures.  The on-disk layout is
// written out byte by byte in slices_host.cpp instead of through packed structs, so it does not depend on the compiler:
//
//   offset 0   u32 LE  magic 0x534C4943 ("SLIC")          include/metadata.h:14
//          4   u32 LE  number of records                   :15
//          8   u32 LE  width of the source image           :16
//         12   u32 LE  height of the source image          :17
//   then per record, 22 bytes + the name:
//          0   i32 LE  label                               :21
//          4   i32 LE  x of the slice in the source image  :22
//          8   i32 LE  y                                   :23
//         12   i32 LE  width of the slice                  :24
//         16   i32 LE  height of the slice                 :25
//         20   u16 LE  length of the file name             :26
//         22   the file name, not terminated               src/metadata.cpp:32
#ifndef ICR_HOST_METADATA_H
#define ICR_HOST_METADATA_H

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace icr_sidecar {
constexpr std::uint32_t kMagic = 0x534C4943u;
constexpr std::size_t kHeaderBytes = 16;
constexpr std::size_t kRecordBytes = 22;      // without the name
}  // namespace icr_sidecar

// one slice: which label it shows, the file that holds it, and where its bbox sits in the source image
struct SliceMetadata {
    int label;
    std::string filename;
    int x;
    int y;
    int width;
    int height;
};

// Writes `path` (truncating); like the reference it does not report I/O errors.
void write_metadata_binary(const std::vector<SliceMetadata>& metadata, const std::string& path,
                           uint32_t image_width, uint32_t image_height);

// Throws std::runtime_error("Cannot open metadata file") or ("Invalid metadata file (magic mismatch)"), the messages of
// src/metadata.cpp:39-49.
std::vector<SliceMetadata> read_metadata_binary(const std::string& path, uint32_t& image_width, uint32_t& image_height);

#endif
