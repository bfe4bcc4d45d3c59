// The decoder of src/reassemble.cpp:30-120 on the device: read metadata.bin, decode each slice file, paste it onto a
// transparent canvas where its alpha > 0 (in record order, clipped to the canvas), write the canvas as PNG.
// Differences, on purpose: slice files are looked up next to metadata.bin first (the reference reads m.filename relative
// to the working directory, src/reassemble.cpp:58, which only works when that is the slice directory), then as given.
#pragma once
#include <torch/torch.h>

#include <filesystem>

// canvas u8 [H,W,4] on `device` (RGBA byte order of the files); throws std::runtime_error when metadata.bin is unreadable
torch::Tensor reassemble_to_tensor(const std::filesystem::path& image_path, int device = 0);
bool reassemble(std::filesystem::path image_path, std::filesystem::path out_filename);
