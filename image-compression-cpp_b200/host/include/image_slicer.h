// Drop-in header for the reference's include/image_slicer.h:6-8, with device tensors in place of cv::Mat (this build has
// no OpenCV): the image arrives the way the rest of the path holds it -- f32 [3,H,W] planes in 0..255 on the GPU
// (src/compress.cpp:112-128 has that tensor before it converts to a cv::Mat) -- and the label map stays on the GPU instead
// of the .to(kCPU) of src/compress.cpp:142.  All slices of the image are cut in one pass on the device
// (icr_slices_prepare / _plan / _emit), copied to the host once, and encoded to PNG by a pool of host threads.
#pragma once
#include <torch/torch.h>

#include <filesystem>
#include <vector>

#include "metadata.h"

// What the device pass produces for a batch: every slice of every label map back to back.
struct SliceBatch {
    torch::Tensor bytes;     // u8 [total] on the device; slice f is bytes[offsets[f] : offsets[f+1]] viewed as [h,w,4]
    torch::Tensor offsets;   // int64 [segments+1] on the device
    torch::Tensor K;         // int32 [V] segments per label map
    torch::Tensor bboxes;    // int32 [V,Kmax,4] (x0,y0,x1,y1) inclusive
    torch::Tensor keys;      // int64 [V,Kmax] label value of compact id k (ascending)
};

// images f32 [B,3,H,W] CUDA, labels int32/int64 [B,H,W] or [B,S,H,W] CUDA.  swap_rb: write (c2,c1,c0,255) -- BGRA from
// RGB planes, the cv::Mat layout of src/compress.cpp:145 -- instead of (c0,c1,c2,255).
SliceBatch slice_images_batched(const torch::Tensor& images_bchw_f32, const torch::Tensor& labels, bool swap_rb = false);

// slice_image of include/image_slicer.h:6 / src/image_slicer.cpp:48-74 for ONE label VALUE: the bbox crop of the image with
// the pixels outside the label zeroed, u8 [h,w,4] = (c0,c1,c2,255) inside / (0,0,0,0) outside, on the input's device.
// input: f32 [3,H,W] or [1,3,H,W] CUDA; mask: integer [H,W] CUDA.  out_box = {x, y, width, height} (cv::Rect order).
// A label that does not occur returns an empty tensor and leaves out_box alone, as the reference returns cv::Mat().
struct SliceBox { int x = 0, y = 0, width = 0, height = 0; };
torch::Tensor slice_image(const torch::Tensor& input_chw_f32, const torch::Tensor& mask, int64_t label, SliceBox& out_box);

// write_slices of src/image_slicer.cpp:81-130: output_path/file_directory_name/slice_<label>.png for every label present
// plus metadata.bin; records are in ascending label order.  input: f32 [3,H,W] or [1,3,H,W] CUDA, channel planes in the
// order the PNG should store them (R,G,B); mask: integer [H,W] CUDA.  Returns false when a file could not be written.
bool write_slices(const torch::Tensor& input_chw_f32, const torch::Tensor& mask, const std::filesystem::path& output_path,
                  const std::filesystem::path& file_directory_name);
