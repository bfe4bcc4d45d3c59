// Drop-in header for the reference's include/png_size_estimator.cuh:61-71 (the pointer-level estimator).
// The reference's PngEstimatorWorkspace / get_png_ws (:12-58) have no counterpart to include: scratch lives
// behind the C ABI (icr_workspace, and a per-(thread, device) filter-id buffer for this entry).
#pragma once
#include <torch/torch.h>
#include <cstdint>

// Writes one size estimate into out_dev[0] (device pointer) on the current PyTorch CUDA stream.
void estimate_png_size_masked_segment_to_output(
    const uint8_t* img_hwc_u8, int full_W, int full_H, int channels,
    const int64_t* labels_compact_hw, const int32_t* counts_k, int64_t seg_id_k,
    int x0, int y0, int w, int h,
    int32_t min_pixels, int L_min, float beta, float b_match_token, float gamma, double overhead_base,
    bool adaptive_filter, double* out_dev);
