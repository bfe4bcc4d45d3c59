// Drop-in header for the reference's include/segment_stats.cuh:6-10.
#pragma once
#include <torch/torch.h>

// bboxes: int32 [K,4] = (x0,y0,x1,y1) inclusive, counts: int32 [K]; K is counts.numel()
void compute_counts_bboxes_from_compact_labels_cuda(
    const torch::Tensor& labels_compact_hw_i64,   // [H,W] int64, values 0..K-1
    torch::Tensor& counts_k_i32,                  // [K] int32
    torch::Tensor& bboxes_k4_i32);                // [K,4] int32

// Extension: Rule L + Rule S for a whole batch in one launch.  labels [V,H,W] int32/int64 CUDA with arbitrary
// values -> {K [V] i32, counts [V,Kmax] i32, bboxes [V,Kmax,4] i32, keys [V,Kmax] i64 (sorted unique labels)}.
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor> segment_stats_batched(
    const torch::Tensor& labels_vhw, int64_t Kmax = 0);
