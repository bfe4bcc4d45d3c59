// The label-map handoff of the reference (src/fcn/training.cpp:169-172): rama_torch_batched
// (include/rama_wrapper.cuh:15-18, src/rama_wrapper.cu:52-127) returns node labels int32 [B, N=H*W], CUDA,
// contiguous; training.cpp views them as [B,H,W] and compute_rewards_batched widens them to int64
// (src/compute_rewards.cu:97).  Here the int32 plane is consumed as delivered: no cast, no copy.
#pragma once
#include <torch/torch.h>
#include <vector>

// Checks the producer contract and returns the [B,H,W] view compute_rewards_batched takes.
torch::Tensor labels_from_rama(const torch::Tensor& node_labels_bn_i32, int64_t H, int64_t W);

// Producer side: what rama_torch_batched does after each solve (src/rama_wrapper.cu:102-124), batched.  `node_mappings`
// are the solver's per-sample results (int32 CUDA, N = H*W entries each, same device); returns the [B,H,W] int32 plane,
// written by one launch per 256 samples instead of one copy per sample.
torch::Tensor labels_from_node_mappings(const std::vector<torch::Tensor>& node_mappings, int64_t H, int64_t W);
