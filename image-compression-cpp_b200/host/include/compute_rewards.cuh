// Drop-in header for the reference's include/compute_rewards.cuh:5-17: same name, same parameter order and types, same
// default values.  The implementation (host/reward_host.cpp) is host C++ over libtorch tensors that reaches the GPU only
// through the C ABI of libicr_b200.so (include/icr_b200.h).
//
//   images       [B,3,H,W] float32 in 0..255, any device (moved to the labels' device when needed)
//   labels       [B,H,W] integer, CUDA; the int32 maps rama_torch_batched hands over are read as they are
//   image_sizes  [B] numeric, CUDA: the size the whole image compresses to
//   the eight estimator/reward knobs and their defaults are the reference's (:9-16)
// Returns a fresh [B] float64 CUDA tensor.
#ifndef ICR_HOST_COMPUTE_REWARDS_CUH
#define ICR_HOST_COMPUTE_REWARDS_CUH

#include <torch/torch.h>

torch::Tensor compute_rewards_batched(const torch::Tensor& images, const torch::Tensor& labels, const torch::Tensor& image_sizes,
                                      int32_t min_pixels_per_segment = 1, int L_min = 4, float beta = 0.012167f,
                                      float b_match_token = 18.0f, float gamma = 0.1f, double overhead_base = 9.308622,
                                      bool adaptive_filter = true, double lambda = 0.5);

// Extension (BASELINE config 4): S label maps per image, labels [B,S,H,W]; each image is read once.  Equals
// compute_rewards_batched(images.repeat_interleave(S,0), labels.view({B*S,H,W}), sizes.repeat_interleave(S)).view({B,S}).
torch::Tensor compute_rewards_multi(const torch::Tensor& images, const torch::Tensor& labels_bshw, const torch::Tensor& image_sizes,
                                    int32_t min_pixels_per_segment = 1, int L_min = 4, float beta = 0.012167f,
                                    float b_match_token = 18.0f, float gamma = 0.1f, double overhead_base = 9.308622,
                                    bool adaptive_filter = true, double lambda = 0.5);

#endif
