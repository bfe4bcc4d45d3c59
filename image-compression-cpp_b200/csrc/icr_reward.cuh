// Rule W (src/compute_rewards.cu:183-192) for one label map by one warp, fixed summation order.  Shared by the reward
// kernel and by the big-segment kernel's last CTA of a map.  VOLATILE: segment sizes may have been written by other CTAs
// of the running kernel, so they are read past L1.
#pragma once
#include "icr_internal.h"

namespace icr {

template <bool VOLATILE>
__device__ __forceinline__ void reward_of_map(const Tables& t, const EstParams& p, const double* __restrict__ image_sizes, int S,
                                              double* __restrict__ rewards, int v, int lane) {
    if (t.status[v] != ST_OK) { if (lane == 0) rewards[v] = __longlong_as_double(0x7ff8000000000000ll); return; }
    const int K = t.K[v];
    const double* sz = t.seg_sizes + (int64_t)v * t.Kmax;
    const int32_t* cnt = t.counts + (int64_t)v * t.Kmax;
    double sum = 0.0;
    int valid = 0;
    for (int k = lane; k < K; k += 32) { sum += VOLATILE ? __ldcg(sz + k) : sz[k]; valid += cnt[k] >= p.min_pixels ? 1 : 0; }
    for (int o = 16; o; o >>= 1) { sum += __shfl_xor_sync(0xffffffffu, sum, o); valid += __shfl_xor_sync(0xffffffffu, valid, o); }
    if (lane == 0) {
        if (K <= 0) { rewards[v] = 0.0; return; }                   // :136-139
        const double s_img = image_sizes[v / S];
        const double P = valid == 1 ? 1.0 : 0.0;
        rewards[v] = (s_img - sum) / s_img - p.lambda * P;
    }
}

}  // namespace icr
