// Rule W (src/compute_rewards.cu:183-192) for one label map by one warp, fixed summation order.  Called by whichever
// warp stored the last segment size of the map (narrow.cu, estimate.cu, big_segment.cu count finished segments in
// Tables::done).  VOLATILE: segment sizes may have been written by other CTAs of the running kernel, so they are read past L1.
#pragma once
#include "icr_internal.h"

namespace icr {

// one reward to the caller's buffer and, when the call gathers, to the same slot of every rank's gathered vector
__device__ __forceinline__ void store_reward(const Tables& t, double* __restrict__ rewards, int v, double r) {
    if (rewards) rewards[v] = r;
    if (t.peer_rewards) {
        for (int g = 0; g < t.peer_count; ++g) t.peer_rewards[g][t.peer_offset + v] = r;
        __threadfence_system();         // ordered before the flag the last CTA of the batch's last kernel raises (big_segment.cu)
    }
}

// Called by the thread that has just stored the size of one segment of label map v: counts it in, ordered after the
// store (release at device scope: no L1 invalidation, unlike a full fence), and tells whether it was the map's last.
// The caller of the last one fences (acquire side) before it reads the other sizes past L1 in reward_of_map.
__device__ __forceinline__ int segment_done(const Tables& t, int v, int K) {
    int old;
    asm volatile("atom.add.release.gpu.global.s32 %0, [%1], 1;" : "=r"(old) : "l"(t.done + v) : "memory");
    return old == K - 1;
}

template <bool VOLATILE>
__device__ __forceinline__ void reward_of_map(const Tables& t, const EstParams& p, const double* __restrict__ image_sizes, int S,
                                              double* __restrict__ rewards, int v, int lane) {
    if (t.status[v] != ST_OK) { if (lane == 0) store_reward(t, rewards, v, __longlong_as_double(0x7ff8000000000000ll)); return; }
    const int K = t.K[v];
    const double* sz = t.seg_sizes + (int64_t)v * t.Kmax;
    const int32_t* cnt = t.counts + (int64_t)v * t.Kmax;
    double sum = 0.0;
    int valid = 0;
    for (int k = lane; k < K; k += 32) { sum += VOLATILE ? __ldcg(sz + k) : sz[k]; valid += cnt[k] >= p.min_pixels ? 1 : 0; }
    for (int o = 16; o; o >>= 1) { sum += __shfl_xor_sync(0xffffffffu, sum, o); valid += __shfl_xor_sync(0xffffffffu, valid, o); }
    if (lane == 0) {
        if (K <= 0) { store_reward(t, rewards, v, 0.0); return; }   // :136-139
        const double s_img = image_sizes[v / S];
        const double P = valid == 1 ? 1.0 : 0.0;
        store_reward(t, rewards, v, (s_img - sum) / s_img - p.lambda * P);
    }
}

}  // namespace icr
