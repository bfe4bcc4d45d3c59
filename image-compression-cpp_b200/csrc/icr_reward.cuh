// Rule W (src/compute_rewards.cu:183-192) for one label map by one warp, fixed summation order.  Shared by the reward
// kernel and by the big-segment kernel's last CTA of a map.  VOLATILE: segment sizes may have been written by other CTAs
// of the running kernel, so they are read past L1.
#pragma once
#include "icr_internal.h"

namespace icr {

// one reward to the caller's buffer and, when the call gathers, to the same slot of every rank's gathered vector
__device__ __forceinline__ void store_reward(const Tables& t, double* __restrict__ rewards, int v, double r) {
    if (rewards) rewards[v] = r;
    if (t.peer_rewards)
        for (int g = 0; g < t.peer_count; ++g) t.peer_rewards[g][t.peer_offset + v] = r;
}

// Called by every CTA of a reward-writing kernel as its last action.  Each CTA makes its peer stores visible system-wide
// and counts itself in; the CTA that finds all others done tells every rank that this rank's rewards of step peer_step
// have landed.  Release/acquire at system scope pairs with the wait in the EMA kernel.
__device__ __forceinline__ void signal_peers_when_grid_done(const Tables& t, int ctas) {
    if (!t.peer_flags) return;
    __syncthreads();
    if (threadIdx.x != 0) return;
    __threadfence_system();
    if (atomicAdd(&t.sched[2], 1) != ctas - 1) return;
    t.sched[2] = 0;
    __threadfence_system();
    for (int g = 0; g < t.peer_count; ++g)
        asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(t.peer_flags[g] + t.peer_rank), "l"(t.peer_step) : "memory");
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
