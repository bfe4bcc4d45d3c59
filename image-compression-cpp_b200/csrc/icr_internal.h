// Internal declarations shared by the torch-free translation units of libicr_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/icr_b200.h"

namespace icr {

// status bits per label map
enum : int { ST_OK = 0, ST_LABEL_RANGE = 1, ST_K_OVERFLOW = 2 };

constexpr int KS_SMEM = 2048;          // per-CTA shared-memory stats table entries in prep
constexpr int64_t LCAP_MAX = 1 << 19;  // label bitmap capacity (bits) that fits shared memory

struct Tables {            // device pointers into the workspace, V label maps
    int32_t* K;            // [V]
    int32_t* status;       // [V]
    int32_t* counts;       // [V,Kmax]
    int4* bboxes;          // [V,Kmax] (x0,y0,x1,y1) inclusive
    long long* keys;       // [V,Kmax] original label value of compact id k (ascending: Rule L)
    double* seg_sizes;     // [V,Kmax]
    int32_t* err;          // [2] {any-error flag, 0x7fffffff - first bad map}
    int32_t* seg_prefix;   // [V+1] exclusive scan of the segment counts of the maps with status OK
    int32_t* sched;        // [4] {prep CTAs done, estimate work cursor, -, -}
    uint8_t* big_filter;   // [V*BIG_CTAS, H] filter ids of the segment a big-kernel CTA is working on
    uint2* runs;           // [V, run_cap] label runs of a map found by prep's first sweep {slot | (len-1)<<20, y<<16 | x}
    int run_cap;           // 0 = no run list (very wide maps)
    int2* map_max;         // [V] {largest bbox area, widest bbox} of a map (INT_MAX when not known): lets the big-segment kernel skip maps
    int32_t* big_done;     // [V] big-segment CTAs of a map that have finished (self-resetting)
    unsigned long long* slice_map_off;   // [V] byte offset of a map's first slice in the packed slice buffer (slices.cu)
    const double* clog2;   // [CLOG2_N] c*log2(c)
    // multi-GPU gather fused into the reward writer (icr_compute_rewards_gather): the reward of map v also goes to
    // peer_rewards[g][peer_offset + v] for every rank g -- plain stores into peer memory over NVLink.  NULL otherwise.
    double* const* peer_rewards;
    int peer_count, peer_offset;
    // optional signalling in place of a barrier: when the last CTA of the reward-writing kernel is done, flag
    // peer_flags[g][peer_rank] = peer_step is stored into every rank g (icr_reward.cuh: signal_peers_when_grid_done);
    // the consumer (icr_ema_update_gathered) waits for the flags of all ranks.
    unsigned long long* const* peer_flags;
    unsigned long long peer_step;
    int peer_rank;
    int V, H, W, Kmax;
};

struct Inputs {            // the caller's device buffers of the current call
    const float* images;   // [B,3,H,W] planar f32
    const void* labels;    // [B*S,H,W]
    int label_dtype;       // ICR_LABELS_I32 / ICR_LABELS_I64
    int S;                 // label maps per image
};

constexpr int BIG_CTAS = 2;       // big-segment CTAs per label map
constexpr int CLOG2_N = 4096;
constexpr int EST_MAX_W = 224;    // widest bbox of the warp path: a one-row group (two row slots) must fit the warp's tile
constexpr int EST_MAX_AREA = 1 << 20;

// Segments the warp path leaves to the big-segment kernel: Rule M can fire (chunk = ceil(N/65536) >= L_min,
// src/png_size_estimator.cu:416-451), or the bbox is wider than 32 register strips, or very large.
__host__ __device__ inline bool for_big_kernel(int w, int h, int L_min) {
    if (w <= 0 || h <= 0) return false;
    long long N = 4ll * w * h;
    return (N + 65535) / 65536 >= (long long)L_min || w > EST_MAX_W || (long long)w * h > EST_MAX_AREA;
}

// can a map whose largest bbox has `area` pixels and whose widest bbox has `maxw` columns hold a big-kernel segment?
__host__ __device__ inline bool map_may_need_big_kernel(int area, int maxw, int L_min) {
    return (4ll * area + 65535) / 65536 >= (long long)L_min || maxw > EST_MAX_W || (long long)area > EST_MAX_AREA;
}

struct EstParams {         // icr_params widened the way the reference does (:588)
    int32_t min_pixels, L_min, adaptive;
    double beta, token, gamma, overhead, lambda;
};

// label capacity (bits of the circular presence bitmap) for an HxW map: a power of two >= max(H*W, 65536), at most
// LCAP_MAX (what shared memory holds).  Maps above LCAP_MAX pixels therefore work as long as their labels spread over
// less than LCAP_MAX values; beyond that the map is reported (ST_LABEL_RANGE) and the host layer ranks the labels first.
inline int64_t label_capacity(int H, int W) {
    int64_t hw = (int64_t)H * W;
    int64_t cap = 65536;
    while (cap < hw && cap < LCAP_MAX) cap <<= 1;
    return cap;
}

// run-list capacity per map: an average run of 8 pixels or more fits; beyond it prep falls back to a second pixel sweep
inline int run_capacity(int H, int W) {
    if (W > 4096 || H > 65535) return 0;          // record packing: 12 bits of length, 16 bits of x and y
    int64_t c = (int64_t)H * W / 8;
    return (int)(c < 1024 ? 1024 : c);
}

// launchers (each returns the cudaError_t of its launches)
cudaError_t launch_prep(const void* labels, int label_dtype, const Tables& t, int32_t* compact_out, cudaStream_t st);
cudaError_t launch_estimate(const Tables& t, const Inputs& in, const EstParams& p, cudaStream_t st);
// image_sizes/rewards non-null: the last big-segment CTA of each map also computes the map's reward (Rule W), so the
// reward kernel is not launched; *launched tells whether the kernel ran at all
cudaError_t launch_big(const Tables& t, const Inputs& in, const EstParams& p, const double* image_sizes, double* rewards,
                       bool* launched, cudaStream_t st);
cudaError_t launch_reward(const Tables& t, const EstParams& p, const double* image_sizes, int S,
                          double* rewards, cudaStream_t st);
cudaError_t launch_estimate_debug(const Tables& t, const Inputs& in, const EstParams& p, int v, int k, uint32_t* row_costs,
                                  uint8_t* filter_id, uint32_t* hist, unsigned long long* match,
                                  double* size_out, cudaStream_t st);
cudaError_t launch_big_debug(const Tables& t, const Inputs& in, const EstParams& p, int v, int k, uint32_t* row_costs,
                             uint8_t* filter_id, uint32_t* hist, unsigned long long* match,
                             double* size_out, cudaStream_t st);
cudaError_t launch_slice(const Tables& t, const Inputs& in, int v, int k, uint8_t* out, size_t cap, cudaStream_t st);
cudaError_t launch_slice_plan(const Tables& t, unsigned long long* offsets, long long capacity, unsigned long long* summary,
                              cudaStream_t st);
cudaError_t launch_slice_emit(const Tables& t, const Inputs& in, const unsigned long long* offsets, uint8_t* out, size_t cap,
                              int swap, cudaStream_t st);
cudaError_t launch_reassemble(const Tables& t, int v0, int nmaps, const unsigned long long* offsets, const uint8_t* slices,
                              uint8_t* canvases, cudaStream_t st);
cudaError_t launch_paste(uint8_t* canvas, int W, int H, const uint8_t* slice, int x, int y, int w, int h, cudaStream_t st);
cudaError_t launch_compact_stats(const int64_t* labels, int H, int W, int K, int32_t* counts,
                                 int32_t* bboxes, cudaStream_t st);
cudaError_t launch_single_segment(const uint8_t* img, int W, int H, int channels, const int64_t* labels,
                                  int64_t k, int x0, int y0, int w, int h, const EstParams& p,
                                  double* out, uint32_t* row_costs, uint8_t* filter_id, uint32_t* hist,
                                  unsigned long long* match, uint8_t* filter_scratch, cudaStream_t st);
cudaError_t launch_ema(const double* rewards, int n, const double* sum_count, double momentum,
                       double* state, double* adv, const unsigned long long* flags, int ranks, unsigned long long step,
                       cudaStream_t st);
cudaError_t launch_sum_count(const double* rewards, int n, double* sum_count, cudaStream_t st);
cudaError_t run_selftest(int* mismatches);

void count_launch(int n = 1);

}  // namespace icr
