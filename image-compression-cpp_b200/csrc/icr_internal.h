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
    uint2* pixlab;         // [V,H,W] {RGBA bytes, compact id}
    int32_t* K;            // [V]
    int32_t* status;       // [V]
    int32_t* counts;       // [V,Kmax]
    int4* bboxes;          // [V,Kmax] (x0,y0,x1,y1) inclusive
    double* seg_sizes;     // [V,Kmax]
    int32_t* err;          // [2] {any-error flag, first bad map + 1}
    uint8_t* big_filter;   // [V*BIG_CTAS, H] filter ids of the segment a big-kernel CTA is working on
    const double* clog2;   // [CLOG2_N] c*log2(c)
    int V, H, W, Kmax;
};

constexpr int BIG_CTAS = 2;       // big-segment CTAs per label map
constexpr int CLOG2_N = 4096;

struct EstParams {         // icr_params widened the way the reference does (:588)
    int32_t min_pixels, L_min, adaptive;
    double beta, token, gamma, overhead, lambda;
};

// label capacity (bits of the presence bitmap) for an HxW map
inline int64_t label_capacity(int H, int W) {
    int64_t hw = (int64_t)H * W;
    int64_t cap = hw < 65536 ? 65536 : hw;
    cap = (cap + 1023) / 1024 * 1024;
    return cap;
}

// launchers (each returns the cudaError_t of its launches)
cudaError_t launch_prep(const float* images, const void* labels, int label_dtype, int B, int S,
                        const Tables& t, bool convert_pixels, cudaStream_t st);
cudaError_t launch_estimate(const Tables& t, const EstParams& p, cudaStream_t st);
cudaError_t launch_big(const Tables& t, const EstParams& p, cudaStream_t st);
cudaError_t launch_reward(const Tables& t, const EstParams& p, const double* image_sizes, int S,
                          double* rewards, cudaStream_t st);
cudaError_t launch_estimate_debug(const Tables& t, const EstParams& p, int v, int k, uint32_t* row_costs,
                                  uint8_t* filter_id, uint32_t* hist, unsigned long long* match,
                                  double* size_out, cudaStream_t st);
cudaError_t launch_big_debug(const Tables& t, const EstParams& p, int v, int k, uint32_t* row_costs,
                             uint8_t* filter_id, uint32_t* hist, unsigned long long* match,
                             double* size_out, cudaStream_t st);
cudaError_t launch_slice(const Tables& t, int v, int k, uint8_t* out, size_t cap, cudaStream_t st);
cudaError_t launch_compact_stats(const int64_t* labels, int H, int W, int K, int32_t* counts,
                                 int32_t* bboxes, cudaStream_t st);
cudaError_t launch_single_segment(const uint8_t* img, int W, int H, int channels, const int64_t* labels,
                                  int64_t k, int x0, int y0, int w, int h, const EstParams& p,
                                  double* out, uint32_t* row_costs, uint8_t* filter_id, uint32_t* hist,
                                  unsigned long long* match, uint8_t* filter_scratch, cudaStream_t st);
cudaError_t launch_ema(const double* rewards, int n, const double* sum_count, double momentum,
                       double* state, double* adv, cudaStream_t st);
cudaError_t launch_sum_count(const double* rewards, int n, double* sum_count, cudaStream_t st);
cudaError_t run_selftest(int* mismatches);

void count_launch(int n = 1);

}  // namespace icr
