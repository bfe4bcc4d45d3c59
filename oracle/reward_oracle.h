/* TEST INFRASTRUCTURE ONLY -- never linked into or called from the product path.
 *
 * CPU restatement (plain C, scalar, single thread) of the compression-reward hot path of
 * moritzschramm/image-compression-cpp.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.
 *
 * Parity pin: the reference ships NO tests, golden vectors or fixtures for this path
 * (SURVEY.md section 4).  The restatement is pinned instead against the reference's own
 * three .cu files compiled unchanged (oracle/build_ref.sh -> oracle/_ref/libref_reward.so)
 * and run on the GPU box by tests/test_reference_parity.py; outputs of that run are
 * committed under tests/golden/.  Until that test has run green the status is
 * "parity unpinned".
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference).
 */
#ifndef REWARD_ORACLE_H
#define REWARD_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int32_t min_pixels;     /* include/compute_rewards.cuh:9  default 1 */
    int32_t L_min;          /* :10 default 4 */
    float beta;             /* :11 default 0.012167f */
    float b_match_token;    /* :12 default 18.0f */
    float gamma;            /* :13 default 0.1f */
    double overhead_base;   /* :14 default 9.308622 */
    int32_t adaptive_filter;/* :15 default 1 */
    double lambda;          /* :16 default 0.5 */
} orc_params;

void orc_default_params(orc_params* p);

/* Rule C  src/compute_rewards.cu:15-53 : planar f32 [3,H,W] -> packed RGBA u8 [H,W,4], A=255 */
void orc_chw3_f32_to_rgba(const float* chw, int H, int W, uint8_t* rgba);

/* Rule L  src/compute_rewards.cu:131-141 (at::_unique sorted+inverse): returns K, writes
 * compact ids (rank among sorted unique values) and, if uniq != NULL, the K sorted values. */
int64_t orc_compact_labels(const int64_t* labels, int64_t n, int64_t* compact, int64_t* uniq);

/* Rule S  src/segment_stats.cu:11-42 : counts[K], bboxes[K*4] = (x0,y0,x1,y1) inclusive */
void orc_segment_stats(const int64_t* compact, int H, int W, int K, int32_t* counts, int32_t* bboxes);

/* Rule P  src/image_slicer.cpp:48-74 == src/png_size_estimator.cu:88-117 : masked bbox crop,
 * out[h*w*4]; pixels whose id != k are zero in all four channels. */
void orc_slice(const uint8_t* rgba, const int64_t* compact, int W, int64_t k,
               int x0, int y0, int w, int h, uint8_t* out);

/* Debug tables of one segment estimate; any pointer may be NULL. */
typedef struct {
    uint32_t* row_costs;   /* [h*5]  Rule F costs (only written when adaptive) */
    uint8_t* filter_id;    /* [h]    Rule F choice */
    uint8_t* residuals;    /* [w*h*4] Rule R stream */
    uint32_t* hist;        /* [4*256] Rule H */
    uint64_t* match;       /* [3] = symbols, count, len_sum  Rule M */
    double* Hbar;          /* [1] */
} orc_debug;

/* Rules P,F,R,H,M,Z for one segment: src/png_size_estimator.cu:472-591. Returns size in bytes. */
double orc_estimate_segment(const uint8_t* rgba, const int64_t* compact, int W, int H,
                            int64_t k, int x0, int y0, int w, int h,
                            const orc_params* p, orc_debug* dbg);

/* Rules C..W for a batch: src/compute_rewards.cu:55-196.
 * images f32 [B,3,H,W]; labels int64 [B,H,W]; image_sizes f64 [B]; rewards f64 [B].
 * Optional per-image outputs with row stride Kstride (skipped when NULL):
 * K_out[B], counts[B*Kstride], bboxes[B*Kstride*4], seg_sizes[B*Kstride]. */
void orc_compute_rewards(const float* images, const int64_t* labels, const double* image_sizes,
                         int B, int H, int W, const orc_params* p, double* rewards,
                         int32_t* K_out, int Kstride, int32_t* counts, int32_t* bboxes,
                         double* seg_sizes);

/* Rule E  include/ema_baseline.hpp:11-21 : returns the new value; *initialized is updated. */
double orc_ema_update(double value, int* initialized, double momentum, const double* rewards, int n);

#ifdef __cplusplus
}
#endif
#endif
