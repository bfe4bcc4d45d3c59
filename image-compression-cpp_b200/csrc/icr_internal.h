// Internal declarations shared by the torch-free translation units of libicr_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/icr_b200.h"

// Bounds checks of our own (compute-sanitizer is not available on the GPU pool): a build with -DICR_CHECKED asserts every
// index into the shared-memory tables and the workspace lists (tools/build_variants.sh all:checked:-DICR_CHECKED, then
// ICR_LIB=tools/_variants/libicr_checked.so python -m pytest tests -m gpu).  A failed check aborts the kernel with file:line.
// Programmatic dependent launch (-DICR_PDL=0 turns it off): a kernel lets the next one of the call start its CTAs while its
// own last ones still run; the next one sets up its shared memory and then waits for the whole predecessor before it reads
// what that wrote.  4-6 us per call: 43 -> 37 us for a batch of 8 images, 228 -> 223 us for 512.
#ifndef ICR_PDL
#define ICR_PDL 1
#endif
#if ICR_PDL
#define ICR_GRID_DEP_WAIT() asm volatile("griddepcontrol.wait;" ::: "memory")
#define ICR_GRID_DEP_LAUNCH() asm volatile("griddepcontrol.launch_dependents;" ::: "memory")
#else
#define ICR_GRID_DEP_WAIT() ((void)0)
#define ICR_GRID_DEP_LAUNCH() ((void)0)
#endif
// launch `kern` as a programmatic dependent of the kernel before it in the stream (plain launch without ICR_PDL)
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_dependent(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
#if ICR_PDL
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
#else
    kern<<<grid, block, smem, st>>>(args...);
    return cudaSuccess;
#endif
}

#ifdef ICR_CHECKED
#include <cassert>
#define ICR_DCHECK(c) assert(c)
#else
#define ICR_DCHECK(c) ((void)0)
#endif

namespace icr {

// status bits per label map
enum : int { ST_OK = 0, ST_LABEL_RANGE = 1, ST_K_OVERFLOW = 2 };

constexpr int KS_SMEM = 2048;          // per-CTA shared-memory stats table entries in prep
constexpr int64_t LCAP_MAX = 1 << 19;  // label bitmap capacity (bits) that fits shared memory

// Merge slot of a big segment whose bands are shared out among several CTAs (big_segment.cu): the parts add their
// histograms and match statistics here, the part that finishes last reads them back, clears the slot and computes the size.
struct BigMerge { uint32_t hist[4 * 256]; unsigned long long ms, mc; uint32_t done, pad[3]; };
constexpr int BIG_MAX_PARTS = 16;      // most CTAs that share one big segment

struct Tables {            // device pointers into the workspace, V label maps
    int32_t* K;            // [V]
    int32_t* status;       // [V]
    int32_t* counts;       // [V,Kmax]
    int4* bboxes;          // [V,Kmax] (x0,y0,x1,y1) inclusive
    long long* keys;       // [V,Kmax] original label value of compact id k (ascending: Rule L)
    double* seg_sizes;     // [V,Kmax]
    int32_t* err;          // [2] {any-error flag, 0x7fffffff - first bad map}
    int32_t* seg_prefix;   // [V+1] exclusive scan of the segment counts of the maps with status OK (built on demand for the slices)
    int32_t* sched;        // [SCHED_N] work-list lengths, cursors and kernel-internal counters, see the SC_* indices
    uint8_t* big_filter;   // [BIG_GRID_MAX, H] filter ids of the segment a big-kernel CTA is working on
    uint32_t* heads;       // [V, head_cap] pixel index of every label-run head of a map, ascending (prep)
    uint32_t* hslot;       // [V, head_cap] bitmap slot of the head's label
    int head_cap;          // run heads a map's list holds (an average run of 8 pixels fits); beyond it prep walks the change bitmap
    uint8_t* chg_global;   // [V, chg_stride] change bitmap of a map when it does not fit shared memory, else NULL
    int64_t chg_stride;
    // Work lists prep leaves for the estimator kernels: entries {label map v, compact id k}.  The segments of the register
    // path (bbox no wider than NARROW_MAX_W, no match statistics) are kept apart by the code variant that will walk them,
    // widest first: list_a holds bboxes wider than 64 from the front and those up to 16 wide from the back, list_b those of
    // 33..64 from the front and 17..32 from the back.  The kernel takes them in that order, so the warps of an SM run the
    // same variant most of the time (its instruction cache holds one variant, not all of them) and the long segments
    // start first.  list_big: Rule M, wider or very large segments (CTA per segment).
    uint2* list_a;         // [list_cap]
    uint2* list_b;         // [list_cap]
    uint2* list_big;       // [list_cap]
    int64_t list_cap;
    int32_t* nbig;         // [V] segments of a map on list_big (prep)
    int32_t* done;         // [V] big-kernel segments of a map whose size has been written; the writer of the last one computes Rule W
    BigMerge* big_merge;   // [merge_cap] all zero between batch calls
    int merge_cap;
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

constexpr int CLOG2_N = 4096;
constexpr int NARROW_MAX_W = 224; // widest bbox of the warp paths (a one-row group must fit the tile path's shared-memory tile)
constexpr int BIG_GRID_MAX = 148 * 8;

// t.sched[]: list lengths are written by prep, cursors by the consuming kernels; the tail kernel's last CTA zeroes them
enum : int { SC_N_W128 = 0, SC_N_W64, SC_N_W32, SC_N_W16, SC_N_BIG, SC_CUR_NARROW, SC_CUR_BIG, SC_TAIL_DONE, SCHED_N = 8 };

// work list of a narrow segment by bbox width: 0: 65..NARROW_MAX_W, 1: 33..64, 2: 17..32, 3: 1..16
__host__ __device__ inline int width_class(int w) { return w > 64 ? 0 : (w > 32 ? 1 : (w > 16 ? 2 : 3)); }
constexpr int EST_MAX_AREA = 1 << 20;

// Segments the warp path leaves to the big-segment kernel: Rule M can fire (chunk = ceil(N/65536) >= L_min,
// src/png_size_estimator.cu:416-451), or the bbox is wider than four register chunks, or very large.
__host__ __device__ inline bool for_big_kernel(int w, int h, int L_min);

struct DebugOut {          // per-segment tables the debug entry fills (device pointers, may be NULL)
    int v, k;
    uint32_t* row_costs; uint8_t* filter_id; uint32_t* hist; unsigned long long* match; double* size;
};

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

// run-head list capacity per map: an average run of 8 pixels or more fits; beyond it prep walks the change bitmap instead
inline int head_capacity(int H, int W) {
    int64_t c = (int64_t)H * W / 8;
    return (int)(c < 1024 ? 1024 : c);
}

// bytes of a map's change bitmap (one bit per pixel, padded to whole 32-bit words)
inline int64_t chg_bytes(int H, int W) { return (((int64_t)H * W + 31) / 32) * 4; }

// shared memory prep needs next to its static arrays: presence bitmap + max(change bitmap, popcount prefix) + stats table.
// The change bitmap moves to global memory when that exceeds what a CTA may have.
constexpr int64_t PREP_SMEM_MAX = 200 * 1024;
inline bool chg_in_smem(int H, int W) {
    const int64_t cap_b = label_capacity(H, W) / 8, cb = chg_bytes(H, W);
    return cap_b + (cb > cap_b ? cb : cap_b) + (int64_t)KS_SMEM * 20 <= PREP_SMEM_MAX;
}

// launchers (each returns the cudaError_t of its launches)
// what the kernels that write segment sizes need for Rule W: whoever writes the last size of a map computes its reward
struct RewardArgs { const double* image_sizes; double* rewards; int S; };

// lists: 0 = tables only (icr_segment_stats, icr_slices_prepare); 1 = the estimator's work lists with one warp-path list in
// map order (tile path); 2 = warp-path lists by bbox width class (register path).
// Maps that cannot be processed get a NaN reward from prep itself when rw.rewards (or the peer vectors) are set.
cudaError_t launch_prep(const void* labels, int label_dtype, const Tables& t, int lists, int L_min, const RewardArgs& rw,
                        cudaStream_t st);
cudaError_t launch_compact_ids(const void* labels, int label_dtype, const Tables& t, int32_t* compact_out, cudaStream_t st);
cudaError_t launch_seg_prefix(const Tables& t, cudaStream_t st);
// the two estimators for the warp-path work lists: rows in registers (narrow.cu) or in a shared-memory tile (estimate.cu)
cudaError_t launch_narrow(const Tables& t, const Inputs& in, const EstParams& p, const RewardArgs& rw, cudaStream_t st);
cudaError_t launch_tile(const Tables& t, const Inputs& in, const EstParams& p, const RewardArgs& rw, cudaStream_t st);
cudaError_t launch_estimate_debug(const Tables& t, const Inputs& in, const EstParams& p, int v, int k, uint32_t* row_costs,
                                  uint8_t* filter_id, uint32_t* hist, unsigned long long* match,
                                  double* size_out, cudaStream_t st);
// always the last kernel of a batch call: Rule M segments, then its last CTA resets the work lists and signals the peers
cudaError_t launch_big(const Tables& t, const Inputs& in, const EstParams& p, const RewardArgs& rw, cudaStream_t st);
cudaError_t launch_narrow_debug(const Tables& t, const Inputs& in, const EstParams& p, int v, int k, uint32_t* row_costs,
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
// same launch with the filter-id scratch in a file-scope __device__ array (big_segment.cu): no allocation of any kind.
// h above NOALLOC_MAX_ROWS is the caller's error (checked in api.cu).
constexpr int NOALLOC_MAX_ROWS = 16384;
cudaError_t launch_single_segment_noalloc(const uint8_t* img, int W, int H, int channels, const int64_t* labels,
                                          int64_t k, int x0, int y0, int w, int h, const EstParams& p,
                                          double* out, uint32_t* row_costs, uint8_t* filter_id, uint32_t* hist,
                                          unsigned long long* match, cudaStream_t st);
cudaError_t launch_ema(const double* rewards, int n, const double* sum_count, double momentum,
                       double* state, double* adv, const unsigned long long* flags, int ranks, unsigned long long step,
                       cudaStream_t st);
cudaError_t launch_expand(const uint8_t* images_u8, int hwc, float* images_f32, int64_t B, int64_t HW, const uint16_t* labels_u16,
                          int32_t* labels_i32, int64_t n_labels, cudaStream_t st);
constexpr int PACK_MAX = 256;      // node_mapping pointers per launch of the label packer (kernel-parameter space)
cudaError_t launch_pack_labels(const int32_t* const* ptrs_host, int B, int64_t N, int32_t* out, cudaStream_t st);
cudaError_t launch_sum_count(const double* rewards, int n, double* sum_count, cudaStream_t st);
cudaError_t run_selftest(int* mismatches);

void count_launch(int n = 1);
int sm_count();            // multiprocessors of the current device (cached per device)

__host__ __device__ inline bool for_big_kernel(int w, int h, int L_min) {
    if (w <= 0 || h <= 0) return false;
    long long N = 4ll * w * h;
    return (N + 65535) / 65536 >= (long long)L_min || w > NARROW_MAX_W || (long long)w * h > EST_MAX_AREA;
}

// which kernel takes a segment with a w x h bbox
enum : int { CLS_NARROW = 0, CLS_BIG = 1 };
__host__ __device__ inline int segment_class(int w, int h, int L_min) { return for_big_kernel(w, h, L_min) ? CLS_BIG : CLS_NARROW; }

}  // namespace icr
