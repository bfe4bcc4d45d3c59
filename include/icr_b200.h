/* icr_b200.h -- C ABI of the B200-native compression-reward path (libicr_b200.so).
 *
 * Drop-in boundary for ONE hot path of moritzschramm/image-compression-cpp:
 *   segment_stats -> masked slice -> png_size_estimator -> compute_rewards.
 * Plain pointers and sizes only; no torch types.  Every device pointer must belong to the
 * device the workspace was created on; every call is asynchronous on `stream` unless noted.
 * All entry points return 0 on success and a non-zero ICR_E* code otherwise; the message of
 * the last failure on the calling thread is icr_last_error().  (The reference reports errors
 * as C++ exceptions: TORCH_CHECK / std::runtime_error, src/compute_rewards.cu:70-75,
 * src/png_size_estimator.cu:483-484; the C++ shim in host/ maps codes back to those.)
 *
 * Citations are relative to the reference checkout (/root/reference).
 */
#ifndef ICR_B200_H
#define ICR_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ICR_OK 0
#define ICR_EINVAL 1      /* bad argument (null pointer, bad shape, channels outside 1..4) */
#define ICR_ECUDA 2       /* CUDA runtime error */
#define ICR_ECAPACITY 3   /* workspace too small for this call */
#define ICR_EDATA 4       /* data-dependent failure reported by icr_check (label range, K > Kmax) */

#define ICR_LABELS_I32 0  /* what rama_torch_batched hands over: src/rama_wrapper.cu:99 */
#define ICR_LABELS_I64 1  /* what compute_rewards_batched converts to: src/compute_rewards.cu:97 */
#define ICR_LABELS_U16 2  /* host buffers of icr_compute_rewards_host_u8 only */

typedef void* icr_stream;               /* a cudaStream_t */
typedef struct icr_workspace icr_workspace;

/* The eight estimator/reward knobs: default arguments of include/compute_rewards.cuh:9-16. */
typedef struct {
    int32_t min_pixels;      /* 1 */
    int32_t L_min;           /* 4 */
    float beta;              /* 0.012167f */
    float b_match_token;     /* 18.0f */
    float gamma;             /* 0.1f */
    double overhead_base;    /* 9.308622 */
    int32_t adaptive_filter; /* 1 */
    double lambda;           /* 0.5 */
} icr_params;

const char* icr_last_error(void);
int icr_version(void);
void icr_default_params(icr_params* p);

/* ---- workspace: replaces the hidden thread_local PngEstimatorWorkspace
 * (include/png_size_estimator.cuh:12-58, src/png_size_estimator.cu:466-469) and the per-image
 * at::empty calls of src/compute_rewards.cu:143-149.  V = B*S label maps of HxW pixels with at
 * most Kmax segments each.  Grow-only is the caller's business: create one per (thread, device,
 * shape). */
size_t icr_workspace_bytes(int V, int H, int W, int Kmax);
int icr_workspace_create(icr_workspace** ws, int device, int V, int H, int W, int Kmax);
int icr_workspace_destroy(icr_workspace* ws);

/* ---- the batch entry: compute_rewards_batched, src/compute_rewards.cu:55-196.
 * images  f32 [B,3,H,W] device, planar, byte values 0..255 (Rule C applied inside)
 * labels  [B,S,H,W] device, int32 or int64, any values whose per-map spread (max-min) is below
 *         the bitmap capacity icr_label_capacity(H,W); S label maps share one image (S=1 is the
 *         reference's call; S>1 equals images.repeat_interleave(S,0))
 * image_sizes f64 [B] device;  rewards f64 [B*S] device (out)
 * Asynchronous: no host synchronisation, no allocation. */
int icr_compute_rewards(const float* images, const void* labels, int label_dtype,
                        const double* image_sizes, int B, int S, int H, int W,
                        const icr_params* params, double* rewards,
                        icr_workspace* ws, icr_stream stream);

/* Sharded batches (one process per GPU; the reference is single-GPU and has no equivalent): the same call with the
 * path's only exchange fused into the reward writer.  gathered_of_rank_dev is a DEVICE array of `ranks` pointers, entry g
 * being rank g's gathered reward vector f64 [ranks*B*S] mapped into this process (CUDA IPC / symmetric memory over
 * NVLink); the reward of local map v is stored at [rank*B*S + v] of every one of them, this rank's own included.
 * rewards_local may be NULL.  flags_of_rank_dev (DEVICE array of `ranks` pointers to each rank's u64 [ranks] flag
 * array, peer-mapped the same way; may be NULL): when this rank's last reward has been stored, flag [rank] of every
 * rank is set to `step` (release at system scope), and icr_ema_update_gathered on the consumer side waits for the flags
 * of all ranks -- no barrier kernel, no collective.  `step` must grow by one per call; the caller alternates two
 * gathered vectors so that a fast rank cannot overwrite what a slow one still reads.  With flags NULL the caller
 * separates steps with its own barrier across ranks. */
int icr_compute_rewards_gather(const float* images, const void* labels, int label_dtype,
                               const double* image_sizes, int B, int S, int H, int W,
                               const icr_params* params, double* rewards_local,
                               double* const* gathered_of_rank_dev,
                               unsigned long long* const* flags_of_rank_dev, unsigned long long step,
                               int ranks, int rank, icr_workspace* ws, icr_stream stream);

/* Same through HOST buffers (pinned or pageable): H2D of images/labels/sizes, the path, D2H of
 * rewards, then a stream synchronise.  This is the call bench.py times as `e2e`.  Batches of 64 images
 * and more go over in four pieces, each copied on a second stream while the piece before it is
 * computed; the workspace tables then describe the last piece only, so the debug/slice entries refuse
 * to run after this call. */
int icr_compute_rewards_host(const float* images_h, const void* labels_h, int label_dtype,
                             const double* image_sizes_h, int B, int S, int H, int W,
                             const icr_params* params, double* rewards_h,
                             icr_workspace* ws, icr_stream stream);

/* The host-buffer entry for what a data loader holds BEFORE its float conversion
 * (include/fcn/EdgeDataset.h:73 makes the f32 tensor from 8-bit pixels): u8 images, planar [B,3,H,W]
 * (ICR_U8_PLANAR) or interleaved [B,H,W,3] in plane order (ICR_U8_INTERLEAVED), and labels as int32,
 * int64 or -- ICR_LABELS_U16, this entry only -- uint16 ids (any map with H*W <= 65536 fits after
 * ranking).  5 to 7 bytes per pixel cross the link instead of 16; a device kernel widens them into the
 * f32 planes / int32 labels of icr_compute_rewards.  (float)u8 is exact and Rule C maps it back to the
 * same byte, so the rewards equal those of icr_compute_rewards_host on the converted tensors bit for bit. */
#define ICR_U8_PLANAR 0
#define ICR_U8_INTERLEAVED 1
int icr_compute_rewards_host_u8(const uint8_t* images_u8_h, int image_layout, const void* labels_h, int label_dtype,
                                const double* image_sizes_h, int B, int S, int H, int W,
                                const icr_params* params, double* rewards_h,
                                icr_workspace* ws, icr_stream stream);

/* Synchronises `stream` and reports data-dependent failures of the calls issued so far on this
 * workspace: returns ICR_EDATA (and the first offending label map in *first_bad, may be NULL)
 * if a label map's value spread exceeded the capacity or its K exceeded Kmax; rewards of such
 * maps are NaN.  The reference has the same one-sync-per-call behaviour, only 2 per image. */
int icr_check(icr_workspace* ws, icr_stream stream, int* first_bad);
/* Same, with the reason: *status_bits = OR over the failed maps of 1 (label spread beyond the capacity: rank the labels first,
 * as at::_unique at src/compute_rewards.cu:132 would) and 2 (more segments than Kmax: *max_segments tells how many table
 * rows the largest such map needs).  Outputs may be NULL. */
int icr_check_detail(icr_workspace* ws, icr_stream stream, int* first_bad, int* status_bits, int* max_segments);
int64_t icr_label_capacity(int H, int W);

/* ---- segment_stats batched, with Rule L folded in (src/compute_rewards.cu:131-141 +
 * src/segment_stats.cu:11-77).  Fills the workspace tables and, when non-NULL, copies them out:
 * K_out int32 [V]; counts int32 [V,Kmax]; bboxes int32 [V,Kmax,4] = (x0,y0,x1,y1) inclusive;
 * compact_out int32 [V,H,W] = the compact ids at::_unique's inverse would hold (the batch path never
 * materialises them: it masks by comparing raw labels with the segment's key). */
int icr_segment_stats(const void* labels, int label_dtype, int V, int H, int W,
                      int32_t* K_out, int32_t* counts, int32_t* bboxes, int32_t* compact_out,
                      icr_workspace* ws, icr_stream stream);

/* Pointer-level drop-in for compute_counts_bboxes_from_compact_labels_cuda
 * (include/segment_stats.cuh:6-10): compact int64 labels [H,W] with values 0..K-1. */
int icr_counts_bboxes_from_compact_labels(const int64_t* labels_hw, int H, int W, int K,
                                          int32_t* counts_k, int32_t* bboxes_k4, icr_stream stream);

/* Pointer-level drop-in for estimate_png_size_masked_segment_to_output
 * (include/png_size_estimator.cuh:61-71): one segment, image u8 [H,W,channels], compact int64
 * labels, writes one double to out_dev.  w<=0 or h<=0 writes 0.0 (src/png_size_estimator.cu:485-488).
 * dbg_* (device, may be NULL): row costs u32 [h*5], filter ids u8 [h], histogram u32 [channels*256],
 * match triple u64 [3] -- tables the reference keeps private in its workspace. */
int icr_estimate_png_size_segment(const uint8_t* img_hwc_u8, int full_W, int full_H, int channels,
                                  const int64_t* labels_compact_hw, const int32_t* counts_k,
                                  int64_t seg_id_k, int x0, int y0, int w, int h,
                                  int32_t min_pixels, int L_min, float beta, float b_match_token,
                                  float gamma, double overhead_base, int adaptive_filter,
                                  double* out_dev,
                                  uint32_t* dbg_row_costs, uint8_t* dbg_filter_id,
                                  uint32_t* dbg_hist, unsigned long long* dbg_match,
                                  icr_stream stream);

/* The same estimate for callers that may not allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, same result as icr_estimate_png_size_segment; the h bytes
 * of filter-id scratch that entry allocates (cudaMallocAsync, grow-only) live here in a static device
 * array of ICR_NOALLOC_MAX_ROWS bytes that every GPU loading the library holds until exit.  Calls
 * no cudaMalloc / cudaFree or any allocating routine; h > ICR_NOALLOC_MAX_ROWS returns ICR_ECAPACITY.
 * The first call on a device resolves the array's address (cudaGetSymbolAddress): make one call
 * before capturing.  All calls on a device share the array (nothing in it has to be cleared: every
 * row's filter id is written before it is read), so calls that overlap on the GPU -- different
 * streams, or graph branches -- race; order them on one stream. */
#define ICR_NOALLOC_MAX_ROWS 16384
int icr_estimate_png_size_segment_noalloc(const uint8_t* img_hwc_u8, int full_W, int full_H, int channels,
                                          const int64_t* labels_compact_hw, const int32_t* counts_k,
                                          int64_t seg_id_k, int x0, int y0, int w, int h,
                                          int32_t min_pixels, int L_min, float beta, float b_match_token,
                                          float gamma, double overhead_base, int adaptive_filter,
                                          double* out_dev,
                                          uint32_t* dbg_row_costs, uint8_t* dbg_filter_id,
                                          uint32_t* dbg_hist, unsigned long long* dbg_match,
                                          icr_stream stream);

/* ---- read-back of the tables the last icr_compute_rewards / icr_segment_stats left in the
 * workspace (device pointers, valid until the next call): K int32 [V], counts int32 [V,Kmax],
 * bboxes int32 [V,Kmax,4], seg_sizes f64 [V,Kmax], keys i64 [V,Kmax] = original label of compact id k. */
int icr_workspace_tables(icr_workspace* ws, const int32_t** K, const int32_t** counts,
                         const int32_t** bboxes, const double** seg_sizes, const long long** keys,
                         int* Kmax);

/* Debug tables of ONE segment of the last batch call (label map v, compact id k; the caller's images and
 * labels of that call must still be alive), produced by the same device code the batch uses: row costs u32 [h*5], filter ids u8 [h], histogram u32 [4*256],
 * match u64 [3] (device pointers, may be NULL). */
int icr_debug_segment(icr_workspace* ws, int v, int k, const icr_params* params,
                      uint32_t* row_costs, uint8_t* filter_id, uint32_t* hist,
                      unsigned long long* match, double* size_out, icr_stream stream);

/* Masked slice of segment k of label map v (Rule P; src/image_slicer.cpp:48-74): u8 [h,w,4]. */
int icr_slice_segment(icr_workspace* ws, int v, int k, uint8_t* out_hw4, size_t out_capacity,
                      icr_stream stream);

/* ---- all slices of the last icr_compute_rewards batch in one packed buffer: what write_slices materialises one label at
 * a time on the CPU (src/image_slicer.cpp:81-125: K x {full-image mask, bbox scan, masked crop}).  Flat segment
 * f = (segments of the maps before v) + k, maps and compact ids ascending -- within a map that is ascending label value.
 * icr_slices_plan: offsets_dev u64 [capacity] receives the byte offset of slice f at [f] and the end at [segments];
 * summary_dev u64 [2] = {segments, total bytes}.  The caller reads summary back, sizes the buffer, then
 * icr_slices_emit writes slice f as u8 [h,w,4] at offsets[f]: (c0,c1,c2,255) inside the segment, (0,0,0,0) outside.
 * order: ICR_ORDER_PLANES keeps the plane order, ICR_ORDER_SWAPPED writes (c2,c1,c0,255) -- BGRA from RGB planes, the
 * cv::Mat layout of src/compress.cpp:145.  The caller's images and labels of that batch call must still be alive.
 * icr_slices_prepare runs only the label pass (Rules L and S) for a batch and records its buffers, for callers that
 * want slices without rewards (the `compress` executable, src/compress.cpp:138-147); follow it with icr_check. */
#define ICR_ORDER_PLANES 0
#define ICR_ORDER_SWAPPED 1
int icr_slices_prepare(const float* images_b3hw, const void* labels_vhw, int label_dtype, int B, int S, int H, int W,
                       icr_workspace* ws, icr_stream stream);
int icr_slices_plan(icr_workspace* ws, unsigned long long* offsets_dev, long long capacity,
                    unsigned long long* summary_dev, icr_stream stream);
int icr_slices_emit(icr_workspace* ws, const unsigned long long* offsets_dev, uint8_t* out_dev, size_t out_capacity,
                    int order, icr_stream stream);

/* ---- decoder side (src/reassemble.cpp:50-107): canvas u8 [H,W,4] starts at (0,0,0,0) and receives a slice where its
 * alpha > 0.  icr_reassemble_map clears the canvas and pastes every slice of label map v of the planned buffer in one
 * launch; icr_reassemble_batch does the same for all V label maps of the workspace's last batch at once (canvases u8
 * [V,H,W,4]: one clear and one launch instead of V of each; the reference decodes one image per process run);
 * icr_paste_slice pastes one slice from anywhere (a decoded file), clipped to the canvas, and leaves the rest of
 * the canvas alone -- pastes issued on one stream keep the reference's sequential order. */
int icr_reassemble_map(icr_workspace* ws, int v, const unsigned long long* offsets_dev, const uint8_t* slices_dev,
                       uint8_t* canvas_hw4, icr_stream stream);
int icr_reassemble_batch(icr_workspace* ws, const unsigned long long* offsets_dev, const uint8_t* slices_dev,
                         uint8_t* canvases_vhw4, icr_stream stream);
int icr_paste_slice(uint8_t* canvas_hw4, int W, int H, const uint8_t* slice_hw4, int x, int y, int w, int h,
                    icr_stream stream);

/* ---- Rule E + advantage (include/ema_baseline.hpp:11-21, src/fcn/training.cpp:178-181).
 * state (device, 2 doubles): {value, initialized flag}.  sum_count (device, 2 doubles) holds the
 * global {sum of rewards, count} -- after an all-reduce when the batch is sharded; pass NULL to
 * have it computed from rewards[0..n).  advantages (device, may be NULL) f64 [n]: (r - value)
 * standardised with the population std clamped at 1e-6. */
int icr_ema_update(const double* rewards, int n, const double* sum_count, double momentum,
                   double* state, double* advantages, icr_stream stream);
int icr_reward_sum_count(const double* rewards, int n, double* sum_count, icr_stream stream);
/* Rule E on a gathered vector filled through peer memory: first waits (on the device, bounded to ~10 s, after which the
 * baseline becomes NaN) until flags_local[g] >= step for every rank g, then does what icr_ema_update does. */
int icr_ema_update_gathered(const double* gathered, int n, const unsigned long long* flags_local, int ranks,
                            unsigned long long step, double momentum, double* state, double* advantages,
                            icr_stream stream);

/* ---- producer side of the label handoff: the copy loop of rama_torch_batched (src/rama_wrapper.cu:102-124, one
 * thrust::copy launch per sample from the solver's node_mapping vector into out[b]) as one launch per 256 samples.
 * node_mapping_dev_ptrs: HOST array of B device pointers to int32 [N]; node_mapping_sizes (host, may be NULL): their
 * lengths, each checked against N like the reference does (:116-117); labels_bn_out int32 [B,N] device, which
 * icr_compute_rewards takes as [B,H,W] with N = H*W.  Asynchronous, no allocation. */
int icr_pack_node_labels(const int32_t* const* node_mapping_dev_ptrs, const int64_t* node_mapping_sizes, int B, int64_t N,
                         int32_t* labels_bn_out, icr_stream stream);

/* ---- the exchange of a sharded batch (one process per GPU).  The reference is single-GPU:
 * include/ema_baseline.hpp:13 takes the mean over the whole batch, which on N GPUs needs every rank's rewards.
 * icr_exchange_rewards queues, on `stream` and inside one ncclGroupStart/End, an all-gather of the ranks' rewards
 * (gathered f64 [ranks*n_local] device, in rank order = image order for contiguous shards; may be NULL) and an
 * all-reduce of {sum, count} (sum_count f64 [2] device, out; may be NULL; this rank's part is computed here first).
 * Every rank passes the same n_local.  Asynchronous, no allocation.  `comm` is an ncclComm_t of the NCCL the process
 * already has loaded (libnccl.so.2 is bound with dlopen at the first call, ICR_NCCL_LIB overrides the path): the
 * caller's own communicator, or one made by icr_nccl_comm_create from an id that rank 0 obtained with
 * icr_nccl_unique_id and sent to the others by whatever means it has (128 bytes).  Create with the rank's device current. */
typedef void* icr_nccl_comm;
int icr_nccl_unique_id(unsigned char* id128);
int icr_nccl_comm_create(icr_nccl_comm* comm, const unsigned char* id128, int ranks, int rank);
int icr_nccl_comm_destroy(icr_nccl_comm comm);
int icr_nccl_comm_ranks(icr_nccl_comm comm, int* ranks);
int icr_exchange_rewards(icr_nccl_comm comm, const double* rewards_local, int n_local, double* gathered,
                         double* sum_count, icr_stream stream);

/* Exhaustive device self-test of the packed-byte filter arithmetic against scalar code
 * (all 2^24 Paeth triples, all Avg pairs, all cost bytes): *mismatches is a host int. Synchronous. */
int icr_selftest(int* mismatches);

/* Measurement hook: with profiling on, icr_compute_rewards records CUDA events on `stream` between its
 * kernels; icr_profile_read waits for the last call and returns the device time in ms of its three launches
 * {prep (Rules L, S + work lists), the estimator that ran (estimate_kernel, or narrow_kernel with ICR_ESTIMATOR=reg:
 *  bboxes up to 224 wide without match statistics), big_kernel (Rule M and wider segments, Rule W of every map,
 *  end-of-call bookkeeping), 0 (reserved)}. */
int icr_workspace_set_profile(icr_workspace* ws, int on);
int icr_profile_read(icr_workspace* ws, float* ms4);

/* Number of kernels this library launched on the calling thread since the last reset. */
long long icr_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif
