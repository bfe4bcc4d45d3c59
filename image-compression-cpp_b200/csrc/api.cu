// C ABI of libicr_b200.so (include/icr_b200.h): workspace, argument checks, launch sequencing.
// No torch, no allocation and no host synchronisation inside icr_compute_rewards.
#include "icr_internal.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

constexpr int HOST_CHUNKS = 4;     // pieces the host-buffer entry cuts a batch into

struct icr_workspace {
    int device, V, H, W, Kmax;
    char* base;            // one device allocation
    size_t bytes;
    icr::Tables t;
    icr::Inputs in;        // the caller's buffers of the last batch call (debug/slice entries read them)
    // device staging for the host-buffer entry
    float* st_images; void* st_labels; double* st_sizes; double* st_rewards;
    size_t st_images_bytes, st_labels_bytes;
    uint8_t* st_raw_images; uint16_t* st_raw_labels;      // u8 pixels / u16 labels as they came over the link (icr_compute_rewards_host_u8)
    size_t st_raw_images_bytes, st_raw_labels_bytes;
    cudaStream_t copy_stream;      // the host-buffer entry copies the next chunk on it while the current one is computed
    cudaEvent_t copy_ev[HOST_CHUNKS + 1];
    int profile;           // record events between the kernels of icr_compute_rewards
    cudaEvent_t ev[5];
    int dirty;             // a batch call failed between prep and the tail kernel: the work lists must be reset by hand
    int prefix_valid;      // seg_prefix matches the tables (built on demand for the slice entries)
};

namespace icr {

static thread_local std::string g_err;
static thread_local long long g_launches = 0;
void count_launch(int n) { g_launches += n; }

static int fail(int code, const std::string& msg) { g_err = msg; return code; }
int api_fail(int code, const std::string& msg) { return fail(code, msg); }      // for the other translation units (exchange.cu)
static int cuda_fail(cudaError_t e, const char* what) {
    return fail(ICR_ECUDA, std::string("CUDA error in ") + what + ": " + cudaGetErrorString(e));
}
#define ICR_CUDA(x, what) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return cuda_fail(e_, what); } while (0)

static size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }
// Big segments are only shared out among CTAs when the grid holds at least two CTAs per segment (at most BIG_GRID_MAX / 2
// of them); four slots per label map cover a background and a few more (a batch with more big segments runs them one CTA each)
static int merge_capacity(int V) { return 4 * (int64_t)V < BIG_GRID_MAX / 2 ? 4 * V : BIG_GRID_MAX / 2; }

// Which estimator takes the warp-path work lists.  Both produce the same tables (the GPU suite runs under either);
// ICR_ESTIMATOR=reg selects the register path of narrow.cu, anything else the shared-memory tile path of estimate.cu.
static bool use_register_path() {
    static const int v = [] { const char* e = std::getenv("ICR_ESTIMATOR"); return (e && std::string(e) == "reg") ? 1 : 0; }();
    return v != 0;
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && cached[dev]) return cached[dev];
    int n = 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) n = 148;
    if (dev >= 0 && dev < 64) cached[dev] = n;
    return n;
}

struct Layout { size_t K, status, counts, bboxes, keys, sizes, err, seg_prefix, sched, big_filter, clog2, heads, hslot, chg, slice_map_off,
                nbig, done, big_merge, list_a, list_b, list_big, total; };

static Layout layout(int V, int H, int W, int Kmax) {
    Layout l; size_t o = 0;
    l.K = o; o = align_up(o + (size_t)V * 4);
    l.status = o; o = align_up(o + (size_t)V * 4);
    l.counts = o; o = align_up(o + (size_t)V * Kmax * 4);
    l.bboxes = o; o = align_up(o + (size_t)V * Kmax * 16);
    l.keys = o; o = align_up(o + (size_t)V * Kmax * 8);
    l.sizes = o; o = align_up(o + (size_t)V * Kmax * 8);
    l.err = o; o = align_up(o + 16);
    l.seg_prefix = o; o = align_up(o + ((size_t)V + 1) * 4);
    l.sched = o; o = align_up(o + (size_t)SCHED_N * 4);
    l.big_filter = o; o = align_up(o + (size_t)BIG_GRID_MAX * H);
    l.clog2 = o; o = align_up(o + (size_t)CLOG2_N * 8);
    l.heads = o; o = align_up(o + (size_t)V * head_capacity(H, W) * 4);
    l.hslot = o; o = align_up(o + (size_t)V * head_capacity(H, W) * 4);
    l.chg = o; o = align_up(o + (chg_in_smem(H, W) ? 0 : (size_t)V * (size_t)chg_bytes(H, W)));
    l.slice_map_off = o; o = align_up(o + (size_t)V * 8);
    l.nbig = o; o = align_up(o + (size_t)V * 4);
    l.done = o; o = align_up(o + (size_t)V * 4);
    l.big_merge = o; o = align_up(o + (size_t)merge_capacity(V) * sizeof(BigMerge));
    l.list_a = o; o = align_up(o + (size_t)V * Kmax * 8);
    l.list_b = o; o = align_up(o + (size_t)V * Kmax * 8);
    l.list_big = o; o = align_up(o + (size_t)V * Kmax * 8);
    l.total = o;
    return l;
}

static EstParams widen(const icr_params* p) {
    EstParams e;
    e.min_pixels = p->min_pixels; e.L_min = p->L_min; e.adaptive = p->adaptive_filter ? 1 : 0;
    e.beta = (double)p->beta; e.token = (double)p->b_match_token; e.gamma = (double)p->gamma;   // float -> double as :588
    e.overhead = p->overhead_base; e.lambda = p->lambda;
    return e;
}

static int check_shape(const icr_workspace* ws, int V, int H, int W) {
    if (!ws) return fail(ICR_EINVAL, "workspace is null");
    if (V <= 0 || H <= 0 || W <= 0) return fail(ICR_EINVAL, "B, S, H and W must be positive");
    if (V > ws->V || H != ws->H || W != ws->W)
        return fail(ICR_ECAPACITY, "workspace was created for V=" + std::to_string(ws->V) + " maps of " + std::to_string(ws->H) +
                                       "x" + std::to_string(ws->W) + ", call needs V=" + std::to_string(V) + " of " +
                                       std::to_string(H) + "x" + std::to_string(W));
    return ICR_OK;
}

}  // namespace icr

using namespace icr;

extern "C" {

const char* icr_last_error(void) { return g_err.c_str(); }
int icr_version(void) { return 100; }
long long icr_launch_count(int reset) { long long n = g_launches; if (reset) g_launches = 0; return n; }

void icr_default_params(icr_params* p) {
    if (!p) return;
    p->min_pixels = 1; p->L_min = 4; p->beta = 0.012167f; p->b_match_token = 18.0f; p->gamma = 0.1f;
    p->overhead_base = 9.308622; p->adaptive_filter = 1; p->lambda = 0.5;
}

int64_t icr_label_capacity(int H, int W) { return label_capacity(H, W); }

size_t icr_workspace_bytes(int V, int H, int W, int Kmax) {
    if (V <= 0 || H <= 0 || W <= 0 || Kmax <= 0) return 0;
    return layout(V, H, W, Kmax).total;
}

int icr_workspace_create(icr_workspace** out, int device, int V, int H, int W, int Kmax) {
    if (!out) return fail(ICR_EINVAL, "null workspace out-pointer");
    *out = nullptr;
    if (V <= 0 || H <= 0 || W <= 0 || Kmax <= 0) return fail(ICR_EINVAL, "V, H, W and Kmax must be positive");
    if ((int64_t)H * W >= (1ll << 31))
        return fail(ICR_ECAPACITY, "label maps of 2^31 pixels and more are not supported (32-bit pixel offsets inside a map)");
    if ((int64_t)V * (Kmax < (int64_t)H * W ? Kmax : (int64_t)H * W) >= (1ll << 31))
        return fail(ICR_ECAPACITY, "V * Kmax must stay below 2^31 (32-bit positions in the work lists)");
    if ((int64_t)Kmax > (int64_t)H * W) Kmax = H * W;
    ICR_CUDA(cudaSetDevice(device), "cudaSetDevice");
    icr_workspace* ws = new icr_workspace();
    std::memset(ws, 0, sizeof *ws);
    ws->device = device; ws->V = V; ws->H = H; ws->W = W; ws->Kmax = Kmax;
    const Layout l = layout(V, H, W, Kmax);
    ws->bytes = l.total;
    cudaError_t e = cudaMalloc((void**)&ws->base, l.total);
    if (e != cudaSuccess) { delete ws; return cuda_fail(e, "cudaMalloc(workspace)"); }
    Tables& t = ws->t;
    t.K = (int32_t*)(ws->base + l.K); t.status = (int32_t*)(ws->base + l.status);
    t.keys = (long long*)(ws->base + l.keys); t.seg_prefix = (int32_t*)(ws->base + l.seg_prefix); t.sched = (int32_t*)(ws->base + l.sched);
    t.counts = (int32_t*)(ws->base + l.counts); t.bboxes = (int4*)(ws->base + l.bboxes); t.seg_sizes = (double*)(ws->base + l.sizes);
    t.err = (int32_t*)(ws->base + l.err); t.big_filter = (uint8_t*)(ws->base + l.big_filter); t.clog2 = (const double*)(ws->base + l.clog2);
    t.V = V; t.H = H; t.W = W; t.Kmax = Kmax;
    t.heads = (uint32_t*)(ws->base + l.heads); t.hslot = (uint32_t*)(ws->base + l.hslot); t.head_cap = head_capacity(H, W);
    t.chg_global = chg_in_smem(H, W) ? nullptr : (uint8_t*)(ws->base + l.chg); t.chg_stride = chg_bytes(H, W);
    t.slice_map_off = (unsigned long long*)(ws->base + l.slice_map_off);
    t.nbig = (int32_t*)(ws->base + l.nbig); t.done = (int32_t*)(ws->base + l.done);
    t.big_merge = (BigMerge*)(ws->base + l.big_merge); t.merge_cap = merge_capacity(V);
    t.list_a = (uint2*)(ws->base + l.list_a); t.list_b = (uint2*)(ws->base + l.list_b); t.list_big = (uint2*)(ws->base + l.list_big);
    t.list_cap = (int64_t)V * Kmax;
    std::vector<double> tab(CLOG2_N);
    tab[0] = 0.0;
    for (int c = 1; c < CLOG2_N; ++c) tab[c] = (double)c * std::log2((double)c);
    e = cudaMemcpy(ws->base + l.clog2, tab.data(), CLOG2_N * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemset(ws->base + l.K, 0, (size_t)V * 4);
    if (e == cudaSuccess) e = cudaMemset(ws->base + l.status, 0, (size_t)V * 4);
    if (e == cudaSuccess) e = cudaMemset(ws->base + l.err, 0, 16);
    if (e == cudaSuccess) e = cudaMemset(ws->base + l.seg_prefix, 0, ((size_t)V + 1) * 4);
    if (e == cudaSuccess) e = cudaMemset(ws->base + l.sched, 0, (size_t)SCHED_N * 4);
    if (e == cudaSuccess) e = cudaMemset(ws->base + l.done, 0, (size_t)V * 4);
    if (e == cudaSuccess) e = cudaMemset(ws->base + l.big_merge, 0, (size_t)merge_capacity(V) * sizeof(BigMerge));
    if (e != cudaSuccess) { cudaFree(ws->base); delete ws; return cuda_fail(e, "workspace init"); }
    *out = ws;
    return ICR_OK;
}

int icr_workspace_destroy(icr_workspace* ws) {
    if (!ws) return ICR_OK;
    cudaSetDevice(ws->device);
    cudaFree(ws->base);
    cudaFree(ws->st_images); cudaFree(ws->st_labels); cudaFree(ws->st_sizes); cudaFree(ws->st_rewards);
    cudaFree(ws->st_raw_images); cudaFree(ws->st_raw_labels);
    if (ws->ev[0]) for (int i = 0; i < 5; ++i) cudaEventDestroy(ws->ev[i]);
    if (ws->copy_stream) {
        for (int i = 0; i <= HOST_CHUNKS; ++i) cudaEventDestroy(ws->copy_ev[i]);
        cudaStreamDestroy(ws->copy_stream);
    }
    delete ws;
    return ICR_OK;
}

static int compute_rewards_impl(const float* images, const void* labels, int label_dtype, const double* image_sizes, int B, int S,
                                int H, int W, const icr_params* params, double* rewards, double* const* peer_rewards, int peer_count,
                                int peer_offset, unsigned long long* const* peer_flags, unsigned long long peer_step, int peer_rank,
                                icr_workspace* ws, icr_stream stream) {
    if (!images || !labels || !image_sizes || !params || (!rewards && !peer_rewards)) return fail(ICR_EINVAL, "null ptr");
    if (label_dtype != ICR_LABELS_I32 && label_dtype != ICR_LABELS_I64) return fail(ICR_EINVAL, "labels must be int32 or int64");
    if (B <= 0 || S <= 0) return fail(ICR_EINVAL, "B and S must be positive");
    if (int rc = check_shape(ws, B * S, H, W)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    Tables t = ws->t;
    t.V = B * S;
    t.peer_rewards = peer_rewards; t.peer_count = peer_count; t.peer_offset = peer_offset;
    t.peer_flags = peer_flags; t.peer_step = peer_step; t.peer_rank = peer_rank;
    const EstParams p = widen(params);
    const Inputs in{images, labels, label_dtype, S};
    ws->in = in;
    ws->t.V = t.V;
    const bool prof = ws->profile != 0;
    const RewardArgs rw{image_sizes, rewards, S};
    ws->prefix_valid = 0;
    if (ws->dirty) {             // an earlier call died between prep and the tail kernel: its work lists were never reset
        ICR_CUDA(cudaMemsetAsync(ws->t.sched, 0, (size_t)SCHED_N * 4, st), "reset work lists");
        ws->dirty = 0;
    }
    ws->dirty = 1;
    if (prof) cudaEventRecord(ws->ev[0], st);
    ICR_CUDA(launch_prep(labels, label_dtype, t, use_register_path() ? 2 : 1, p.L_min, rw, st), "prep");
    if (prof) cudaEventRecord(ws->ev[1], st);
    if (use_register_path()) ICR_CUDA(launch_narrow(t, in, p, rw, st), "narrow");
    else ICR_CUDA(launch_tile(t, in, p, rw, st), "estimate");
    if (prof) cudaEventRecord(ws->ev[2], st);
    ICR_CUDA(launch_big(t, in, p, rw, st), "big_segment");       // also resets the work lists and signals the peers
    if (prof) { cudaEventRecord(ws->ev[3], st); cudaEventRecord(ws->ev[4], st); }
    ws->dirty = 0;
    return ICR_OK;
}

int icr_compute_rewards(const float* images, const void* labels, int label_dtype, const double* image_sizes, int B, int S, int H,
                        int W, const icr_params* params, double* rewards, icr_workspace* ws, icr_stream stream) {
    if (!rewards) return fail(ICR_EINVAL, "null ptr");
    return compute_rewards_impl(images, labels, label_dtype, image_sizes, B, S, H, W, params, rewards, nullptr, 0, 0, nullptr, 0, 0, ws, stream);
}

int icr_compute_rewards_gather(const float* images, const void* labels, int label_dtype, const double* image_sizes, int B, int S,
                               int H, int W, const icr_params* params, double* rewards_local, double* const* gathered_of_rank_dev,
                               unsigned long long* const* flags_of_rank_dev, unsigned long long step, int ranks, int rank,
                               icr_workspace* ws, icr_stream stream) {
    if (!gathered_of_rank_dev) return fail(ICR_EINVAL, "null ptr");
    if (ranks <= 0 || rank < 0 || rank >= ranks) return fail(ICR_EINVAL, "rank must lie in [0, ranks)");
    return compute_rewards_impl(images, labels, label_dtype, image_sizes, B, S, H, W, params, rewards_local, gathered_of_rank_dev, ranks,
                                rank * B * S, flags_of_rank_dev, step, rank, ws, stream);
}

// The host-buffer entries.  image_kind: 0 f32 planar (the reference's tensor), 1 u8 planar, 2 u8 interleaved.
static int grow(void** p, size_t* have, size_t want, const char* what) {
    if (*have >= want) return ICR_OK;
    cudaFree(*p); *p = nullptr; *have = 0;
    ICR_CUDA(cudaMalloc(p, want), what);
    *have = want;
    return ICR_OK;
}

static int host_entry(const void* images_h, int image_kind, const void* labels_h, int label_dtype, const double* image_sizes_h,
                      int B, int S, int H, int W, const icr_params* params, double* rewards_h, icr_workspace* ws, icr_stream stream) {
    if (!images_h || !labels_h || !image_sizes_h || !rewards_h || !params) return fail(ICR_EINVAL, "null ptr");
    const bool u16 = label_dtype == ICR_LABELS_U16;
    if (label_dtype != ICR_LABELS_I32 && label_dtype != ICR_LABELS_I64 && !(u16 && image_kind != 0))
        return fail(ICR_EINVAL, image_kind ? "labels must be int32, int64 or uint16" : "labels must be int32 or int64");
    if (B <= 0 || S <= 0) return fail(ICR_EINVAL, "B and S must be positive");
    if (int rc = check_shape(ws, B * S, H, W)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t HW = (size_t)H * W;
    const size_t isz = image_kind ? 1 : 4;                                    // bytes per image element on the host
    const size_t lsz = label_dtype == ICR_LABELS_I64 ? 8 : (u16 ? 2 : 4);     // bytes per label on the host
    const size_t dev_lsz = label_dtype == ICR_LABELS_I64 ? 8 : 4;             // and on the device
    const int dev_dtype = u16 ? ICR_LABELS_I32 : label_dtype;
    if (int rc = grow((void**)&ws->st_images, &ws->st_images_bytes, (size_t)B * 3 * HW * 4, "cudaMalloc(staging images)")) return rc;
    if (int rc = grow(&ws->st_labels, &ws->st_labels_bytes, (size_t)B * S * HW * dev_lsz, "cudaMalloc(staging labels)")) return rc;
    if (image_kind)
        if (int rc = grow((void**)&ws->st_raw_images, &ws->st_raw_images_bytes, (size_t)B * 3 * HW, "cudaMalloc(staging u8 images)")) return rc;
    if (u16)
        if (int rc = grow((void**)&ws->st_raw_labels, &ws->st_raw_labels_bytes, (size_t)B * S * HW * 2, "cudaMalloc(staging u16 labels)")) return rc;
    if (!ws->st_sizes) {
        ICR_CUDA(cudaMalloc((void**)&ws->st_sizes, (size_t)ws->V * 8), "cudaMalloc(staging sizes)");
        ICR_CUDA(cudaMalloc((void**)&ws->st_rewards, (size_t)ws->V * 8), "cudaMalloc(staging rewards)");
    }
    if (!ws->copy_stream) {
        ICR_CUDA(cudaStreamCreateWithFlags(&ws->copy_stream, cudaStreamNonBlocking), "cudaStreamCreate(copy)");
        for (int i = 0; i <= HOST_CHUNKS; ++i) ICR_CUDA(cudaEventCreateWithFlags(&ws->copy_ev[i], cudaEventDisableTiming), "cudaEventCreate(copy)");
    }
    // The batch goes over in HOST_CHUNKS pieces: the copy of piece c+1 runs on a second stream under the kernels of piece
    // c, so the call costs the transfer plus the kernels of the last piece only.  The copies start after whatever the
    // caller's stream still has queued (an earlier call may be reading the staging buffers).
    const int nchunk = B >= 16 * HOST_CHUNKS ? HOST_CHUNKS : 1;
    char* img_dst = image_kind ? (char*)ws->st_raw_images : (char*)ws->st_images;
    char* lab_dst = u16 ? (char*)ws->st_raw_labels : (char*)ws->st_labels;
    ICR_CUDA(cudaEventRecord(ws->copy_ev[HOST_CHUNKS], st), "event record");
    ICR_CUDA(cudaStreamWaitEvent(ws->copy_stream, ws->copy_ev[HOST_CHUNKS], 0), "stream wait");
    ICR_CUDA(cudaMemcpyAsync(ws->st_sizes, image_sizes_h, (size_t)B * 8, cudaMemcpyHostToDevice, st), "H2D sizes");
    int b0[HOST_CHUNKS + 1];
    for (int c = 0; c <= nchunk; ++c) b0[c] = (int)((long long)B * c / nchunk);
    for (int c = 0; c < nchunk; ++c) {
        const size_t io = (size_t)b0[c] * 3 * HW, lo = (size_t)b0[c] * S * HW, n = (size_t)(b0[c + 1] - b0[c]);
        ICR_CUDA(cudaMemcpyAsync(img_dst + io * isz, (const char*)images_h + io * isz, n * 3 * HW * isz, cudaMemcpyHostToDevice,
                                 ws->copy_stream), "H2D images");
        ICR_CUDA(cudaMemcpyAsync(lab_dst + lo * lsz, (const char*)labels_h + lo * lsz, n * S * HW * lsz, cudaMemcpyHostToDevice,
                                 ws->copy_stream), "H2D labels");
        ICR_CUDA(cudaEventRecord(ws->copy_ev[c], ws->copy_stream), "event record");
    }
    for (int c = 0; c < nchunk; ++c) {
        const size_t io = (size_t)b0[c] * 3 * HW, lo = (size_t)b0[c] * S * HW, n = (size_t)(b0[c + 1] - b0[c]);
        ICR_CUDA(cudaStreamWaitEvent(st, ws->copy_ev[c], 0), "stream wait");
        if (image_kind || u16)
            ICR_CUDA(launch_expand(image_kind ? ws->st_raw_images + io : nullptr, image_kind == 2, ws->st_images + io, (int64_t)n, (int64_t)HW,
                                   u16 ? ws->st_raw_labels + lo : nullptr, (int32_t*)ws->st_labels + lo, (int64_t)(n * S * HW), st), "expand");
        if (int rc = icr_compute_rewards(ws->st_images + io, (const char*)ws->st_labels + lo * dev_lsz, dev_dtype, ws->st_sizes + b0[c],
                                         (int)n, S, H, W, params, ws->st_rewards + (size_t)b0[c] * S, ws, stream)) return rc;
    }
    if (nchunk > 1) ws->in = Inputs{nullptr, nullptr, 0, 1};      // the tables describe the last piece only: no debug/slice entry after this call
    ICR_CUDA(cudaMemcpyAsync(rewards_h, ws->st_rewards, (size_t)B * S * 8, cudaMemcpyDeviceToHost, st), "D2H rewards");
    ICR_CUDA(cudaStreamSynchronize(st), "stream sync");
    return ICR_OK;
}

int icr_compute_rewards_host(const float* images_h, const void* labels_h, int label_dtype, const double* image_sizes_h,
                             int B, int S, int H, int W, const icr_params* params, double* rewards_h,
                             icr_workspace* ws, icr_stream stream) {
    return host_entry(images_h, 0, labels_h, label_dtype, image_sizes_h, B, S, H, W, params, rewards_h, ws, stream);
}

int icr_compute_rewards_host_u8(const uint8_t* images_u8_h, int image_layout, const void* labels_h, int label_dtype,
                                const double* image_sizes_h, int B, int S, int H, int W, const icr_params* params,
                                double* rewards_h, icr_workspace* ws, icr_stream stream) {
    if (image_layout != ICR_U8_PLANAR && image_layout != ICR_U8_INTERLEAVED) return fail(ICR_EINVAL, "image_layout must be ICR_U8_PLANAR or ICR_U8_INTERLEAVED");
    return host_entry(images_u8_h, image_layout == ICR_U8_INTERLEAVED ? 2 : 1, labels_h, label_dtype, image_sizes_h, B, S, H, W, params,
                      rewards_h, ws, stream);
}

int icr_pack_node_labels(const int32_t* const* node_mapping_dev_ptrs, const int64_t* node_mapping_sizes, int B, int64_t N,
                         int32_t* labels_bn_out, icr_stream stream) {
    if (!node_mapping_dev_ptrs || !labels_bn_out) return fail(ICR_EINVAL, "null ptr");
    if (B <= 0 || N <= 0) return fail(ICR_EINVAL, "B and N must be positive");
    for (int b = 0; b < B; ++b) {
        if (!node_mapping_dev_ptrs[b]) return fail(ICR_EINVAL, "node_mapping " + std::to_string(b) + " is null");
        if (node_mapping_sizes && node_mapping_sizes[b] != N)                                            // src/rama_wrapper.cu:116-117
            return fail(ICR_EINVAL, "node_mapping " + std::to_string(b) + " has " + std::to_string(node_mapping_sizes[b]) +
                                        " entries, expected N = " + std::to_string(N));
    }
    ICR_CUDA(launch_pack_labels(node_mapping_dev_ptrs, B, N, labels_bn_out, (cudaStream_t)stream), "pack labels");
    return ICR_OK;
}

int icr_workspace_set_profile(icr_workspace* ws, int on) {
    if (!ws) return fail(ICR_EINVAL, "workspace is null");
    if (on && !ws->ev[0]) for (int i = 0; i < 5; ++i) ICR_CUDA(cudaEventCreate(&ws->ev[i]), "cudaEventCreate");
    ws->profile = on ? 1 : 0;
    return ICR_OK;
}

int icr_profile_read(icr_workspace* ws, float* ms4) {
    if (!ws || !ms4) return fail(ICR_EINVAL, "null ptr");
    if (!ws->ev[0]) return fail(ICR_EINVAL, "profiling was never enabled on this workspace");
    ICR_CUDA(cudaEventSynchronize(ws->ev[4]), "cudaEventSynchronize");
    for (int i = 0; i < 4; ++i) ICR_CUDA(cudaEventElapsedTime(&ms4[i], ws->ev[i], ws->ev[i + 1]), "cudaEventElapsedTime");
    return ICR_OK;
}

int icr_check_detail(icr_workspace* ws, icr_stream stream, int* first_bad, int* status_bits, int* max_segments) {
    if (!ws) return fail(ICR_EINVAL, "workspace is null");
    int32_t h[4] = {0, 0, 0, 0};
    ICR_CUDA(cudaMemcpyAsync(h, ws->t.err, 16, cudaMemcpyDeviceToHost, (cudaStream_t)stream), "D2H error flag");
    ICR_CUDA(cudaStreamSynchronize((cudaStream_t)stream), "stream sync");
    if (first_bad) *first_bad = h[0] ? 0x7fffffff - h[1] : -1;
    if (status_bits) *status_bits = h[0] ? h[2] : 0;
    if (max_segments) *max_segments = h[0] ? h[3] : 0;
    if (!h[0]) return ICR_OK;
    ICR_CUDA(cudaMemsetAsync(ws->t.err, 0, 16, (cudaStream_t)stream), "reset error flag");
    std::string why;
    if (h[2] & ST_LABEL_RANGE) why += "label spread exceeds icr_label_capacity(H,W)";
    if (h[2] & ST_K_OVERFLOW) why += std::string(why.empty() ? "" : "; ") + "segment count (up to " + std::to_string(h[3]) + ") exceeds Kmax";
    return fail(ICR_EDATA, "label map " + std::to_string(0x7fffffff - h[1]) + ": " + why + "; the rewards of such maps are NaN");
}

int icr_check(icr_workspace* ws, icr_stream stream, int* first_bad) { return icr_check_detail(ws, stream, first_bad, nullptr, nullptr); }

int icr_segment_stats(const void* labels, int label_dtype, int V, int H, int W, int32_t* K_out, int32_t* counts,
                      int32_t* bboxes, int32_t* compact_out, icr_workspace* ws, icr_stream stream) {
    if (!labels) return fail(ICR_EINVAL, "null ptr");
    if (label_dtype != ICR_LABELS_I32 && label_dtype != ICR_LABELS_I64) return fail(ICR_EINVAL, "labels must be int32 or int64");
    if (int rc = check_shape(ws, V, H, W)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    Tables t = ws->t;
    t.V = V;
    ws->in = Inputs{nullptr, labels, label_dtype, 1};
    ws->t.V = V;
    ws->prefix_valid = 0;
    ICR_CUDA(launch_prep(labels, label_dtype, t, 0, 4, RewardArgs{nullptr, nullptr, 1}, st), "prep");
    if (compact_out) ICR_CUDA(launch_compact_ids(labels, label_dtype, t, compact_out, st), "compact ids");
    if (K_out) ICR_CUDA(cudaMemcpyAsync(K_out, t.K, (size_t)V * 4, cudaMemcpyDeviceToDevice, st), "copy K");
    if (counts) ICR_CUDA(cudaMemcpyAsync(counts, t.counts, (size_t)V * t.Kmax * 4, cudaMemcpyDeviceToDevice, st), "copy counts");
    if (bboxes) ICR_CUDA(cudaMemcpyAsync(bboxes, t.bboxes, (size_t)V * t.Kmax * 16, cudaMemcpyDeviceToDevice, st), "copy bboxes");
    return ICR_OK;
}

int icr_counts_bboxes_from_compact_labels(const int64_t* labels_hw, int H, int W, int K, int32_t* counts_k,
                                          int32_t* bboxes_k4, icr_stream stream) {
    if (!labels_hw || !counts_k || !bboxes_k4) return fail(ICR_EINVAL, "null ptr");
    if (H <= 0 || W <= 0 || K < 0) return fail(ICR_EINVAL, "bad shape");
    ICR_CUDA(launch_compact_stats(labels_hw, H, W, K, counts_k, bboxes_k4, (cudaStream_t)stream), "compact stats");
    return ICR_OK;
}

int icr_estimate_png_size_segment(const uint8_t* img, int full_W, int full_H, int channels, const int64_t* labels,
                                  const int32_t* counts_k, int64_t seg_id_k, int x0, int y0, int w, int h,
                                  int32_t min_pixels, int L_min, float beta, float b_match_token, float gamma,
                                  double overhead_base, int adaptive_filter, double* out_dev, uint32_t* dbg_row_costs,
                                  uint8_t* dbg_filter_id, uint32_t* dbg_hist, unsigned long long* dbg_match,
                                  icr_stream stream) {
    if (!img || !labels || !counts_k || !out_dev) return fail(ICR_EINVAL, "null ptr");                 // :483
    if (channels < 1 || channels > 4) return fail(ICR_EINVAL, "channels must be 1..4");                // :484
    cudaStream_t st = (cudaStream_t)stream;
    if (w <= 0 || h <= 0) {                                                                             // :485-488
        ICR_CUDA(cudaMemsetAsync(out_dev, 0, sizeof(double), st), "memset out");
        return ICR_OK;
    }
    if (x0 < 0 || y0 < 0 || x0 + w > full_W || y0 + h > full_H) return fail(ICR_EINVAL, "bbox outside the image");
    icr_params prm{min_pixels, L_min, beta, b_match_token, gamma, overhead_base, adaptive_filter, 0.0};
    // per-(thread, device) grow-only filter-id scratch, like the reference's hidden workspace (:466-469)
    static thread_local uint8_t* scratch[16] = {nullptr};
    static thread_local int scratch_cap[16] = {0};
    int dev = 0;
    ICR_CUDA(cudaGetDevice(&dev), "cudaGetDevice");
    if (dev < 0 || dev >= 16) return fail(ICR_EINVAL, "device index above 15");
    if (scratch_cap[dev] < h) {
        // stream-ordered: the old scratch is released behind the launches that use it, nothing waits on the host
        if (scratch[dev]) ICR_CUDA(cudaFreeAsync(scratch[dev], st), "cudaFreeAsync(filter scratch)");
        scratch[dev] = nullptr; scratch_cap[dev] = 0;
        int cap = h < 4096 ? 4096 : h;
        ICR_CUDA(cudaMallocAsync((void**)&scratch[dev], (size_t)cap, st), "cudaMallocAsync(filter scratch)");
        scratch_cap[dev] = cap;
    }
    ICR_CUDA(launch_single_segment(img, full_W, full_H, channels, labels, seg_id_k, x0, y0, w, h, widen(&prm), out_dev,
                                   dbg_row_costs, dbg_filter_id, dbg_hist, dbg_match, scratch[dev], st), "single segment");
    return ICR_OK;
}

static_assert(NOALLOC_MAX_ROWS == ICR_NOALLOC_MAX_ROWS, "the header's limit is the array's size");

int icr_estimate_png_size_segment_noalloc(const uint8_t* img, int full_W, int full_H, int channels, const int64_t* labels,
                                          const int32_t* counts_k, int64_t seg_id_k, int x0, int y0, int w, int h,
                                          int32_t min_pixels, int L_min, float beta, float b_match_token, float gamma,
                                          double overhead_base, int adaptive_filter, double* out_dev, uint32_t* dbg_row_costs,
                                          uint8_t* dbg_filter_id, uint32_t* dbg_hist, unsigned long long* dbg_match,
                                          icr_stream stream) {
    if (!img || !labels || !counts_k || !out_dev) return fail(ICR_EINVAL, "null ptr");
    if (channels < 1 || channels > 4) return fail(ICR_EINVAL, "channels must be 1..4");
    cudaStream_t st = (cudaStream_t)stream;
    if (w <= 0 || h <= 0) {
        ICR_CUDA(cudaMemsetAsync(out_dev, 0, sizeof(double), st), "memset out");
        return ICR_OK;
    }
    if (x0 < 0 || y0 < 0 || x0 + w > full_W || y0 + h > full_H) return fail(ICR_EINVAL, "bbox outside the image");
    if (h > NOALLOC_MAX_ROWS)
        return fail(ICR_ECAPACITY, "icr_estimate_png_size_segment_noalloc: bbox of " + std::to_string(h) + " rows, the static scratch holds " +
                                       std::to_string(NOALLOC_MAX_ROWS) + " (ICR_NOALLOC_MAX_ROWS); use icr_estimate_png_size_segment");
    icr_params prm{min_pixels, L_min, beta, b_match_token, gamma, overhead_base, adaptive_filter, 0.0};
    ICR_CUDA(launch_single_segment_noalloc(img, full_W, full_H, channels, labels, seg_id_k, x0, y0, w, h, widen(&prm), out_dev,
                                           dbg_row_costs, dbg_filter_id, dbg_hist, dbg_match, st), "single segment (static scratch)");
    return ICR_OK;
}

int icr_workspace_tables(icr_workspace* ws, const int32_t** K, const int32_t** counts, const int32_t** bboxes,
                         const double** seg_sizes, const long long** keys, int* Kmax) {
    if (!ws) return fail(ICR_EINVAL, "workspace is null");
    if (K) *K = ws->t.K;
    if (counts) *counts = ws->t.counts;
    if (bboxes) *bboxes = (const int32_t*)ws->t.bboxes;
    if (seg_sizes) *seg_sizes = ws->t.seg_sizes;
    if (keys) *keys = ws->t.keys;
    if (Kmax) *Kmax = ws->t.Kmax;
    return ICR_OK;
}

int icr_debug_segment(icr_workspace* ws, int v, int k, const icr_params* params, uint32_t* row_costs,
                      uint8_t* filter_id, uint32_t* hist, unsigned long long* match, double* size_out,
                      icr_stream stream) {
    if (!ws || !params) return fail(ICR_EINVAL, "null ptr");
    if (v < 0 || v >= ws->t.V || k < 0 || k >= ws->Kmax) return fail(ICR_EINVAL, "segment index out of range");
    const EstParams p = widen(params);
    cudaStream_t st = (cudaStream_t)stream;
    // both kernels classify the segment on the device (segment_class); exactly one of them acts
    if (!ws->in.images) return fail(ICR_EINVAL, "no batch call has run on this workspace");
    if (use_register_path()) ICR_CUDA(launch_narrow_debug(ws->t, ws->in, p, v, k, row_costs, filter_id, hist, match, size_out, st), "narrow debug");
    else ICR_CUDA(launch_estimate_debug(ws->t, ws->in, p, v, k, row_costs, filter_id, hist, match, size_out, st), "estimate debug");
    ICR_CUDA(launch_big_debug(ws->t, ws->in, p, v, k, row_costs, filter_id, hist, match, size_out, st), "big debug");
    return ICR_OK;
}

int icr_slice_segment(icr_workspace* ws, int v, int k, uint8_t* out_hw4, size_t out_capacity, icr_stream stream) {
    if (!ws || !out_hw4) return fail(ICR_EINVAL, "null ptr");
    if (v < 0 || v >= ws->t.V || k < 0 || k >= ws->Kmax) return fail(ICR_EINVAL, "segment index out of range");
    if (!ws->in.images) return fail(ICR_EINVAL, "no batch call has run on this workspace");
    ICR_CUDA(launch_slice(ws->t, ws->in, v, k, out_hw4, out_capacity, (cudaStream_t)stream), "slice");
    return ICR_OK;
}

int icr_slices_prepare(const float* images, const void* labels, int label_dtype, int B, int S, int H, int W, icr_workspace* ws,
                       icr_stream stream) {
    if (!images || !labels) return fail(ICR_EINVAL, "null ptr");
    if (label_dtype != ICR_LABELS_I32 && label_dtype != ICR_LABELS_I64) return fail(ICR_EINVAL, "labels must be int32 or int64");
    if (B <= 0 || S <= 0) return fail(ICR_EINVAL, "B and S must be positive");
    if (int rc = check_shape(ws, B * S, H, W)) return rc;
    Tables t = ws->t;
    t.V = B * S;
    ws->in = Inputs{images, labels, label_dtype, S};
    ws->t.V = t.V;
    ws->prefix_valid = 0;
    ICR_CUDA(launch_prep(labels, label_dtype, t, 0, 4, RewardArgs{nullptr, nullptr, 1}, (cudaStream_t)stream), "prep");
    return ICR_OK;
}

int icr_slices_plan(icr_workspace* ws, unsigned long long* offsets_dev, long long capacity, unsigned long long* summary_dev,
                    icr_stream stream) {
    if (!ws || !offsets_dev || !summary_dev) return fail(ICR_EINVAL, "null ptr");
    if (capacity < 1) return fail(ICR_EINVAL, "offsets capacity must be at least 1");
    if (!ws->in.labels) return fail(ICR_EINVAL, "no batch call has run on this workspace");
    if (!ws->prefix_valid) {
        ICR_CUDA(launch_seg_prefix(ws->t, (cudaStream_t)stream), "segment prefix");
        ws->prefix_valid = 1;
    }
    ICR_CUDA(launch_slice_plan(ws->t, offsets_dev, capacity, summary_dev, (cudaStream_t)stream), "slice plan");
    return ICR_OK;
}

int icr_slices_emit(icr_workspace* ws, const unsigned long long* offsets_dev, uint8_t* out_dev, size_t out_capacity, int order,
                    icr_stream stream) {
    if (!ws || !offsets_dev || !out_dev) return fail(ICR_EINVAL, "null ptr");
    if (order != ICR_ORDER_PLANES && order != ICR_ORDER_SWAPPED) return fail(ICR_EINVAL, "order must be ICR_ORDER_PLANES or ICR_ORDER_SWAPPED");
    if (!ws->in.images) return fail(ICR_EINVAL, "no batch call with images has run on this workspace");
    ICR_CUDA(launch_slice_emit(ws->t, ws->in, offsets_dev, out_dev, out_capacity, order == ICR_ORDER_SWAPPED, (cudaStream_t)stream),
             "slice emit");
    return ICR_OK;
}

int icr_reassemble_map(icr_workspace* ws, int v, const unsigned long long* offsets_dev, const uint8_t* slices_dev,
                       uint8_t* canvas_hw4, icr_stream stream) {
    if (!ws || !offsets_dev || !slices_dev || !canvas_hw4) return fail(ICR_EINVAL, "null ptr");
    if (v < 0 || v >= ws->t.V) return fail(ICR_EINVAL, "label map index out of range");
    ICR_CUDA(launch_reassemble(ws->t, v, 1, offsets_dev, slices_dev, canvas_hw4, (cudaStream_t)stream), "reassemble");
    return ICR_OK;
}

int icr_reassemble_batch(icr_workspace* ws, const unsigned long long* offsets_dev, const uint8_t* slices_dev,
                         uint8_t* canvases_vhw4, icr_stream stream) {
    if (!ws || !offsets_dev || !slices_dev || !canvases_vhw4) return fail(ICR_EINVAL, "null ptr");
    ICR_CUDA(launch_reassemble(ws->t, 0, ws->t.V, offsets_dev, slices_dev, canvases_vhw4, (cudaStream_t)stream), "reassemble batch");
    return ICR_OK;
}

int icr_paste_slice(uint8_t* canvas_hw4, int W, int H, const uint8_t* slice_hw4, int x, int y, int w, int h, icr_stream stream) {
    if (!canvas_hw4 || !slice_hw4) return fail(ICR_EINVAL, "null ptr");
    if (W <= 0 || H <= 0 || w <= 0 || h <= 0 || x < 0 || y < 0) return fail(ICR_EINVAL, "bad slice or canvas geometry");
    ICR_CUDA(launch_paste(canvas_hw4, W, H, slice_hw4, x, y, w, h, (cudaStream_t)stream), "paste");
    return ICR_OK;
}

int icr_ema_update(const double* rewards, int n, const double* sum_count, double momentum, double* state,
                   double* advantages, icr_stream stream) {
    if (!rewards || !state) return fail(ICR_EINVAL, "null ptr");
    if (n <= 0) return fail(ICR_EINVAL, "n must be positive");
    ICR_CUDA(launch_ema(rewards, n, sum_count, momentum, state, advantages, nullptr, 0, 0, (cudaStream_t)stream), "ema");
    return ICR_OK;
}

int icr_ema_update_gathered(const double* gathered, int n, const unsigned long long* flags_local, int ranks, unsigned long long step,
                            double momentum, double* state, double* advantages, icr_stream stream) {
    if (!gathered || !state || !flags_local) return fail(ICR_EINVAL, "null ptr");
    if (n <= 0 || ranks <= 0 || ranks > 1024) return fail(ICR_EINVAL, "n and ranks must be positive (ranks at most 1024)");
    ICR_CUDA(launch_ema(gathered, n, nullptr, momentum, state, advantages, flags_local, ranks, step, (cudaStream_t)stream), "ema");
    return ICR_OK;
}

int icr_reward_sum_count(const double* rewards, int n, double* sum_count, icr_stream stream) {
    if (!rewards || !sum_count) return fail(ICR_EINVAL, "null ptr");
    if (n <= 0) return fail(ICR_EINVAL, "n must be positive");
    ICR_CUDA(launch_sum_count(rewards, n, sum_count, (cudaStream_t)stream), "sum_count");
    return ICR_OK;
}

int icr_selftest(int* mismatches) {
    if (!mismatches) return fail(ICR_EINVAL, "null ptr");
    ICR_CUDA(run_selftest(mismatches), "selftest");
    return ICR_OK;
}

}  // extern "C"
