// Host C++ side of the reward path: the reference's call surface over libtorch tensors, implemented on
// the C ABI of libicr_b200.so (include/icr_b200.h).  Replaces, as drop-ins:
//   compute_rewards_batched                          src/compute_rewards.cu:55-196
//   compute_counts_bboxes_from_compact_labels_cuda   src/segment_stats.cu:44-77
//   estimate_png_size_masked_segment_to_output       src/png_size_estimator.cu:472-591
//   EMABaseline::update                              include/ema_baseline.hpp:11-21
// and the label-map handoff after rama_torch_batched  src/fcn/training.cpp:169-172.
// Tensors are device-memory handles here; every byte of arithmetic happens in the CUDA kernels behind the
// C ABI, on the current PyTorch stream.  Compiled by g++ (no nvcc): this file holds no device code.
//
// Error convention, as in the reference: argument errors are TORCH_CHECK failures (c10::Error), CUDA and
// null-pointer errors of the pointer-level estimator are std::runtime_error.
#include <torch/torch.h>
#include <ATen/ops/_unique.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <tuple>

#include "host_internal.hpp"
#include "include/compute_rewards.cuh"
#include "include/ema_baseline.hpp"
#include "include/png_size_estimator.cuh"
#include "include/rama_handoff.hpp"
#include "include/segment_stats.cuh"

using icrh::WsHolder;
using icrh::current_stream;
using icrh::label_code;
using icrh::throw_runtime;
using icrh::workspace_for;

namespace icrh {

using WsKey = std::tuple<int, int, int>;
thread_local std::map<WsKey, std::unique_ptr<WsHolder>> g_ws;

WsHolder& workspace_for(int device, int V, int H, int W, int Kmax) {
    auto& slot = g_ws[WsKey(device, H, W)];
    if (!slot) slot = std::make_unique<WsHolder>();
    WsHolder& h = *slot;
    const int64_t hw = (int64_t)H * W;
    if (Kmax > hw) Kmax = (int)hw;
    if (!h.ws || h.V < V || h.Kmax < Kmax) {
        if (h.ws) {
            // kernels of earlier calls may still read the old tables
            c10::cuda::getCurrentCUDAStream().synchronize();
            icr_workspace_destroy(h.ws);
            h.ws = nullptr;
        }
        const int nV = std::max(V, h.V), nK = std::max(Kmax, h.Kmax);
        if (icr_workspace_create(&h.ws, device, nV, H, W, nK) != ICR_OK) { h.V = h.Kmax = 0; TORCH_CHECK(false, icr_last_error()); }
        h.V = nV; h.Kmax = nK;
    }
    return h;
}

}  // namespace icrh

namespace icrh {

// labels whose value spread defeats the presence bitmap: the reference accepts any integers (it sorts,
// src/compute_rewards.cu:131-134), so fall back to ranking them with at::_unique per map and rerun on the ranks
torch::Tensor rank_labels(const torch::Tensor& labs_vhw) {
    auto out = torch::empty(labs_vhw.sizes(), labs_vhw.options().dtype(torch::kInt64));
    for (int64_t v = 0; v < labs_vhw.size(0); ++v) {
        auto inv = std::get<1>(at::_unique(labs_vhw[v].reshape({-1}), /*sorted=*/true, /*return_inverse=*/true));
        out[v].copy_(inv.view({labs_vhw.size(1), labs_vhw.size(2)}));
    }
    return out;
}

// What to change before the next attempt after icr_check_detail reported data the fast path could not take: more table
// rows when a map held more segments than Kmax (as many as the largest such map needs, not H*W), ranked labels when a
// map's label spread exceeded the bitmap.  Returns false when neither applies (give up).
bool plan_retry(int status_bits, int max_segments, int64_t HW, int& Kmax, bool& rank) {
    bool changed = false;
    if ((status_bits & 2) && Kmax < HW) {
        int64_t want = std::max<int64_t>(max_segments, 2 * (int64_t)Kmax);
        Kmax = (int)std::min<int64_t>(HW, want);
        changed = true;
    }
    if ((status_bits & 1) && !rank) { rank = true; changed = true; }
    return changed;
}

}  // namespace icrh

namespace {

using icrh::KMAX_FIRST;
using icrh::rank_labels;


icr_params make_params(int32_t min_pixels, int L_min, float beta, float token, float gamma, double overhead, bool adaptive,
                       double lambda) {
    icr_params p;
    p.min_pixels = min_pixels; p.L_min = L_min; p.beta = beta; p.b_match_token = token; p.gamma = gamma;
    p.overhead_base = overhead; p.adaptive_filter = adaptive ? 1 : 0; p.lambda = lambda;
    return p;
}

torch::Tensor rewards_impl(const torch::Tensor& images_bchw_f32, const torch::Tensor& labels, const torch::Tensor& image_sizes,
                           int64_t S, const icr_params& prm) {
    torch::NoGradGuard ng;
    const auto dev = labels.device();
    c10::cuda::CUDAGuard dg(dev);

    torch::Tensor images = images_bchw_f32;
    if (!images.is_cuda() || images.device() != dev) images = images.to(dev, /*non_blocking=*/true);
    images = images.contiguous();
    const int64_t B = images.size(0), H = images.size(2), W = images.size(3);

    // int32 (the RAMA handoff) and int64 planes are read as delivered; anything else is widened like the reference does
    torch::Tensor labs = labels;
    if (labs.scalar_type() != torch::kInt64 && labs.scalar_type() != torch::kInt32) labs = labs.to(torch::kInt64);
    labs = labs.contiguous();

    torch::Tensor sizes = image_sizes.view({-1});
    if (sizes.scalar_type() != torch::kFloat64) sizes = sizes.to(torch::kFloat64);
    sizes = sizes.contiguous();

    auto rewards = torch::empty({B * S}, torch::TensorOptions().device(dev).dtype(torch::kFloat64));
    if (B == 0) return rewards;
    const int V = (int)(B * S);
    const icr_stream st = current_stream();

    int Kmax = (int)std::min<int64_t>(H * W, KMAX_FIRST);
    bool rank = false, ranked = false;
    for (int attempt = 0; attempt < 4; ++attempt) {
        if (rank && !ranked) { labs = rank_labels(labs.view({(int64_t)V, H, W})).view(labs.sizes()); ranked = true; }
        WsHolder& h = workspace_for(dev.index(), V, (int)H, (int)W, Kmax);
        const int rc = icr_compute_rewards(images.data_ptr<float>(), labs.data_ptr(), label_code(labs), sizes.data_ptr<double>(),
                                           (int)B, (int)S, (int)H, (int)W, &prm, rewards.data_ptr<double>(), h.ws, st);
        if (rc == ICR_ECUDA) throw_runtime();
        TORCH_CHECK(rc == ICR_OK, icr_last_error());
        // the reference synchronises inside this call too (one blocking D2H per image, src/compute_rewards.cu:152); here it is
        // one 16-byte read per batch, which also tells why a label map could not be taken by the fast path
        int bits = 0, maxk = 0;
        const int chk = icr_check_detail(h.ws, st, nullptr, &bits, &maxk);
        if (chk == ICR_OK) return rewards;
        if (chk == ICR_ECUDA) throw_runtime();
        if (!icrh::plan_retry(bits, maxk, H * W, Kmax, rank)) break;
    }
    TORCH_CHECK(false, "compute_rewards_batched: ", icr_last_error());
}

}  // namespace

// ------------------------------------------------------------------------------------------------ compute_rewards
torch::Tensor compute_rewards_batched(const torch::Tensor& images_bchw_f32, const torch::Tensor& labels_bhw,
                                      const torch::Tensor& image_sizes_b, int32_t min_pixels_per_segment, int L_min, float beta,
                                      float b_match_token, float gamma, double overhead_base, bool adaptive_filter, double lambda) {
    TORCH_CHECK(images_bchw_f32.dim() == 4 && images_bchw_f32.size(1) == 3, "images must be [B,3,H,W]");
    TORCH_CHECK(images_bchw_f32.scalar_type() == torch::kFloat32, "images must be float32 (CV_32FC3)");
    TORCH_CHECK(labels_bhw.is_cuda(), "labels must be CUDA");
    TORCH_CHECK(image_sizes_b.is_cuda(), "image_sizes must be CUDA");
    TORCH_CHECK(labels_bhw.dim() == 3, "labels must be [B,H,W]");
    TORCH_CHECK(image_sizes_b.numel() == images_bchw_f32.size(0), "image_sizes must have B elements");
    TORCH_CHECK(labels_bhw.size(0) == images_bchw_f32.size(0) && labels_bhw.size(1) == images_bchw_f32.size(2) &&
                    labels_bhw.size(2) == images_bchw_f32.size(3),
                "labels must match images: [B,H,W]");
    TORCH_CHECK(c10::isIntegralType(labels_bhw.scalar_type(), /*includeBool=*/false), "labels must be an integer tensor");
    return rewards_impl(images_bchw_f32, labels_bhw, image_sizes_b, 1,
                        make_params(min_pixels_per_segment, L_min, beta, b_match_token, gamma, overhead_base, adaptive_filter, lambda));
}

torch::Tensor compute_rewards_multi(const torch::Tensor& images_bchw_f32, const torch::Tensor& labels_bshw,
                                    const torch::Tensor& image_sizes_b, int32_t min_pixels_per_segment, int L_min, float beta,
                                    float b_match_token, float gamma, double overhead_base, bool adaptive_filter, double lambda) {
    TORCH_CHECK(images_bchw_f32.dim() == 4 && images_bchw_f32.size(1) == 3, "images must be [B,3,H,W]");
    TORCH_CHECK(images_bchw_f32.scalar_type() == torch::kFloat32, "images must be float32 (CV_32FC3)");
    TORCH_CHECK(labels_bshw.is_cuda(), "labels must be CUDA");
    TORCH_CHECK(image_sizes_b.is_cuda(), "image_sizes must be CUDA");
    TORCH_CHECK(labels_bshw.dim() == 4, "labels must be [B,S,H,W]");
    TORCH_CHECK(image_sizes_b.numel() == images_bchw_f32.size(0), "image_sizes must have B elements");
    TORCH_CHECK(labels_bshw.size(0) == images_bchw_f32.size(0) && labels_bshw.size(2) == images_bchw_f32.size(2) &&
                    labels_bshw.size(3) == images_bchw_f32.size(3),
                "labels must match images: [B,S,H,W]");
    TORCH_CHECK(c10::isIntegralType(labels_bshw.scalar_type(), false), "labels must be an integer tensor");
    const int64_t S = labels_bshw.size(1);
    TORCH_CHECK(S > 0, "labels must hold at least one map per image");
    return rewards_impl(images_bchw_f32, labels_bshw, image_sizes_b, S,
                        make_params(min_pixels_per_segment, L_min, beta, b_match_token, gamma, overhead_base, adaptive_filter, lambda))
        .view({images_bchw_f32.size(0), S});
}

// ------------------------------------------------------------------------------------------------ segment_stats
void compute_counts_bboxes_from_compact_labels_cuda(const torch::Tensor& labels_compact_hw_i64, torch::Tensor& counts_k_i32,
                                                    torch::Tensor& bboxes_k4_i32) {
    TORCH_CHECK(labels_compact_hw_i64.is_cuda(), "labels must be CUDA");
    TORCH_CHECK(labels_compact_hw_i64.scalar_type() == torch::kInt64, "labels must be int64");
    TORCH_CHECK(labels_compact_hw_i64.dim() == 2, "labels must be [H,W]");
    TORCH_CHECK(counts_k_i32.is_cuda() && bboxes_k4_i32.is_cuda(), "outputs must be CUDA");
    TORCH_CHECK(counts_k_i32.scalar_type() == torch::kInt32, "counts must be int32");
    TORCH_CHECK(bboxes_k4_i32.scalar_type() == torch::kInt32, "bboxes must be int32");
    TORCH_CHECK(bboxes_k4_i32.dim() == 2 && bboxes_k4_i32.size(1) == 4, "bboxes must be [K,4]");
    // the reference hands data_ptr()s to its kernels, i.e. assumes dense layouts; make that explicit
    TORCH_CHECK(labels_compact_hw_i64.is_contiguous() && counts_k_i32.is_contiguous() && bboxes_k4_i32.is_contiguous(),
                "labels, counts and bboxes must be contiguous");
    c10::cuda::CUDAGuard dg(labels_compact_hw_i64.device());
    const int rc = icr_counts_bboxes_from_compact_labels(labels_compact_hw_i64.data_ptr<int64_t>(), (int)labels_compact_hw_i64.size(0),
                                                         (int)labels_compact_hw_i64.size(1), (int)counts_k_i32.numel(),
                                                         counts_k_i32.data_ptr<int32_t>(), bboxes_k4_i32.data_ptr<int32_t>(),
                                                         current_stream());
    if (rc != ICR_OK) throw_runtime();       // CUDA_CHECK -> std::runtime_error in the reference (src/segment_stats.cu:7-8)
}

std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor> segment_stats_batched(const torch::Tensor& labels_vhw,
                                                                                              int64_t Kmax) {
    TORCH_CHECK(labels_vhw.is_cuda(), "labels must be CUDA");
    TORCH_CHECK(labels_vhw.dim() == 3, "labels must be [V,H,W]");
    TORCH_CHECK(c10::isIntegralType(labels_vhw.scalar_type(), false), "labels must be an integer tensor");
    torch::NoGradGuard ng;
    c10::cuda::CUDAGuard dg(labels_vhw.device());
    torch::Tensor labs = labels_vhw;
    if (labs.scalar_type() != torch::kInt64 && labs.scalar_type() != torch::kInt32) labs = labs.to(torch::kInt64);
    labs = labs.contiguous();
    const int64_t V = labs.size(0), H = labs.size(1), W = labs.size(2);
    if (Kmax <= 0 || Kmax > H * W) Kmax = H * W;
    WsHolder& h = workspace_for(labs.device().index(), (int)V, (int)H, (int)W, (int)Kmax);
    const icr_stream st = current_stream();
    int rc = icr_segment_stats(labs.data_ptr(), label_code(labs), (int)V, (int)H, (int)W, nullptr, nullptr, nullptr, nullptr, h.ws, st);
    if (rc == ICR_ECUDA) throw_runtime();
    TORCH_CHECK(rc == ICR_OK, icr_last_error());
    rc = icr_check(h.ws, st, nullptr);
    if (rc == ICR_ECUDA) throw_runtime();
    TORCH_CHECK(rc == ICR_OK, icr_last_error());
    const int32_t *pK, *pc, *pb; const long long* pk; int km = 0;
    icr_workspace_tables(h.ws, &pK, &pc, &pb, nullptr, &pk, &km);
    auto i32 = torch::TensorOptions().device(labs.device()).dtype(torch::kInt32);
    // views of the workspace tables, cloned so they survive the next call
    auto K = torch::from_blob((void*)pK, {V}, i32).clone();
    auto counts = torch::from_blob((void*)pc, {V, (int64_t)km}, i32).clone();
    auto bboxes = torch::from_blob((void*)pb, {V, (int64_t)km, 4}, i32).clone();
    auto keys = torch::from_blob((void*)pk, {V, (int64_t)km}, i32.dtype(torch::kInt64)).clone();
    return {K, counts, bboxes, keys};
}

// ------------------------------------------------------------------------------------------------ png_size_estimator
void estimate_png_size_masked_segment_to_output(const uint8_t* img_hwc_u8, int full_W, int full_H, int channels,
                                                const int64_t* labels_compact_hw, const int32_t* counts_k, int64_t seg_id_k, int x0,
                                                int y0, int w, int h, int32_t min_pixels, int L_min, float beta, float b_match_token,
                                                float gamma, double overhead_base, bool adaptive_filter, double* out_dev) {
    const int rc = icr_estimate_png_size_segment(img_hwc_u8, full_W, full_H, channels, labels_compact_hw, counts_k, seg_id_k, x0, y0, w,
                                                 h, min_pixels, L_min, beta, b_match_token, gamma, overhead_base, adaptive_filter ? 1 : 0,
                                                 out_dev, nullptr, nullptr, nullptr, nullptr, current_stream());
    if (rc != ICR_OK) throw_runtime();       // "null ptr", "channels must be 1..4", "CUDA error: ..." (src/png_size_estimator.cu:483-484)
}

// ------------------------------------------------------------------------------------------------ label handoff
torch::Tensor labels_from_rama(const torch::Tensor& node_labels_bn_i32, int64_t H, int64_t W) {
    TORCH_CHECK(node_labels_bn_i32.device().is_cuda(), "node_labels must be a CUDA tensor");
    TORCH_CHECK(node_labels_bn_i32.is_contiguous(), "node_labels must be contiguous");
    TORCH_CHECK(node_labels_bn_i32.scalar_type() == torch::kInt32, "node_labels must be int32");
    TORCH_CHECK(node_labels_bn_i32.dim() == 2 && node_labels_bn_i32.size(1) == H * W, "node_labels must be [B, H*W]");
    return node_labels_bn_i32.view({node_labels_bn_i32.size(0), H, W});
}

torch::Tensor labels_from_node_mappings(const std::vector<torch::Tensor>& node_mappings, int64_t H, int64_t W) {
    TORCH_CHECK(!node_mappings.empty(), "no node_mapping given");
    TORCH_CHECK(H > 0 && W > 0, "Invalid number of nodes");                                  // :96
    const int64_t B = (int64_t)node_mappings.size(), N = H * W;
    const auto dev = node_mappings[0].device();
    std::vector<const int32_t*> ptrs(B);
    std::vector<int64_t> sizes(B);
    for (int64_t b = 0; b < B; ++b) {
        const auto& m = node_mappings[b];
        TORCH_CHECK(m.is_cuda() && m.device() == dev, "node_mappings must be on one CUDA device");
        TORCH_CHECK(m.scalar_type() == torch::kInt32 && m.is_contiguous(), "node_mappings must be contiguous int32");
        TORCH_CHECK(m.numel() == N, "rama_cuda returned node_mapping of unexpected size");    // :116-117
        ptrs[b] = m.data_ptr<int32_t>();
        sizes[b] = m.numel();
    }
    c10::cuda::CUDAGuard dg(dev);
    auto out = torch::empty({B, H, W}, node_mappings[0].options());
    const int rc = icr_pack_node_labels(ptrs.data(), sizes.data(), (int)B, N, out.data_ptr<int32_t>(), current_stream());
    if (rc == ICR_ECUDA) throw_runtime();
    TORCH_CHECK(rc == ICR_OK, icr_last_error());
    return out;
}

// ------------------------------------------------------------------------------------------------ EMA baseline
torch::Tensor EMABaseline::run(const torch::Tensor& rewards, const torch::Tensor* sum_count, torch::Tensor* adv) {
    torch::NoGradGuard ng;
    TORCH_CHECK(rewards.is_cuda(), "rewards must be CUDA");
    TORCH_CHECK(rewards.numel() > 0, "rewards must not be empty");
    c10::cuda::CUDAGuard dg(rewards.device());
    torch::Tensor r = rewards.reshape({-1});
    if (r.scalar_type() != torch::kFloat64) r = r.to(torch::kFloat64);
    r = r.contiguous();
    if (!state_.defined() || state_.device() != r.device()) {
        auto s = torch::zeros({2}, r.options());
        if (state_.defined()) s.copy_(state_);
        state_ = s;
    }
    if (!initialized) state_.zero_();        // the host flag is the reference's contract; keep the device copy in step
    const double* sc = nullptr;
    torch::Tensor scc;
    if (sum_count) {
        TORCH_CHECK(sum_count->is_cuda() && sum_count->scalar_type() == torch::kFloat64 && sum_count->numel() == 2,
                    "sum_count must be float64 CUDA [2]");
        scc = sum_count->contiguous();
        sc = scc.data_ptr<double>();
    }
    torch::Tensor a;
    if (adv) a = torch::empty_like(r);
    const int rc = icr_ema_update(r.data_ptr<double>(), (int)r.numel(), sc, momentum, state_.data_ptr<double>(),
                                  adv ? a.data_ptr<double>() : nullptr, current_stream());
    if (rc != ICR_OK) throw_runtime();
    if (adv) *adv = a.view(rewards.sizes());
    initialized = true;
    // a fresh tensor per update, as the reference's update() produces (include/ema_baseline.hpp:15-19): a caller that keeps
    // last step's baseline must not see it change (the mean is computed in float64 whatever the rewards' dtype)
    value = state_[0].clone();
    return value;
}

torch::Tensor EMABaseline::update(const torch::Tensor& rewards) { return run(rewards, nullptr, nullptr); }

torch::Tensor EMABaseline::update_global(const torch::Tensor& rewards_local, const torch::Tensor& sum_count) {
    return run(rewards_local, &sum_count, nullptr);
}

torch::Tensor EMABaseline::update_global(const torch::Tensor& rewards_local, void* nccl_comm, torch::Tensor* gathered_out) {
    torch::NoGradGuard ng;
    TORCH_CHECK(rewards_local.is_cuda(), "rewards must be CUDA");
    TORCH_CHECK(rewards_local.numel() > 0, "rewards must not be empty");
    TORCH_CHECK(nccl_comm != nullptr, "nccl_comm is null");
    c10::cuda::CUDAGuard dg(rewards_local.device());
    torch::Tensor r = rewards_local.reshape({-1});
    if (r.scalar_type() != torch::kFloat64) r = r.to(torch::kFloat64);
    r = r.contiguous();
    int ranks = 0;
    TORCH_CHECK(icr_nccl_comm_ranks(nccl_comm, &ranks) == ICR_OK, icr_last_error());
    auto gathered = torch::empty({(int64_t)ranks * r.numel()}, r.options());
    auto sum_count = torch::empty({2}, r.options());
    const int rc = icr_exchange_rewards(nccl_comm, r.data_ptr<double>(), (int)r.numel(), gathered.data_ptr<double>(),
                                        sum_count.data_ptr<double>(), current_stream());
    TORCH_CHECK(rc == ICR_OK, icr_last_error());
    if (gathered_out) *gathered_out = gathered;
    return run(gathered, nullptr, nullptr);      // the sum over the gathered vector in image order: the single-GPU value bit for bit
}

std::pair<torch::Tensor, torch::Tensor> EMABaseline::update_with_advantages(const torch::Tensor& rewards) {
    torch::Tensor adv;
    auto v = run(rewards, nullptr, &adv);
    return {v, adv};
}

// ------------------------------------------------------------------------------------------------ C hooks for the tests
// Raw-pointer forwarders with the same argument lists as oracle/ref_shim.cpp, so one ctypes test drives this
// library and the reference's own sources the same way.  They exist for tests/ only; product callers use the C++
// functions above.
namespace {
thread_local std::string g_hook_err;
}  // namespace
namespace icrh { void set_hook_error(const std::string& e) { g_hook_err = e; } }
using icrh::guarded;

extern "C" {

const char* icrh_last_error() { return g_hook_err.c_str(); }

int icrh_compute_rewards(const float* images, int images_on_device, const void* labels_dev, int labels_i64, const double* sizes_dev,
                         int B, int H, int W, int device, int min_pixels, int L_min, float beta, float token, float gamma,
                         double overhead_base, int adaptive, double lambda, double* rewards_out_dev) {
    return guarded([&] {
        c10::cuda::CUDAGuard g(device);
        auto dev = torch::Device(torch::kCUDA, device);
        auto img = torch::from_blob((void*)images, {B, 3, H, W},
                                    torch::TensorOptions().dtype(torch::kFloat32).device(images_on_device ? dev : torch::Device(torch::kCPU)));
        auto lab = torch::from_blob((void*)labels_dev, {B, H, W},
                                    torch::TensorOptions().dtype(labels_i64 ? torch::kInt64 : torch::kInt32).device(dev));
        auto sz = torch::from_blob((void*)sizes_dev, {B}, torch::TensorOptions().dtype(torch::kFloat64).device(dev));
        auto r = compute_rewards_batched(img, lab, sz, min_pixels, L_min, beta, token, gamma, overhead_base, adaptive != 0, lambda);
        torch::from_blob((void*)rewards_out_dev, {B}, torch::TensorOptions().dtype(torch::kFloat64).device(dev)).copy_(r);
        torch::cuda::synchronize(device);
    });
}

int icrh_compute_rewards_multi(const float* images_dev, const void* labels_dev, int labels_i64, const double* sizes_dev, int B, int S,
                               int H, int W, int device, double* rewards_out_dev) {
    return guarded([&] {
        c10::cuda::CUDAGuard g(device);
        auto dev = torch::Device(torch::kCUDA, device);
        auto img = torch::from_blob((void*)images_dev, {B, 3, H, W}, torch::TensorOptions().dtype(torch::kFloat32).device(dev));
        auto lab = torch::from_blob((void*)labels_dev, {B, S, H, W},
                                    torch::TensorOptions().dtype(labels_i64 ? torch::kInt64 : torch::kInt32).device(dev));
        auto sz = torch::from_blob((void*)sizes_dev, {B}, torch::TensorOptions().dtype(torch::kFloat64).device(dev));
        auto r = compute_rewards_multi(img, lab, sz);
        torch::from_blob((void*)rewards_out_dev, {B, S}, torch::TensorOptions().dtype(torch::kFloat64).device(dev)).copy_(r);
        torch::cuda::synchronize(device);
    });
}

int icrh_segment_stats(const int64_t* labels_dev, int H, int W, int K, int device, int32_t* counts_dev, int32_t* bboxes_dev) {
    return guarded([&] {
        c10::cuda::CUDAGuard g(device);
        auto dev = torch::Device(torch::kCUDA, device);
        auto lab = torch::from_blob((void*)labels_dev, {H, W}, torch::TensorOptions().dtype(torch::kInt64).device(dev));
        auto cnt = torch::from_blob((void*)counts_dev, {K}, torch::TensorOptions().dtype(torch::kInt32).device(dev));
        auto bb = torch::from_blob((void*)bboxes_dev, {K, 4}, torch::TensorOptions().dtype(torch::kInt32).device(dev));
        compute_counts_bboxes_from_compact_labels_cuda(lab, cnt, bb);
        torch::cuda::synchronize(device);
    });
}

int icrh_estimate_segment(const uint8_t* img_dev, int W, int H, int channels, const int64_t* labels_dev, const int32_t* counts_dev,
                          int64_t k, int x0, int y0, int w, int h, int device, int min_pixels, int L_min, float beta, float token,
                          float gamma, double overhead_base, int adaptive, double* out_dev) {
    return guarded([&] {
        c10::cuda::CUDAGuard g(device);
        estimate_png_size_masked_segment_to_output(img_dev, W, H, channels, labels_dev, counts_dev, k, x0, y0, w, h, min_pixels, L_min,
                                                   beta, token, gamma, overhead_base, adaptive != 0, out_dev);
        torch::cuda::synchronize(device);
    });
}

// rewards_dev f64 [n]; state_io host {value, initialized}; advantages_dev may be NULL.  One EMABaseline step.
int icrh_ema_step(const double* rewards_dev, int n, int device, double momentum, double* state_io, double* advantages_dev) {
    return guarded([&] {
        c10::cuda::CUDAGuard g(device);
        auto dev = torch::Device(torch::kCUDA, device);
        auto f64 = torch::TensorOptions().dtype(torch::kFloat64).device(dev);
        auto r = torch::from_blob((void*)rewards_dev, {n}, f64);
        EMABaseline b(momentum);
        if (state_io[1] != 0.0) { b.update(torch::full({1}, state_io[0], f64)); }     // seed: first call copies the mean
        torch::Tensor v;
        if (advantages_dev) {
            auto pr = b.update_with_advantages(r);
            v = pr.first;
            torch::from_blob((void*)advantages_dev, {n}, f64).copy_(pr.second);
        } else {
            v = b.update(r);
        }
        state_io[0] = v.item<double>();
        state_io[1] = b.initialized ? 1.0 : 0.0;
    });
}

// two updates in a row; returns in values_out the baseline tensor of the FIRST update read after the second one, and the
// second one's: a caller that keeps last step's baseline must still see last step's number (include/ema_baseline.hpp:15-19)
int icrh_ema_two_steps(const double* r1_dev, const double* r2_dev, int n, int device, double* values_out) {
    return guarded([&] {
        c10::cuda::CUDAGuard g(device);
        auto f64 = torch::TensorOptions().dtype(torch::kFloat64).device(torch::Device(torch::kCUDA, device));
        EMABaseline b(0.9);
        const torch::Tensor v1 = b.update(torch::from_blob((void*)r1_dev, {n}, f64));
        const torch::Tensor v2 = b.update(torch::from_blob((void*)r2_dev, {n}, f64));
        values_out[0] = v1.item<double>();
        values_out[1] = v2.item<double>();
    });
}

int icrh_ema_update_global_nccl(const double* rewards_local_dev, int n, int device, void* nccl_comm, int ranks, double* value_out,
                                double* gathered_out_dev) {
    return guarded([&] {
        c10::cuda::CUDAGuard g(device);
        auto opt = torch::TensorOptions().dtype(torch::kFloat64).device(torch::Device(torch::kCUDA, device));
        auto r = torch::from_blob((void*)rewards_local_dev, {n}, opt);
        EMABaseline ema;
        torch::Tensor gathered;
        auto v = ema.update_global(r, nccl_comm, &gathered);
        TORCH_CHECK(gathered.numel() == (int64_t)ranks * n, "gathered has the wrong size");
        torch::from_blob((void*)gathered_out_dev, {(int64_t)ranks * n}, opt).copy_(gathered);
        *value_out = v.item<double>();
    });
}

int icrh_labels_from_node_mappings(const int32_t* const* ptrs_dev, const int64_t* sizes, int B, int H, int W, int device, int32_t* out_dev) {
    return guarded([&] {
        c10::cuda::CUDAGuard g(device);
        auto opt = torch::TensorOptions().dtype(torch::kInt32).device(torch::Device(torch::kCUDA, device));
        std::vector<torch::Tensor> maps;
        for (int b = 0; b < B; ++b) maps.push_back(torch::from_blob((void*)ptrs_dev[b], {sizes[b]}, opt));
        auto out = labels_from_node_mappings(maps, H, W);
        TORCH_CHECK(out.dim() == 3 && out.size(0) == B && out.size(1) == H && out.size(2) == W, "wrong shape");
        torch::from_blob((void*)out_dev, {B, H, W}, opt).copy_(out);
        torch::cuda::synchronize(device);
    });
}

int icrh_labels_from_rama(const int32_t* node_labels_dev, int B, int H, int W, int device, int transpose_first) {
    return guarded([&] {
        auto dev = torch::Device(torch::kCUDA, device);
        auto t = torch::from_blob((void*)node_labels_dev, {B, H * W}, torch::TensorOptions().dtype(torch::kInt32).device(dev));
        auto v = labels_from_rama(transpose_first ? t.t() : t, H, W);
        TORCH_CHECK(v.data_ptr() == t.data_ptr(), "handoff must not copy");
    });
}

}  // extern "C"
