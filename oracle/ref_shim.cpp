// TEST INFRASTRUCTURE ONLY. C-ABI shim over the UNMODIFIED reference sources
// (/root/reference/src/{segment_stats,png_size_estimator,compute_rewards}.cu), which
// build_ref.sh compiles where they lie into oracle/_ref/libref_reward.so. No reference
// source is copied into this repo; this file only forwards raw device pointers to the
// reference's three public entry points (include/compute_rewards.cuh:5-17,
// include/segment_stats.cuh:6-10, include/png_size_estimator.cuh:61-71).
#include <torch/torch.h>
#include <c10/cuda/CUDAGuard.h>
#include <string>
#include "compute_rewards.cuh"
#include "segment_stats.cuh"
#include "png_size_estimator.cuh"

static thread_local std::string g_err;

template <class F> static int guarded(F&& f) {
    try { f(); return 0; }
    catch (const std::exception& e) { g_err = e.what(); return 1; }
    catch (...) { g_err = "unknown exception"; return 2; }
}

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

// images: device or host f32 [B,3,H,W]; labels: device int32/int64 [B,H,W]; sizes: device f64 [B]
int ref_compute_rewards(const float* images, int images_on_device,
                        const void* labels_dev, int labels_i64,
                        const double* sizes_dev, int B, int H, int W, int device,
                        int min_pixels, int L_min, float beta, float token, float gamma,
                        double overhead_base, int adaptive, double lambda,
                        double* rewards_out_dev) {
    return guarded([&] {
        c10::cuda::CUDAGuard g(device);
        auto dev = torch::Device(torch::kCUDA, device);
        auto img = torch::from_blob((void*)images, {B, 3, H, W},
            torch::TensorOptions().dtype(torch::kFloat32).device(images_on_device ? dev : torch::Device(torch::kCPU)));
        auto lab = torch::from_blob((void*)labels_dev, {B, H, W},
            torch::TensorOptions().dtype(labels_i64 ? torch::kInt64 : torch::kInt32).device(dev));
        auto sz = torch::from_blob((void*)sizes_dev, {B}, torch::TensorOptions().dtype(torch::kFloat64).device(dev));
        auto r = compute_rewards_batched(img, lab, sz, min_pixels, L_min, beta, token, gamma,
                                         overhead_base, adaptive != 0, lambda);
        auto out = torch::from_blob((void*)rewards_out_dev, {B}, torch::TensorOptions().dtype(torch::kFloat64).device(dev));
        out.copy_(r);
        torch::cuda::synchronize(device);
    });
}

int ref_segment_stats(const int64_t* labels_dev, int H, int W, int K, int device,
                      int32_t* counts_dev, int32_t* bboxes_dev) {
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

int ref_estimate_segment(const uint8_t* img_dev, int W, int H, int channels,
                         const int64_t* labels_dev, const int32_t* counts_dev, int64_t k,
                         int x0, int y0, int w, int h, int device,
                         int min_pixels, int L_min, float beta, float token, float gamma,
                         double overhead_base, int adaptive, double* out_dev) {
    return guarded([&] {
        c10::cuda::CUDAGuard g(device);
        estimate_png_size_masked_segment_to_output(img_dev, W, H, channels, labels_dev, counts_dev, k,
            x0, y0, w, h, min_pixels, L_min, beta, token, gamma, overhead_base, adaptive != 0, out_dev);
        torch::cuda::synchronize(device);
    });
}

}  // extern "C"
