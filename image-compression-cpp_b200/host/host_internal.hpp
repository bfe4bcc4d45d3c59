// Shared by the translation units of libicr_host.so: the grow-only workspace cache and small helpers.
#pragma once
#include <torch/torch.h>
#include <c10/cuda/CUDAStream.h>

#include <stdexcept>
#include <string>

#include "../../include/icr_b200.h"

namespace icrh {

inline icr_stream current_stream() { return (icr_stream)c10::cuda::getCurrentCUDAStream().stream(); }
[[noreturn]] inline void throw_runtime() { throw std::runtime_error(icr_last_error()); }
inline int label_code(const torch::Tensor& t) { return t.scalar_type() == torch::kInt64 ? ICR_LABELS_I64 : ICR_LABELS_I32; }

// one per (host thread, device, H, W): the role of the reference's thread_local PngEstimatorWorkspace table
// (src/png_size_estimator.cu:466-469)
struct WsHolder {
    icr_workspace* ws = nullptr;
    int V = 0, Kmax = 0;
    ~WsHolder() { if (ws) icr_workspace_destroy(ws); }
};
WsHolder& workspace_for(int device, int V, int H, int W, int Kmax);       // reward_host.cpp
constexpr int KMAX_FIRST = 2048;   // table rows per label map to start with; grown on demand (plan_retry)
torch::Tensor rank_labels(const torch::Tensor& labs_vhw);                                      // reward_host.cpp
bool plan_retry(int status_bits, int max_segments, int64_t HW, int& Kmax, bool& rank);        // reward_host.cpp

// extern "C" test hooks report exceptions through these (reward_host.cpp)
void set_hook_error(const std::string& s);
template <class F> int guarded(F&& f) {
    try { f(); return 0; }
    catch (const c10::Error& e) { set_hook_error(e.what_without_backtrace()); return 1; }
    catch (const std::exception& e) { set_hook_error(e.what()); return 2; }
    catch (...) { set_hook_error("unknown exception"); return 3; }
}

}  // namespace icrh
