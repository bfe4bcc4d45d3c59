// Drop-in header for the reference's include/ema_baseline.hpp:4-22: same struct, same members, same update()
// contract (first call copies the batch mean, later calls blend with `momentum`), computed by one kernel of
// libicr_b200.so instead of a chain of ATen scalar ops.  `value` is a 0-dim float64 view of device state.
#pragma once
#include <torch/torch.h>

struct EMABaseline {
    double momentum = 0.99;
    bool initialized = false;
    torch::Tensor value;

    explicit EMABaseline(double m = 0.99) : momentum(m) {}

    // rewards: float64 CUDA, any shape.  Returns `value`.
    torch::Tensor update(const torch::Tensor& rewards);

    // Sharded batch: `sum_count` (float64 CUDA [2]) holds the all-reduced {sum of rewards, count}; every rank
    // then applies the identical update.
    torch::Tensor update_global(const torch::Tensor& rewards_local, const torch::Tensor& sum_count);

    // Sharded batch, exchange included: all-gathers every rank's rewards (same count on every rank) and all-reduces
    // {sum, count} with icr_exchange_rewards on the current stream -- `nccl_comm` is the caller's ncclComm_t -- then applies
    // Rule E to the gathered vector, which is bit-identical to update() on the whole batch on one GPU.  `gathered_out`
    // (optional) receives the rewards of all ranks in rank order.
    torch::Tensor update_global(const torch::Tensor& rewards_local, void* nccl_comm, torch::Tensor* gathered_out = nullptr);

    // update() + the advantage of src/fcn/training.cpp:178-181: (r - value) standardised by the batch mean and the
    // population std clamped at 1e-6.  Returns {value, advantages like rewards}.
    std::pair<torch::Tensor, torch::Tensor> update_with_advantages(const torch::Tensor& rewards);

private:
    torch::Tensor state_;   // float64 [2] on the rewards' device: {value, initialized flag}
    torch::Tensor run(const torch::Tensor& rewards, const torch::Tensor* sum_count, torch::Tensor* adv);
};
