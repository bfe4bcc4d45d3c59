// Small kernels around the estimator: slice materialisation of one segment (Rule P),
// the compact-label stats drop-in (Rule S), EMA baseline + advantage (Rule E) and the self-test.
#include "icr_internal.h"
#include "icr_pixel.cuh"
#include "icr_reward.cuh"

namespace icr {

// ---- Rule P (src/image_slicer.cpp:59-66): masked bbox crop of one segment, u8 [h,w,4]
template <typename LT>
__global__ void slice_kernel(Tables t, Inputs in, int v, int k, uint32_t* out, size_t cap_px) {
    const int4 bb = t.bboxes[(int64_t)v * t.Kmax + k];
    const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
    if (w <= 0 || h <= 0) return;
    const int64_t HW = (int64_t)t.H * t.W;
    const float* img = in.images + (int64_t)(v / in.S) * 3 * HW;
    const LT* lab = static_cast<const LT*>(in.labels) + (int64_t)v * HW;
    const LT key = (LT)t.keys[(int64_t)v * t.Kmax + k];
    const int64_t n = (int64_t)w * h;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n && (size_t)i < cap_px; i += (int64_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / w), x = (int)(i - (int64_t)y * w);
        const int64_t g = (int64_t)(bb.y + y) * t.W + bb.x + x;
        uint32_t px = 0u;
        if (lab[g] == key) {
            auto tb = [](float f) { return (uint32_t)__float2int_rn(fminf(fmaxf(f, 0.f), 255.f)); };
            px = tb(img[g]) | tb(img[HW + g]) << 8 | tb(img[2 * HW + g]) << 16 | 0xff000000u;
        }
        out[i] = px;
    }
}

cudaError_t launch_slice(const Tables& t, const Inputs& in, int v, int k, uint8_t* out, size_t cap, cudaStream_t st) {
    if (in.label_dtype == ICR_LABELS_I64) slice_kernel<int64_t><<<sm_count(), 256, 0, st>>>(t, in, v, k, reinterpret_cast<uint32_t*>(out), cap / 4);
    else slice_kernel<int32_t><<<sm_count(), 256, 0, st>>>(t, in, v, k, reinterpret_cast<uint32_t*>(out), cap / 4);
    count_launch();
    return cudaGetLastError();
}

// ---- Rule S on already-compact int64 labels (src/segment_stats.cu:11-77)
__global__ void compact_init_kernel(int32_t* counts, int32_t* bboxes, int K, int W, int H) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    counts[k] = 0;
    reinterpret_cast<int4*>(bboxes)[k] = make_int4(W, H, -1, -1);
}

__global__ void __launch_bounds__(256) compact_stats_kernel(const int64_t* __restrict__ lab, int H, int W, int K,
                                                             int32_t* counts, int32_t* bboxes) {
    // one thread per pixel, runs of equal ids along a row collapse to one update (warp-aggregated)
    const int64_t HW = (int64_t)H * W;
    const int lane = threadIdx.x & 31;
    for (int64_t base = (int64_t)blockIdx.x * blockDim.x; base < HW; base += (int64_t)gridDim.x * blockDim.x) {
        const int64_t p = base + threadIdx.x;
        const bool in = p < HW;
        const int64_t id = in ? lab[p] : -1;
        const int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
        const int64_t pid = __shfl_up_sync(0xffffffffu, id, 1);
        const int py = __shfl_up_sync(0xffffffffu, y, 1);
        const bool ok = in && id >= 0 && id < K;
        const bool joins = lane > 0 && ok && pid == id && py == y;
        const uint32_t breaks = __ballot_sync(0xffffffffu, !joins);
        if (ok && !joins) {
            uint32_t above = lane == 31 ? 0u : (breaks & (0xffffffffu << (lane + 1)));
            const int n = (above ? __ffs(above) - 1 : 32) - lane;
            int32_t* bb = bboxes + 4 * id;
            atomicAdd(&counts[id], n);
            atomicMin(&bb[0], x); atomicMin(&bb[1], y); atomicMax(&bb[2], x + n - 1); atomicMax(&bb[3], y);
        }
    }
}

cudaError_t launch_compact_stats(const int64_t* labels, int H, int W, int K, int32_t* counts, int32_t* bboxes, cudaStream_t st) {
    if (K > 0) { compact_init_kernel<<<(K + 255) / 256, 256, 0, st>>>(counts, bboxes, K, W, H); count_launch(); }
    const int64_t HW = (int64_t)H * W;
    int blocks = (int)((HW + 255) / 256);
    if (blocks > sm_count() * 16) blocks = sm_count() * 16;
    if (blocks > 0) { compact_stats_kernel<<<blocks, 256, 0, st>>>(labels, H, W, K, counts, bboxes); count_launch(); }
    return cudaGetLastError();
}

// ---- Rule E + advantage (include/ema_baseline.hpp:11-21, src/fcn/training.cpp:178-181)
__device__ double block_sum(double v, double* sh) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = 0.0;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += sh[i];
    return s;
}

__global__ void __launch_bounds__(1024) sum_count_kernel(const double* __restrict__ r, int n, double* out) {
    __shared__ double sh[32];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += r[i];
    s = block_sum(s, sh);
    if (threadIdx.x == 0) { out[0] = s; out[1] = (double)n; }
}

// flags != NULL: r is this rank's gathered vector, filled by all ranks' reward kernels through peer memory; wait until
// every rank has signalled step `step` (bounded: a rank that never arrives poisons the baseline instead of hanging the GPU)
// r is read past L1 and without the read-only path: on the gathered path it is being written by peer GPUs while this kernel
// already runs, and only the acquire on the flags orders those writes before the reads below
__global__ void __launch_bounds__(1024) ema_kernel(const double* r, int n, const double* sum_count,
                                                    double momentum, double* state, double* adv,
                                                    const unsigned long long* flags, int ranks, unsigned long long step) {
    __shared__ double sh[32];
    __shared__ int s_timeout;
    if (flags) {
        if (threadIdx.x == 0) s_timeout = 0;
        __syncthreads();
        if ((int)threadIdx.x < ranks) {
            const long long t0 = clock64();
            unsigned long long seen;
            do {
                asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(flags + threadIdx.x) : "memory");
            } while (seen < step && clock64() - t0 < 20000000000ll);         // ~10 s
            if (seen < step) s_timeout = 1;
        }
        __syncthreads();
        if (s_timeout) { if (threadIdx.x == 0) { state[0] = __longlong_as_double(0x7ff8000000000000ll); state[1] = 1.0; } return; }
    }
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += __ldcg(r + i);
    s = block_sum(s, sh);
    const double local_mean = n > 0 ? s / (double)n : 0.0;
    const double mean = sum_count ? sum_count[0] / sum_count[1] : local_mean;
    __shared__ double s_value;
    if (threadIdx.x == 0) {
        double value = state[1] != 0.0 ? state[0] * momentum + mean * (1.0 - momentum) : mean;
        state[0] = value; state[1] = 1.0;
        s_value = value;
    }
    if (!adv) return;
    // adv = (r - value); standardise: the baseline cancels in (adv - mean(adv)), std is the population std
    double q = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { double d = __ldcg(r + i) - local_mean; q += d * d; }
    q = block_sum(q, sh);
    const double sd = fmax(sqrt(q / (double)n), 1e-6);
    const double value = s_value;
    const double adv_mean = local_mean - value;
    for (int i = threadIdx.x; i < n; i += blockDim.x) adv[i] = ((__ldcg(r + i) - value) - adv_mean) / sd;
}

cudaError_t launch_sum_count(const double* rewards, int n, double* sum_count, cudaStream_t st) {
    sum_count_kernel<<<1, 1024, 0, st>>>(rewards, n, sum_count);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_ema(const double* rewards, int n, const double* sum_count, double momentum, double* state, double* adv,
                       const unsigned long long* flags, int ranks, unsigned long long step, cudaStream_t st) {
    ema_kernel<<<1, 1024, 0, st>>>(rewards, n, sum_count, momentum, state, adv, flags, ranks, step);
    count_launch();
    return cudaGetLastError();
}

// ---- exhaustive self-test of the packed-byte arithmetic against scalar code
__global__ void selftest_kernel(unsigned int* bad) {
    // thread = (a, b) pair; loops over c: Paeth; also Avg, absdiff/cost and vsub4 on the pair
    const uint32_t ab = blockIdx.x * blockDim.x + threadIdx.x;     // 0..65535
    const uint32_t a = ab & 255u, b = ab >> 8;
    unsigned int errs = 0;
    for (uint32_t c = 0; c < 256; c += 4) {
        // four different c values in the four byte lanes, a and b rotated so lanes differ
        uint32_t A = a | ((a ^ 0x55u) << 8) | (((a + 77u) & 255u) << 16) | ((255u - a) << 24);
        uint32_t Bv = b | (((b + 131u) & 255u) << 8) | ((b ^ 0xAAu) << 16) | ((255u - b) << 24);
        uint32_t Cv = c | ((c + 1) << 8) | ((c + 2) << 16) | ((c + 3) << 24);
        uint32_t P = paeth4(A, Bv, Cv), G = avg4(A, Bv), D = absdiff4(A, Bv), R = __vsub4(A, Bv);
        uint32_t cost = cost_acc(D, 7u), cost_ref = 7u;
        for (int l = 0; l < 4; ++l) {
            int aa = (A >> (8 * l)) & 255, bb = (Bv >> (8 * l)) & 255, cc = (Cv >> (8 * l)) & 255;
            if ((int)((P >> (8 * l)) & 255u) != paeth1(aa, bb, cc)) ++errs;
            if ((int)((G >> (8 * l)) & 255u) != ((aa + bb) >> 1)) ++errs;
            if ((int)((D >> (8 * l)) & 255u) != abs(aa - bb)) ++errs;
            if ((int)((R >> (8 * l)) & 255u) != ((aa - bb) & 255)) ++errs;
            cost_ref += (uint32_t)abs((int)(int8_t)(uint8_t)((aa - bb) & 255));
        }
        if (cost != cost_ref) ++errs;
    }
    if (errs) atomicAdd(bad, errs);
}

cudaError_t run_selftest(int* mismatches) {
    unsigned int* d = nullptr;
    cudaError_t e = cudaMalloc(&d, sizeof(unsigned int));
    if (e != cudaSuccess) return e;
    cudaMemset(d, 0, sizeof(unsigned int));
    selftest_kernel<<<256, 256>>>(d);
    count_launch();
    unsigned int h = 0;
    e = cudaMemcpy(&h, d, sizeof h, cudaMemcpyDeviceToHost);
    cudaFree(d);
    *mismatches = (int)h;
    return e;
}

}  // namespace icr
