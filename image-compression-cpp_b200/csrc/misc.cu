// Small kernels around the estimator: slice materialisation of one segment (Rule P),
// the compact-label stats drop-in (Rule S), EMA baseline + advantage (Rule E) and the self-test.
#include "icr_internal.h"
#include <algorithm>
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

// ---- what a data loader holds before its float conversion (include/fcn/EdgeDataset.h:73): u8 pixels, and label ids that
// fit 16 bits whenever H*W <= 65536.  The host-buffer entry can take those over the link (5 instead of 16 bytes per pixel)
// and widen them here into the f32 planes / int32 labels the path reads; (float)u8 is exact, so Rule C gives the byte back.
__global__ void __launch_bounds__(256) expand_planar_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, int64_t n, int vec) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (vec) {
        const uchar4* s4 = reinterpret_cast<const uchar4*>(src);
        float4* d4 = reinterpret_cast<float4*>(dst);
        for (int64_t i = t0; i < (n >> 2); i += stride) {
            const uchar4 u = __ldg(s4 + i);
            d4[i] = make_float4((float)u.x, (float)u.y, (float)u.z, (float)u.w);
        }
    } else {
        for (int64_t i = t0; i < n; i += stride) dst[i] = (float)__ldg(src + i);
    }
}

// interleaved [B,H,W,3] -> planar [B,3,H,W]; a thread per pixel
__global__ void __launch_bounds__(256) expand_hwc_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, int64_t B, int64_t HW) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * HW; i += stride) {
        const int64_t b = i / HW, p = i - b * HW;
        const uint8_t* q = src + 3 * i;
        float* d = dst + b * 3 * HW + p;
        d[0] = (float)__ldg(q); d[HW] = (float)__ldg(q + 1); d[2 * HW] = (float)__ldg(q + 2);
    }
}

__global__ void __launch_bounds__(256) expand_u16_kernel(const uint16_t* __restrict__ src, int32_t* __restrict__ dst, int64_t n, int vec) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (vec) {
        const ushort4* s4 = reinterpret_cast<const ushort4*>(src);
        int4* d4 = reinterpret_cast<int4*>(dst);
        for (int64_t i = t0; i < (n >> 2); i += stride) {
            const ushort4 u = __ldg(s4 + i);
            d4[i] = make_int4(u.x, u.y, u.z, u.w);
        }
    } else {
        for (int64_t i = t0; i < n; i += stride) dst[i] = (int32_t)__ldg(src + i);
    }
}

cudaError_t launch_expand(const uint8_t* images_u8, int hwc, float* images_f32, int64_t B, int64_t HW, const uint16_t* labels_u16,
                          int32_t* labels_i32, int64_t n_labels, cudaStream_t st) {
    const int grid = sm_count() * 8;
    if (images_u8) {
        const int64_t n = B * 3 * HW;
        if (hwc) expand_hwc_kernel<<<grid, 256, 0, st>>>(images_u8, images_f32, B, HW);
        else expand_planar_kernel<<<grid, 256, 0, st>>>(images_u8, images_f32, n,
                                                        (n & 3) == 0 && (reinterpret_cast<uintptr_t>(images_u8) & 3) == 0 &&
                                                            (reinterpret_cast<uintptr_t>(images_f32) & 15) == 0);
        count_launch();
    }
    if (labels_u16) {
        expand_u16_kernel<<<grid, 256, 0, st>>>(labels_u16, labels_i32, n_labels,
                                                (n_labels & 3) == 0 && (reinterpret_cast<uintptr_t>(labels_u16) & 7) == 0 &&
                                                    (reinterpret_cast<uintptr_t>(labels_i32) & 15) == 0);
        count_launch();
    }
    return cudaGetLastError();
}

// ---- the producer side of the label handoff (src/rama_wrapper.cu:102-124): the solver returns one node_mapping vector per
// sample, which the reference copies into out[b] with one thrust::copy (a launch) per sample.  Here up to PACK_MAX vectors
// travel as kernel parameters and one launch writes their rows of the [B,N] plane.
struct PackPtrs { const int32_t* p[PACK_MAX]; };

__global__ void __launch_bounds__(256) pack_labels_kernel(PackPtrs src, int64_t N, int32_t* __restrict__ out) {
    const int32_t* s = src.p[blockIdx.y];
    int32_t* d = out + (int64_t)blockIdx.y * N;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (((reinterpret_cast<uintptr_t>(s) | reinterpret_cast<uintptr_t>(d)) & 15) == 0) {
        const int64_t n4 = N >> 2;
        for (int64_t i = t0; i < n4; i += stride) reinterpret_cast<int4*>(d)[i] = __ldg(reinterpret_cast<const int4*>(s) + i);
        for (int64_t i = (n4 << 2) + t0; i < N; i += stride) d[i] = __ldg(s + i);
    } else {
        for (int64_t i = t0; i < N; i += stride) d[i] = __ldg(s + i);
    }
}

cudaError_t launch_pack_labels(const int32_t* const* ptrs_host, int B, int64_t N, int32_t* out, cudaStream_t st) {
    const int bx = (int)std::min<int64_t>((N + 4 * 256 - 1) / (4 * 256), 64);
    for (int b0 = 0; b0 < B; b0 += PACK_MAX) {
        const int nb = std::min(PACK_MAX, B - b0);
        PackPtrs pp;
        for (int i = 0; i < nb; ++i) pp.p[i] = ptrs_host[b0 + i];
        pack_labels_kernel<<<dim3(bx, nb), 256, 0, st>>>(pp, N, out + (int64_t)b0 * N);
        count_launch();
    }
    return cudaGetLastError();
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
