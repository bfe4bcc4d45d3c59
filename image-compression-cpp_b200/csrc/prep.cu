// prep: one CTA per label map.  Folds three reference steps into one sweep over the inputs:
//   Rule C  byte conversion f32 planar -> RGBA      (src/compute_rewards.cu:15-53)
//   Rule L  label compaction (sorted-unique rank)   (src/compute_rewards.cu:131-141, at::_unique)
//   Rule S  per-segment count + inclusive bbox      (src/segment_stats.cu:11-42)
// Rule L is a presence bitmap over (label - min) in shared memory plus a popcount prefix scan --
// no sort, no host sync on K.  Rule S is warp-aggregated (runs of equal ids along x collapse to
// one update) into a shared-memory table; bbox atomics are skipped when they cannot change the
// value.  Output is the interleaved {RGBA, compact id} plane the estimator reads with 64-bit loads.
#include "icr_internal.h"
#include <climits>

namespace icr {

constexpr int PREP_T = 512;

template <typename T> struct Vec4;
template <> struct Vec4<int32_t> { using type = int4; };
template <> struct Vec4<float> { using type = float4; };

__device__ __forceinline__ uint32_t to_byte(float v) {
    // clamp, round half to even, (second clamp of the reference is a no-op after the first)
    v = fminf(fmaxf(v, 0.f), 255.f);
    return (uint32_t)__float2int_rn(v);
}

template <typename LT, int VEC>
__device__ __forceinline__ void load_labels(const LT* lab, int64_t p, bool inrange, long long (&out)[VEC]) {
    if constexpr (VEC == 4 && sizeof(LT) == 4) {
        int4 q = inrange ? *reinterpret_cast<const int4*>(lab + p) : make_int4(0, 0, 0, 0);
        out[0] = q.x; out[1] = q.y; out[2] = q.z; out[3] = q.w;
    } else if constexpr (VEC == 4 && sizeof(LT) == 8) {
        longlong2 a = make_longlong2(0, 0), b = a;
        if (inrange) { a = *reinterpret_cast<const longlong2*>(lab + p); b = *reinterpret_cast<const longlong2*>(lab + p + 2); }
        out[0] = a.x; out[1] = a.y; out[2] = b.x; out[3] = b.y;
    } else {
        out[0] = inrange ? (long long)lab[p] : 0;
    }
}

struct StatsTab {        // either shared-memory tables (bbox stride 1) or the global output rows (stride 4)
    int32_t* cnt; int32_t* x0; int32_t* y0; int32_t* x1; int32_t* y1;
};

template <typename LT, int VEC, bool CONVERT>
__global__ void __launch_bounds__(PREP_T) prep_kernel(const float* __restrict__ images, const LT* __restrict__ labels,
                                                       int S, Tables t, int cap_words) {
    extern __shared__ uint32_t smem[];
    uint32_t* bitmap = smem;
    uint32_t* prefix = smem + cap_words;
    int32_t* stab = reinterpret_cast<int32_t*>(smem + 2 * cap_words);   // 5 * KS_SMEM
    __shared__ long long red_lo[PREP_T / 32], red_hi[PREP_T / 32];
    __shared__ uint32_t warp_tot[PREP_T / 32];
    __shared__ long long s_min;
    __shared__ int s_bad;

    const int v = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int H = t.H, W = t.W;
    const int64_t HW = (int64_t)H * W;
    const LT* lab = labels + (int64_t)v * HW;
    const float* img = CONVERT ? images + (int64_t)(v / S) * 3 * HW : nullptr;

    // ---- pass 0: min / max label of this map
    long long lo = LLONG_MAX, hi = LLONG_MIN;
    for (int64_t base = 0; base < HW; base += (int64_t)PREP_T * VEC) {
        int64_t p = base + (int64_t)tid * VEC;
        long long l[VEC];
        bool in = p < HW;
        load_labels<LT, VEC>(lab, p, in, l);
        if (in) {
#pragma unroll
            for (int i = 0; i < VEC; ++i) { lo = min(lo, l[i]); hi = max(hi, l[i]); }
        }
    }
    for (int o = 16; o; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0) { red_lo[wid] = lo; red_hi[wid] = hi; }
    for (int i = tid; i < cap_words; i += PREP_T) bitmap[i] = 0;
    __syncthreads();
    if (tid == 0) {
        for (int i = 1; i < PREP_T / 32; ++i) { lo = min(lo, red_lo[i]); hi = max(hi, red_hi[i]); }
        s_min = lo;
        // spread must fit the bitmap (unsigned compare also covers overflow of hi-lo)
        s_bad = ((unsigned long long)hi - (unsigned long long)lo >= (unsigned long long)cap_words * 32ull) ? ST_LABEL_RANGE : ST_OK;
    }
    __syncthreads();
    const long long lmin = s_min;
    if (s_bad) {
        if (tid == 0) {
            t.K[v] = 0; t.status[v] = s_bad;
            atomicExch(&t.err[0], 1); atomicMax(&t.err[1], 0x7fffffff - v);
        }
        return;
    }

    // ---- pass 1: presence bitmap (plain read first: almost every bit is already set)
    for (int64_t base = 0; base < HW; base += (int64_t)PREP_T * VEC) {
        int64_t p = base + (int64_t)tid * VEC;
        long long l[VEC];
        bool in = p < HW;
        load_labels<LT, VEC>(lab, p, in, l);
        if (in) {
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                uint32_t idx = (uint32_t)(l[i] - lmin), bit = 1u << (idx & 31);
                if (i > 0 && l[i] == l[i - 1]) continue;
                if (!(bitmap[idx >> 5] & bit)) atomicOr(&bitmap[idx >> 5], bit);
            }
        }
    }
    __syncthreads();

    // ---- exclusive popcount scan over bitmap words: prefix[w] = rank of the first bit of word w
    const int per = (cap_words + PREP_T - 1) / PREP_T;
    const int w0 = min(tid * per, cap_words), w1 = min(w0 + per, cap_words);
    uint32_t mine = 0;
    for (int w = w0; w < w1; ++w) mine += __popc(bitmap[w]);
    uint32_t inc = mine;
    for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += n; }
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    uint32_t woff = 0, total = 0;
    for (int i = 0; i < PREP_T / 32; ++i) { uint32_t x = warp_tot[i]; if (i < wid) woff += x; total += x; }
    uint32_t run = woff + inc - mine;
    for (int w = w0; w < w1; ++w) { prefix[w] = run; run += __popc(bitmap[w]); }
    const int K = (int)total;
    if (tid == 0) t.K[v] = K;
    if (K > t.Kmax) {
        if (tid == 0) { t.status[v] = ST_K_OVERFLOW; atomicExch(&t.err[0], 1); atomicMax(&t.err[1], 0x7fffffff - v); }
        return;
    }
    if (tid == 0) t.status[v] = ST_OK;

    // ---- stats tables: shared memory when K fits, else straight into the global rows
    int32_t* g_cnt = t.counts + (int64_t)v * t.Kmax;
    int32_t* g_bb = reinterpret_cast<int32_t*>(t.bboxes + (int64_t)v * t.Kmax);
    const bool in_smem = K <= KS_SMEM;
    StatsTab tab;
    if (in_smem) tab = StatsTab{stab, stab + KS_SMEM, stab + 2 * KS_SMEM, stab + 3 * KS_SMEM, stab + 4 * KS_SMEM};
    else tab = StatsTab{g_cnt, g_bb, g_bb + 1, g_bb + 2, g_bb + 3};
    if (in_smem) {
        for (int k = tid; k < K; k += PREP_T) { tab.cnt[k] = 0; tab.x0[k] = W; tab.y0[k] = H; tab.x1[k] = -1; tab.y1[k] = -1; }
    } else {
        for (int k = tid; k < K; k += PREP_T) { g_cnt[k] = 0; t.bboxes[(int64_t)v * t.Kmax + k] = make_int4(W, H, -1, -1); }
    }
    __syncthreads();
    const int bst = in_smem ? 1 : 4;    // bbox element stride

    // ---- pass 2: ids, bytes, stats
    uint2* out = t.pixlab + (int64_t)v * HW;
    for (int64_t base = 0; base < HW; base += (int64_t)PREP_T * VEC) {
        int64_t p = base + (int64_t)tid * VEC;
        const bool in = p < HW;
        long long l[VEC];
        load_labels<LT, VEC>(lab, p, in, l);
        uint32_t id[VEC], px[VEC];
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            uint32_t idx = in ? (uint32_t)(l[i] - lmin) : 0u;
            id[i] = prefix[idx >> 5] + __popc(bitmap[idx >> 5] & ((1u << (idx & 31)) - 1u));
            px[i] = 0;
        }
        if (CONVERT && in) {
            if constexpr (VEC == 4) {
                float4 r = *reinterpret_cast<const float4*>(img + p);
                float4 g = *reinterpret_cast<const float4*>(img + HW + p);
                float4 bl = *reinterpret_cast<const float4*>(img + 2 * HW + p);
                px[0] = to_byte(r.x) | to_byte(g.x) << 8 | to_byte(bl.x) << 16 | 0xff000000u;
                px[1] = to_byte(r.y) | to_byte(g.y) << 8 | to_byte(bl.y) << 16 | 0xff000000u;
                px[2] = to_byte(r.z) | to_byte(g.z) << 8 | to_byte(bl.z) << 16 | 0xff000000u;
                px[3] = to_byte(r.w) | to_byte(g.w) << 8 | to_byte(bl.w) << 16 | 0xff000000u;
            } else {
                px[0] = to_byte(img[p]) | to_byte(img[HW + p]) << 8 | to_byte(img[2 * HW + p]) << 16 | 0xff000000u;
            }
        }
        if (in) {
            if constexpr (VEC == 4) {
                uint4* o4 = reinterpret_cast<uint4*>(out + p);
                o4[0] = make_uint4(px[0], id[0], px[1], id[1]);
                o4[1] = make_uint4(px[2], id[2], px[3], id[3]);
            } else {
                out[p] = make_uint2(px[0], id[0]);
            }
        }
        // warp-aggregated stats: a lane whose VEC pixels share one id is a unit; consecutive
        // units with the same id on the same row form a run, updated once by its head lane
        const int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
        bool uni = in;
#pragma unroll
        for (int i = 1; i < VEC; ++i) uni = uni && id[i] == id[0];
        const uint32_t pid = __shfl_up_sync(0xffffffffu, id[0], 1);
        const int py = __shfl_up_sync(0xffffffffu, y, 1);
        const bool puni = __shfl_up_sync(0xffffffffu, uni, 1);
        const bool joins = lane > 0 && uni && puni && pid == id[0] && py == y;
        const uint32_t breaks = __ballot_sync(0xffffffffu, !joins);
        if (uni && !joins) {
            uint32_t above = lane == 31 ? 0u : (breaks & (0xffffffffu << (lane + 1)));
            int next = above ? __ffs(above) - 1 : 32;
            int n = (next - lane) * VEC;
            int k = (int)id[0];
            atomicAdd(&tab.cnt[k], n);
            if (x < tab.x0[k * bst]) atomicMin(&tab.x0[k * bst], x);
            if (y < tab.y0[k * bst]) atomicMin(&tab.y0[k * bst], y);
            if (x + n - 1 > tab.x1[k * bst]) atomicMax(&tab.x1[k * bst], x + n - 1);
            if (y > tab.y1[k * bst]) atomicMax(&tab.y1[k * bst], y);
        } else if (in && !uni) {
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                int k = (int)id[i];
                atomicAdd(&tab.cnt[k], 1);
                if (x + i < tab.x0[k * bst]) atomicMin(&tab.x0[k * bst], x + i);
                if (y < tab.y0[k * bst]) atomicMin(&tab.y0[k * bst], y);
                if (x + i > tab.x1[k * bst]) atomicMax(&tab.x1[k * bst], x + i);
                if (y > tab.y1[k * bst]) atomicMax(&tab.y1[k * bst], y);
            }
        }
    }
    if (in_smem) {
        __syncthreads();
        for (int k = tid; k < K; k += PREP_T) {
            g_cnt[k] = tab.cnt[k];
            t.bboxes[(int64_t)v * t.Kmax + k] = make_int4(tab.x0[k], tab.y0[k], tab.x1[k], tab.y1[k]);
        }
    }
}

template <typename LT, int VEC>
static cudaError_t launch_t(const float* images, const void* labels, int S, const Tables& t, bool convert, cudaStream_t st) {
    const int cap_words = (int)(label_capacity(t.H, t.W) / 32);
    const size_t smem = (size_t)cap_words * 8 + (size_t)KS_SMEM * 5 * 4;
    auto k1 = prep_kernel<LT, VEC, true>;
    auto k0 = prep_kernel<LT, VEC, false>;
    auto k = convert ? k1 : k0;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k<<<t.V, PREP_T, smem, st>>>(images, static_cast<const LT*>(labels), S, t, cap_words);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_prep(const float* images, const void* labels, int label_dtype, int B, int S,
                        const Tables& t, bool convert, cudaStream_t st) {
    (void)B;
    const int64_t HW = (int64_t)t.H * t.W;
    const bool i64 = label_dtype == ICR_LABELS_I64;
    bool vec = (t.W % 4 == 0) && ((uintptr_t)labels % 16 == 0) && (!convert || (uintptr_t)images % 16 == 0) && HW % 4 == 0;
    if (i64) return vec ? launch_t<int64_t, 4>(images, labels, S, t, convert, st) : launch_t<int64_t, 1>(images, labels, S, t, convert, st);
    return vec ? launch_t<int32_t, 4>(images, labels, S, t, convert, st) : launch_t<int32_t, 1>(images, labels, S, t, convert, st);
}

}  // namespace icr
