// Slice materialisation for the whole batch and the decoder's paste.
// The reference makes K full-image passes per image on the CPU: (mask == label) then a bbox scan then a masked crop
// (src/image_slicer.cpp:48-74, one std::async per label at :98-100).  Here the counts and bboxes of the batch are
// already in the workspace tables, so one launch plans where every slice goes in one packed buffer and one launch
// writes all of them; the decoder side pastes slices where alpha > 0 (src/reassemble.cpp:75-107).
#include <algorithm>
#include "icr_internal.h"
#include "icr_pixel.cuh"

namespace icr {

constexpr int PLAN_T = 1024;
constexpr int EMIT_T = 64;

// ---- plan: offsets[f] = byte offset of the slice of flat segment f = seg_prefix[v] + k, in (v, k) order;
// summary = {segments, total bytes}.  Three small launches: bytes per map (a warp per map), a scan over the maps
// (one CTA), and the offsets inside each map (a warp per map).
__device__ __forceinline__ unsigned long long slice_bytes_of(const int4 bb) {
    const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
    return (w > 0 && h > 0) ? 4ull * (unsigned long long)w * (unsigned long long)h : 0ull;
}

__device__ __forceinline__ unsigned long long warp_inclusive(unsigned long long x, int lane) {
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long up = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += up;
    }
    return x;
}

__global__ void __launch_bounds__(256) plan_map_bytes_kernel(Tables t, unsigned long long* __restrict__ map_off) {
    const int v = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (v >= t.V) return;
    const int K = t.seg_prefix[v + 1] - t.seg_prefix[v];               // 0 for a map that failed
    unsigned long long a = 0ull;
    for (int k = lane; k < K; k += 32) a += slice_bytes_of(__ldg(t.bboxes + (int64_t)v * t.Kmax + k));
    for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if (lane == 0) map_off[v] = a;
}

__global__ void __launch_bounds__(PLAN_T) plan_scan_maps_kernel(Tables t, unsigned long long* __restrict__ map_off,
                                                                 unsigned long long* __restrict__ summary) {
    __shared__ unsigned long long s_warp[PLAN_T / 32];
    __shared__ unsigned long long s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0ull;
    __syncthreads();
    for (int base = 0; base < t.V; base += PLAN_T) {
        const int v = base + tid;
        const unsigned long long a = v < t.V ? map_off[v] : 0ull;
        const unsigned long long incl = warp_inclusive(a, lane);
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) s_warp[lane] = warp_inclusive(s_warp[lane], lane);
        __syncthreads();
        if (v < t.V) map_off[v] = s_carry + (warp ? s_warp[warp - 1] : 0ull) + incl - a;       // exclusive
        __syncthreads();
        if (tid == 0) s_carry += s_warp[PLAN_T / 32 - 1];
        __syncthreads();
    }
    if (tid == 0) {
        summary[0] = (unsigned long long)t.seg_prefix[t.V];
        summary[1] = s_carry;
    }
}

__global__ void __launch_bounds__(256) plan_offsets_kernel(Tables t, const unsigned long long* __restrict__ map_off,
                                                           unsigned long long* __restrict__ offsets, long long capacity) {
    const int v = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (v >= t.V) return;
    const int f0 = t.seg_prefix[v], K = t.seg_prefix[v + 1] - f0;
    unsigned long long run = map_off[v];
    if (v == 0 && lane == 0) offsets[0] = 0ull;
    for (int k0 = 0; k0 < K; k0 += 32) {
        const int k = k0 + lane;
        const unsigned long long a = k < K ? slice_bytes_of(__ldg(t.bboxes + (int64_t)v * t.Kmax + k)) : 0ull;
        const unsigned long long incl = warp_inclusive(a, lane);
        if (k < K && (long long)f0 + k + 1 < capacity) offsets[f0 + k + 1] = run + incl;
        run += __shfl_sync(0xffffffffu, incl, 31);
    }
}

// ---- emit (Rule C + Rule P): one CTA per segment, threads along the slice's pixels in row-major order: the slice is
// contiguous in the packed buffer, so stores are fully coalesced and the loads are row fragments of the three planes and
// the label map.  CTAs are dealt to (map, segment) pairs directly, so no CTA has to search the flat work list for its
// map: with at least as many maps as CTAs a CTA takes whole maps, otherwise the CTAs of one residue class share a map
// and stride over its segments.  The table row of a CTA's next segment is requested before the pixels of the current
// one, so a segment costs one memory round trip (its pixels) instead of a chain of look-ups followed by the pixels.
template <typename LT>
struct SliceRow { int4 bb; LT key; unsigned long long o; };

template <typename LT>
__device__ __forceinline__ SliceRow<LT> slice_row(const Tables& t, const unsigned long long* __restrict__ offsets, int v, int f0, int k) {
    SliceRow<LT> r;
    const int64_t row = (int64_t)v * t.Kmax + k;
    r.bb = __ldg(t.bboxes + row);
    r.key = (LT)__ldg(t.keys + row);
    r.o = __ldg(offsets + f0 + k) >> 2;
    return r;
}

template <typename LT, bool SWAP>
__global__ void __launch_bounds__(EMIT_T, 32) slice_emit_kernel(Tables t, Inputs in, const unsigned long long* __restrict__ offsets,
                                                             uint32_t* __restrict__ out, unsigned long long cap_px) {
    const int64_t HW = (int64_t)t.H * t.W;
    // the grid is P consecutive CTAs for each of the V maps (launch_slice_emit): CTA c strides over the segments of map c / P
    const int nsub = gridDim.x / t.V, sub = blockIdx.x % nsub;
    {
        const int v = blockIdx.x / nsub;
        const int f0 = __ldg(t.seg_prefix + v), K = __ldg(t.seg_prefix + v + 1) - f0;      // K = 0 for a map that failed
        if (K <= 0) return;
        // fewer segments than CTAs (coarse partitions, down to one whole-image segment): `parts` CTAs share a segment
        // and interleave its pixel blocks
        const int parts = K >= nsub ? 1 : nsub / K, kstep = K >= nsub ? nsub : K;
        const int part = sub / kstep, k0 = sub - part * kstep;
        if (part >= parts) return;
        const float* __restrict__ img_v = in.images + (int64_t)(v / in.S) * 3 * HW;
        const LT* __restrict__ lab_v = static_cast<const LT*>(in.labels) + (int64_t)v * HW;
        SliceRow<LT> nx = slice_row<LT>(t, offsets, v, f0, k0);
        for (int k = k0; k < K; k += kstep) {
            const SliceRow<LT> cur = nx;
            if (k + kstep < K) nx = slice_row<LT>(t, offsets, v, f0, k + kstep);
            const int4 bb = cur.bb;
            const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
            if (w <= 0 || h <= 0) continue;
            if (cur.o + (unsigned long long)w * h > cap_px) continue;      // the host entry has checked the capacity already
            const LT key = cur.key;
            const int base = bb.y * t.W + bb.x;                        // a map has fewer than 2^31 pixels (icr_workspace_create)
            const float* __restrict__ img = img_v + base;
            const LT* __restrict__ lab = lab_v + base;
            uint32_t* __restrict__ dst = out + cur.o;
            // lanes take the bbox pixels in row-major order, four independent pixels per lane in flight; all four loads of
            // a pixel are issued together (the planes are read for masked-out pixels too: one latency instead of two)
            const int n = w * h;                                       // the warp path's and the big kernel's bboxes are below 2^31 pixels
            const float rw = 1.0f / (float)w;
            constexpr int U = 4;
            for (int i0 = threadIdx.x + part * (EMIT_T * U); i0 < n; i0 += EMIT_T * U * parts) {
                float c0[U], c1[U], c2[U];
                LT l[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int i = i0 + u * EMIT_T;
                    if (i < n) {
                        int y = __float2int_rz(((float)i + 0.5f) * rw);        // i / w; corrected below where the float estimate is off by one
                        y += (y + 1) * w <= i;
                        y -= y * w > i;
                        const int g = y * t.W + (i - y * w);
                        l[u] = lab[g];
                        c0[u] = __ldg(img + g); c1[u] = __ldg(img + HW + g); c2[u] = __ldg(img + 2 * HW + g);
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int i = i0 + u * EMIT_T;
                    if (i < n) dst[i] = l[u] == key ? (SWAP ? rgba_of(c2[u], c1[u], c0[u]) : rgba_of(c0[u], c1[u], c2[u])) : 0u;
                }
            }
        }
    }
}

cudaError_t launch_slice_plan(const Tables& t, unsigned long long* offsets, long long capacity, unsigned long long* summary,
                              cudaStream_t st) {
    const int grid = (t.V + 7) / 8;
    plan_map_bytes_kernel<<<grid, 256, 0, st>>>(t, t.slice_map_off);
    plan_scan_maps_kernel<<<1, PLAN_T, 0, st>>>(t, t.slice_map_off, summary);
    plan_offsets_kernel<<<grid, 256, 0, st>>>(t, t.slice_map_off, offsets, capacity);
    count_launch(3);
    return cudaGetLastError();
}

cudaError_t launch_slice_emit(const Tables& t, const Inputs& in, const unsigned long long* offsets, uint8_t* out, size_t cap,
                              int swap, cudaStream_t st) {
    if (t.V <= 0) return cudaSuccess;
    const long long resident = (long long)sm_count() * 32;           // CTAs the machine holds at once (32 registers, 64 threads)
    const long long map_bytes = (long long)t.H * t.W * (12 + (in.label_dtype == ICR_LABELS_I64 ? 8 : 4));
    const long long in_flight = std::max(1ll, (48ll << 20) / std::max(1ll, map_bytes));      // maps that share the L2 comfortably
    long long P = std::min(1024ll, (resident + in_flight - 1) / in_flight);                  // CTAs per map
    while (P > 1 && P * t.V > (1ll << 30)) P >>= 1;
    const int grid = (int)(P * t.V);
    uint32_t* o = reinterpret_cast<uint32_t*>(out);
    const unsigned long long cap_px = cap / 4;
    if (in.label_dtype == ICR_LABELS_I64) {
        if (swap) slice_emit_kernel<int64_t, true><<<grid, EMIT_T, 0, st>>>(t, in, offsets, o, cap_px);
        else slice_emit_kernel<int64_t, false><<<grid, EMIT_T, 0, st>>>(t, in, offsets, o, cap_px);
    } else {
        if (swap) slice_emit_kernel<int32_t, true><<<grid, EMIT_T, 0, st>>>(t, in, offsets, o, cap_px);
        else slice_emit_kernel<int32_t, false><<<grid, EMIT_T, 0, st>>>(t, in, offsets, o, cap_px);
    }
    count_launch();
    return cudaGetLastError();
}

// ---- decoder: paste where alpha > 0 (src/reassemble.cpp:75-107).  All slices of one label map in one launch: the
// segments of a partition do not overlap, so the order of the pastes does not matter.
// The grid is P consecutive CTAs per map, as in the emit: `v0` is the first map and canvases lie HW*4 bytes apart, so
// one map (icr_reassemble_map) and the whole batch (icr_reassemble_batch) are the same launch.
__global__ void __launch_bounds__(EMIT_T) reassemble_kernel(Tables t, int v0, int P, const unsigned long long* __restrict__ offsets,
                                                             const uint32_t* __restrict__ slices, uint32_t* __restrict__ canvases) {
    const int v = v0 + blockIdx.x / P;
    uint32_t* __restrict__ canvas = canvases + (int64_t)(blockIdx.x / P) * t.H * t.W;
    const int K = t.status[v] == ST_OK ? t.K[v] : 0;
    const int f0 = t.seg_prefix[v];
    if (K <= 0) return;
    const int sub = blockIdx.x % P;
    const int parts = K >= P ? 1 : P / K, kstep = K >= P ? P : K;          // CTAs that share a slice when there are few, as in the emit
    const int part = sub / kstep;
    if (part >= parts) return;
    for (int k = sub - part * kstep; k < K; k += kstep) {
        const int4 bb = t.bboxes[(int64_t)v * t.Kmax + k];
        const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
        if (w <= 0 || h <= 0) continue;
        const uint32_t* __restrict__ src = slices + (offsets[f0 + k] >> 2);
        uint32_t* __restrict__ dst = canvas + (int64_t)bb.y * t.W + bb.x;
        // threads along the slice's pixels in row-major order (the slice is contiguous: coalesced loads, all lanes busy
        // whatever the width), four independent pixels per thread in flight
        const int n = w * h;
        const float rw = 1.0f / (float)w;
        constexpr int U = 4;
        for (int i0 = threadIdx.x + part * (EMIT_T * U); i0 < n; i0 += EMIT_T * U * parts) {
            uint32_t px[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + u * EMIT_T;
                px[u] = i < n ? __ldg(src + i) : 0u;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + u * EMIT_T;
                if (px[u] >> 24) {
                    int y = __float2int_rz(((float)i + 0.5f) * rw);        // i / w, corrected where the float estimate is off by one
                    y += (y + 1) * w <= i;
                    y -= y * w > i;
                    dst[y * t.W + (i - y * w)] = px[u];
                }
            }
        }
    }
}

// one slice from anywhere (a decoded file), clipped to the canvas as :58-66 does; pastes issued on one stream keep the
// reference's order, so a later slice wins where two overlap
__global__ void __launch_bounds__(256) paste_kernel(uint32_t* __restrict__ canvas, int W, int H, const uint32_t* __restrict__ slice,
                                                     int x0, int y0, int w, int cw, int ch) {
    for (int y = blockIdx.x * 8 + (threadIdx.x >> 5); y < ch; y += gridDim.x * 8)
        for (int x = threadIdx.x & 31; x < cw; x += 32) {
            const uint32_t px = slice[(int64_t)y * w + x];
            if (px >> 24) canvas[(int64_t)(y0 + y) * W + x0 + x] = px;
        }
}

cudaError_t launch_reassemble(const Tables& t, int v0, int nmaps, const unsigned long long* offsets, const uint8_t* slices,
                              uint8_t* canvases, cudaStream_t st) {
    if (nmaps <= 0) return cudaSuccess;
    const size_t map_bytes = (size_t)t.H * t.W * 4;
    cudaError_t e = cudaMemsetAsync(canvases, 0, map_bytes * (size_t)nmaps, st);   // cv::Scalar(0,0,0,0), :50
    if (e != cudaSuccess) return e;
    // CTAs per map: enough that the maps in flight (canvas + slices, ~2.5x the canvas) share the L2, see launch_slice_emit
    const long long resident = (long long)sm_count() * 32;
    const long long in_flight = std::max(1ll, (48ll << 20) / std::max<long long>(1, (long long)map_bytes * 5 / 2));
    long long P = std::min(1024ll, (resident + in_flight - 1) / in_flight);
    if (nmaps == 1) P = sm_count() * 4;
    while (P > 1 && P * nmaps > (1ll << 30)) P >>= 1;
    reassemble_kernel<<<(unsigned)(P * nmaps), EMIT_T, 0, st>>>(t, v0, (int)P, offsets, reinterpret_cast<const uint32_t*>(slices),
                                                                reinterpret_cast<uint32_t*>(canvases));
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_paste(uint8_t* canvas, int W, int H, const uint8_t* slice, int x, int y, int w, int h, cudaStream_t st) {
    const int cw = min(w, W - x), ch = min(h, H - y);                              // :58-59
    if (cw <= 0 || ch <= 0) return cudaSuccess;                                    // "lies outside canvas, skipping", :61-64
    const int grid = min((ch + 7) / 8, sm_count() * 4);
    paste_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<uint32_t*>(canvas), W, H, reinterpret_cast<const uint32_t*>(slice), x, y, w, cw, ch);
    count_launch();
    return cudaGetLastError();
}

}  // namespace icr
