// estimate: Rules C, P, F, R, H, Z for every (label map, segment) of the batch in ONE launch.
// Replaces the reference's 13 stream operations per segment (src/png_size_estimator.cu:505-590),
// its two nested host loops (src/compute_rewards.cu:125,159) and its RGBA staging kernel (:23-53).
//
// Persistent warps pull segments from the flat work list prep built (4 at a time, one global atomic).
// A warp owns one segment.  It walks the bbox in groups of R rows: lanes along x stage the group
// into a shared-memory tile with coalesced loads straight from the caller's buffers -- raw label
// compared with the segment's key (Rule P), three f32 planes clamped/rounded/packed to RGBA
// (Rule C) -- so a masked pixel is one 32-bit word.  Then every lane owns a (row, strip<=8 px) of
// the tile: neighbours come from the tile (the row above the group stays in a rolling slot, a zero
// guard column stands left of x=0), the five filter costs are accumulated with packed-byte
// arithmetic in registers, strips of a row are combined by a butterfly, the filter is chosen, and
// the chosen filter's residual bytes go to a per-warp 4x256 histogram in shared memory.  Zero
// residuals are never sent to the table; bin 0 is recovered as w*h minus the rest.  No residual
// stream, RGBA plane or compact-id plane is ever written to memory.  Segments that need Rule M
// match statistics (chunk >= L_min) or are wider than 256 px are left to the big-segment kernel.
#include "icr_internal.h"
#include "icr_pixel.cuh"

namespace icr {

constexpr int EST_WARPS = 4;
constexpr int TILE_WORDS = 544;       // max over strip counts of (R+1)*(S*sw+2) + read-ahead padding
constexpr int GRAB = 4;               // segments taken per global atomic

struct DebugOut {
    int v, k;
    uint32_t* row_costs; uint8_t* filter_id; uint32_t* hist; unsigned long long* match; double* size;
};

__device__ __forceinline__ double clog2_of(uint32_t c, const double* __restrict__ tab) {
    return c < (uint32_t)CLOG2_N ? __ldg(tab + c) : (double)c * log2((double)c);
}

// Rule Z (src/png_size_estimator.cu:10-28), every operand widened to double as at :588
__device__ __forceinline__ double size_formula(double Hbar, unsigned long long ms, unsigned long long mc,
                                               unsigned long long mls, unsigned long long N, int h, const EstParams& p) {
    double f = (N > 0 && ms > 0) ? (double)ms / (double)N : 0.0;
    double Lbar = mc > 0 ? (double)mls / (double)mc : (double)p.L_min;
    double b_lit = Hbar + p.beta;
    double b_match = p.token / Lbar + p.gamma;
    double b_data = (1.0 - f) * b_lit + f * b_match;
    return (p.overhead + (double)h) + ((double)N * b_data) / 8.0;
}

__device__ __forceinline__ uint32_t to_byte(float v) {   // Rule C (src/compute_rewards.cu:15-20)
    return (uint32_t)__float2int_rn(fminf(fmaxf(v, 0.f), 255.f));
}

// FULL: the segment fills its bbox (count == w*h), so no label has to be read at all
template <typename LT, bool FULL>
__device__ __forceinline__ uint32_t masked_rgba(const float* __restrict__ img, const LT* __restrict__ lab,
                                                int64_t HW, int64_t idx, LT key) {
    // all four loads are issued together; the mask is applied afterwards (one memory round trip, not two)
    const float r = __ldg(img + idx), g = __ldg(img + HW + idx), b = __ldg(img + 2 * HW + idx);
    const uint32_t px = to_byte(r) | to_byte(g) << 8 | to_byte(b) << 16 | 0xff000000u;
    if (FULL) return px;
    return __ldg(lab + idx) == key ? px : 0u;
}

template <typename LT, bool DEBUG>
__device__ void warp_segment(const float* __restrict__ img, const LT* __restrict__ lab, int64_t HW, int W, LT key, int count,
                             int4 bb, uint32_t* hist, uint32_t* tile, const EstParams& p,
                             const double* __restrict__ clog2, double* out, const DebugOut& dbg) {
    const int lane = threadIdx.x & 31;
    const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
    if (w <= 0 || h <= 0) { if (lane == 0) *out = 0.0; return; }   // src/compute_rewards.cu:164

    // strips per row S = 2^lg with strip width sw <= 8; among those minimise (row groups) x (strip width + fixed cost)
    int lg = 0;
    {
        int best = 0x7fffffff;
        for (int l = 0; l <= 5; ++l) {
            const int swl = (w + (1 << l) - 1) >> l;
            if (swl > 8) continue;
            const int cost = (((h << l) + 31) >> 5) * (swl + 3);
            if (cost < best) { best = cost; lg = l; }
        }
    }
    const int S = 1 << lg, R = 32 >> lg;
    const int sr = lane >> lg, stp = lane & (S - 1);
    const int sw = (w + S - 1) >> lg;
    const int xs = stp * sw;
    const int nstrip = max(0, min(sw, w - xs));
    const int stride = (S * sw + 1) | 1;                 // guard column + row, odd so rows land on different banks
    // staging geometry: lanes along x, tw = power of two covering min(w, 32)
    int twl = 0;
    while ((1 << twl) < w && twl < 5) ++twl;
    const int tx = lane & ((1 << twl) - 1), ty = lane >> twl, rpp = 32 >> twl;

    // zero the guard column of every slot and the slot of row -1
    for (int s = lane; s <= R; s += 32) tile[s * stride] = 0u;
    for (int x = lane; x < stride + 8; x += 32) tile[x] = 0u;
    int s0 = 1;                                          // slot of the group's first row; row y lives in slot (y+1) mod (R+1)
    const bool full = count == w * h;
    const float* imgp = img + (int64_t)bb.y * W + bb.x;
    const LT* labp = lab + (int64_t)bb.y * W + bb.x;

    for (int yb = 0; yb < h; yb += R) {
        __syncwarp();
        // ---- stage rows yb .. yb+R-1 (coalesced along x), masked and packed
        for (int r0 = 0; r0 < R; r0 += rpp) {
            const int rr = r0 + ty, yy = yb + rr;
            if (rr < R && yy < h) {
                int slot = s0 + rr; if (slot > R) slot -= R + 1;
                uint32_t* trow = tile + slot * stride + 1;
                const int64_t rowoff = (int64_t)yy * W;
                if (full) { for (int x = tx; x < w; x += (1 << twl)) trow[x] = masked_rgba<LT, true>(imgp, labp, HW, rowoff + x, key); }
                else { for (int x = tx; x < w; x += (1 << twl)) trow[x] = masked_rgba<LT, false>(imgp, labp, HW, rowoff + x, key); }
            }
        }
        __syncwarp();
        // ---- lane = (row sr of the group, strip stp)
        const int y = yb + sr;
        const int n = y < h ? nstrip : 0;
        int slot = s0 + sr; if (slot > R) slot -= R + 1;
        const int uslot = slot == 0 ? R : slot - 1;
        const uint32_t* base = tile + slot * stride + 1 + xs;
        const uint32_t* ub = tile + uslot * stride + 1 + xs;
        uint32_t cur[8], up[8], pa[8], pp[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { cur[i] = base[i]; up[i] = ub[i]; }
        const uint32_t left0 = base[-1], ul0 = ub[-1];
        int f = 4;
        if (p.adaptive) {
            Costs5 c = {0, 0, 0, 0, 0};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t left = i ? cur[i - 1] : left0, ul = i ? up[i - 1] : ul0;
                pa[i] = avg4(left, up[i]);
                pp[i] = paeth4(left, up[i], ul);
                if (i < n) {
                    c.none = cost_acc(cur[i], c.none);
                    c.sub = cost_acc(absdiff4(cur[i], left), c.sub);
                    c.up = cost_acc(absdiff4(cur[i], up[i]), c.up);
                    c.avg = cost_acc(absdiff4(cur[i], pa[i]), c.avg);
                    c.paeth = cost_acc(absdiff4(cur[i], pp[i]), c.paeth);
                }
            }
            for (int o = 1; o < S; o <<= 1) {
                c.none += __shfl_xor_sync(0xffffffffu, c.none, o);
                c.sub += __shfl_xor_sync(0xffffffffu, c.sub, o);
                c.up += __shfl_xor_sync(0xffffffffu, c.up, o);
                c.avg += __shfl_xor_sync(0xffffffffu, c.avg, o);
                c.paeth += __shfl_xor_sync(0xffffffffu, c.paeth, o);
            }
            f = pick_filter(c);
            if (DEBUG && y < h && stp == 0 && dbg.row_costs) {
                uint32_t* rc = dbg.row_costs + 5 * y;
                rc[0] = c.none; rc[1] = c.sub; rc[2] = c.up; rc[3] = c.avg; rc[4] = c.paeth;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) { pa[i] = 0u; pp[i] = paeth4(i ? cur[i - 1] : left0, up[i], i ? up[i - 1] : ul0); }
        }
        if (DEBUG && y < h && stp == 0 && dbg.filter_id) dbg.filter_id[y] = (uint8_t)f;
        // ---- residual bytes of the chosen filter -> histogram (Rules R, H)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (i < n) {
                const uint32_t left = i ? cur[i - 1] : left0;
                uint32_t pred = f == 1 ? left : 0u;
                if (f == 2) pred = up[i];
                if (f == 3) pred = pa[i];
                if (f == 4) pred = pp[i];
                const uint32_t r = __vsub4(cur[i], pred);
                if (r) {
                    const uint32_t a0 = (r << 2) & 0x3fcu, a1 = (r >> 6) & 0x3fcu, a2 = (r >> 14) & 0x3fcu, a3 = (r >> 22) & 0x3fcu;
                    char* hb = reinterpret_cast<char*>(hist);
                    if (a0) atomicAdd(reinterpret_cast<uint32_t*>(hb + a0), 1u);
                    if (a1) atomicAdd(reinterpret_cast<uint32_t*>(hb + 1024 + a1), 1u);
                    if (a2) atomicAdd(reinterpret_cast<uint32_t*>(hb + 2048 + a2), 1u);
                    if (a3) atomicAdd(reinterpret_cast<uint32_t*>(hb + 3072 + a3), 1u);
                }
            }
        }
        s0 = s0 == 0 ? R : s0 - 1;                       // (yb + R + 1) mod (R+1)
    }
    __syncwarp();

    // Rule H: H_c = -sum p log2 p = log2(n) - (sum c log2 c)/n ; lane owns 32 bins of channel lane/8
    const uint32_t npx = (uint32_t)w * (uint32_t)h;
    double acc = 0.0;
    uint32_t nz = 0;
    uint4* h4 = reinterpret_cast<uint4*>(hist) + lane * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        uint4 q = h4[j];
        if (DEBUG && dbg.hist) {
            uint32_t* d = dbg.hist + lane * 32 + j * 4;
            d[0] = q.x; d[1] = q.y; d[2] = q.z; d[3] = q.w;
        }
        if (q.x | q.y | q.z | q.w) {
            if (q.x) { acc += clog2_of(q.x, clog2); nz += q.x; }
            if (q.y) { acc += clog2_of(q.y, clog2); nz += q.y; }
            if (q.z) { acc += clog2_of(q.z, clog2); nz += q.z; }
            if (q.w) { acc += clog2_of(q.w, clog2); nz += q.w; }
            h4[j] = make_uint4(0, 0, 0, 0);      // leave the table clean for the next segment
        }
    }
    for (int o = 1; o < 8; o <<= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
        nz += __shfl_xor_sync(0xffffffffu, nz, o);
    }
    const uint32_t zeros = npx - nz;
    if (zeros) acc += clog2_of(zeros, clog2);
    if (DEBUG && dbg.hist) { __syncwarp(); if ((lane & 7) == 0) dbg.hist[(lane >> 3) * 256] = zeros; }
    const double Hc = log2((double)npx) - acc / (double)npx;
    const double H0 = __shfl_sync(0xffffffffu, Hc, 0), H1 = __shfl_sync(0xffffffffu, Hc, 8);
    const double H2 = __shfl_sync(0xffffffffu, Hc, 16), H3 = __shfl_sync(0xffffffffu, Hc, 24);
    if (lane == 0) {
        const double Hbar = ((H0 + H2) + (H1 + H3)) / 4.0;     // tree of mean_channels_kernel (:314-328)
        const double sz = size_formula(Hbar, 0, 0, 0, 4ull * npx, h, p);
        *out = sz;
        if (DEBUG) {
            if (dbg.match) { dbg.match[0] = 0; dbg.match[1] = 0; dbg.match[2] = 0; }
            if (dbg.size) *dbg.size = sz;
        }
    }
}

template <typename LT>
__global__ void __launch_bounds__(EST_WARPS * 32) estimate_kernel(Tables t, Inputs in, EstParams p) {
    __shared__ __align__(16) uint32_t hist_all[EST_WARPS][1024];
    __shared__ uint32_t tile_all[EST_WARPS][TILE_WORDS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t* hist = hist_all[warp];
    for (int i = lane; i < 1024; i += 32) hist[i] = 0;
    __syncwarp();
    const int64_t HW = (int64_t)t.H * t.W;
    const int total = t.seg_prefix[t.V];
    const LT* labels = static_cast<const LT*>(in.labels);
    DebugOut none{};
    for (;;) {
        int base = lane == 0 ? atomicAdd(&t.sched[1], GRAB) : 0;
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= total) break;
        // lanes 0..GRAB-1 locate their flat index: the map v with seg_prefix[v] <= flat < seg_prefix[v+1]
        int mv = 0, mk = 0;
        if (lane < GRAB && base + lane < total) {
            const int flat = base + lane;
            int lo = 0, hi = t.V;                       // invariant: seg_prefix[lo] <= flat < seg_prefix[hi]
            while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (__ldg(t.seg_prefix + mid) <= flat) lo = mid; else hi = mid; }
            mv = lo; mk = flat - __ldg(t.seg_prefix + lo);
        }
        const int cnt = min(GRAB, total - base);
        for (int j = 0; j < cnt; ++j) {
            const int v = __shfl_sync(0xffffffffu, mv, j), k = __shfl_sync(0xffffffffu, mk, j);
            const int4 bb = __ldg(t.bboxes + (int64_t)v * t.Kmax + k);
            if (for_big_kernel(bb.z - bb.x + 1, bb.w - bb.y + 1, p.L_min)) continue;
            const LT key = (LT)__ldg(t.keys + (int64_t)v * t.Kmax + k);
            warp_segment<LT, false>(in.images + (int64_t)(v / in.S) * 3 * HW, labels + (int64_t)v * HW, HW, t.W, key,
                                    __ldg(t.counts + (int64_t)v * t.Kmax + k), bb,
                                    hist, tile_all[warp], p, t.clog2, t.seg_sizes + (int64_t)v * t.Kmax + k, none);
        }
    }
}

template <typename LT>
__global__ void __launch_bounds__(32) estimate_debug_kernel(Tables t, Inputs in, EstParams p, DebugOut dbg) {
    __shared__ __align__(16) uint32_t hist[1024];
    __shared__ uint32_t tile[TILE_WORDS];
    for (int i = threadIdx.x; i < 1024; i += 32) hist[i] = 0;
    __syncwarp();
    const int64_t HW = (int64_t)t.H * t.W;
    const int4 bb = t.bboxes[(int64_t)dbg.v * t.Kmax + dbg.k];
    if (for_big_kernel(bb.z - bb.x + 1, bb.w - bb.y + 1, p.L_min)) return;     // the big-segment debug kernel acts
    const LT key = (LT)t.keys[(int64_t)dbg.v * t.Kmax + dbg.k];
    double scratch;
    warp_segment<LT, true>(in.images + (int64_t)(dbg.v / in.S) * 3 * HW, static_cast<const LT*>(in.labels) + (int64_t)dbg.v * HW,
                           HW, t.W, key, t.counts[(int64_t)dbg.v * t.Kmax + dbg.k], bb, hist, tile, p, t.clog2,
                           dbg.size ? dbg.size : &scratch, dbg);
}

static int est_grid() {
    static int g = 0;
    if (!g) {
        int dev = 0, sms = 148, a = 0, b = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, estimate_kernel<int32_t>, EST_WARPS * 32, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, estimate_kernel<int64_t>, EST_WARPS * 32, 0);
        int per = a < b ? a : b;
        if (per < 1) per = 1;
        g = sms * per;
    }
    return g;
}

cudaError_t launch_estimate(const Tables& t, const Inputs& in, const EstParams& p, cudaStream_t st) {
    const int grid = est_grid();     // persistent: every CTA resident, warps pull work until the list is empty
    if (in.label_dtype == ICR_LABELS_I64) estimate_kernel<int64_t><<<grid, EST_WARPS * 32, 0, st>>>(t, in, p);
    else estimate_kernel<int32_t><<<grid, EST_WARPS * 32, 0, st>>>(t, in, p);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_estimate_debug(const Tables& t, const Inputs& in, const EstParams& p, int v, int k, uint32_t* row_costs,
                                  uint8_t* filter_id, uint32_t* hist, unsigned long long* match, double* size_out,
                                  cudaStream_t st) {
    DebugOut d{v, k, row_costs, filter_id, hist, match, size_out};
    if (in.label_dtype == ICR_LABELS_I64) estimate_debug_kernel<int64_t><<<1, 32, 0, st>>>(t, in, p, d);
    else estimate_debug_kernel<int32_t><<<1, 32, 0, st>>>(t, in, p, d);
    count_launch();
    return cudaGetLastError();
}

}  // namespace icr
