// estimate: Rules P, F, R, H, Z for every (label map, segment) of the batch in ONE launch.
// Replaces the reference's 13 stream operations per segment (src/png_size_estimator.cu:505-590)
// and its two nested host loops (src/compute_rewards.cu:125,159).
//
// One warp owns one segment.  A lane owns a (row, column-strip) of the segment's bbox and walks
// it left to right: `left` and `up-left` live in registers, `up` comes from the lane that owns
// the row above (one shuffle), so each bbox pixel costs one 64-bit load {RGBA, id}, one compare
// for the mask (Rule P) and the packed-byte filter arithmetic of icr_pixel.cuh.  Row costs are
// lane-local (plus a log2(strips)-step butterfly), the filter is chosen in registers, and the
// row is walked a second time for the residual histogram, which lives in a per-warp 4x256 table
// in shared memory.  Zero residuals (the bulk: every masked-out pixel away from the segment) are
// never sent to the table; bin 0 is recovered as w*h minus the rest.  No residual stream is
// ever written.  Segments large enough for Rule M match statistics (chunk >= L_min) are left to
// the big-segment kernel.
#include "icr_internal.h"
#include "icr_pixel.cuh"

namespace icr {

constexpr int EST_WARPS = 8;

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

__device__ __forceinline__ bool needs_match_stats(int w, int h, int L_min) {
    // Rule M: chunk = ceil(N/65536); a run never leaves its chunk, so nothing matches when chunk < L_min
    long long N = 4ll * w * h;
    long long chunk = (N + 65535) / 65536;
    return chunk >= (long long)L_min;
}

__device__ __forceinline__ uint32_t masked(uint2 q, uint32_t k) { return q.y == k ? q.x : 0u; }

template <bool DEBUG>
__device__ void warp_segment(const uint2* __restrict__ map, int W, uint32_t k, int4 bb, uint32_t* hist,
                             const EstParams& p, const double* __restrict__ clog2, double* out, const DebugOut& dbg) {
    const int lane = threadIdx.x & 31;
    const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
    if (w <= 0 || h <= 0) { if (lane == 0) *out = 0.0; return; }   // src/compute_rewards.cu:164

    // strips per row: the power of two that minimises (row groups) x (strip width)
    int lg = 0;
    {
        int best = 0x7fffffff;
        for (int l = 0; l <= 5; ++l) {
            int S = 1 << l;
            int cost = ((h * S + 31) >> 5) * ((w + S - 1) >> l);
            if (cost < best) { best = cost; lg = l; }
        }
    }
    const int S = 1 << lg, R = 32 >> lg;
    const int sr = lane >> lg, stp = lane & (S - 1);
    const int sw = (w + S - 1) >> lg;
    const int xs = stp * sw, xe = min(xs + sw, w);

    for (int yb = 0; yb < h; yb += R) {
        const int y = yb + sr;
        const bool rowact = y < h;
        const uint2* rowp = map + (int64_t)(bb.y + y) * W + bb.x;
        const uint2* upp = rowp - W;
        const bool has_up = y > 0;
        uint32_t left0 = 0, upleft0 = 0;
        if (rowact && xs > 0 && xs < xe) {
            left0 = masked(__ldg(rowp + xs - 1), k);
            if (has_up) upleft0 = masked(__ldg(upp + xs - 1), k);
        }
        int f = 4;
        if (p.adaptive) {
            Costs5 c = {0, 0, 0, 0, 0};
            uint32_t left = left0, upleft = upleft0;
            for (int i = 0; i < sw; ++i) {
                const int x = xs + i;
                const bool valid = rowact && x < xe;
                uint32_t cur = valid ? masked(__ldg(rowp + x), k) : 0u;
                uint32_t up = __shfl_up_sync(0xffffffffu, cur, S);
                if (sr == 0) up = (valid && has_up) ? masked(__ldg(upp + x), k) : 0u;
                if (valid) add_costs(c, cur, left, up, upleft);
                left = cur; upleft = up;
            }
            for (int o = 1; o < S; o <<= 1) {
                c.none += __shfl_xor_sync(0xffffffffu, c.none, o);
                c.sub += __shfl_xor_sync(0xffffffffu, c.sub, o);
                c.up += __shfl_xor_sync(0xffffffffu, c.up, o);
                c.avg += __shfl_xor_sync(0xffffffffu, c.avg, o);
                c.paeth += __shfl_xor_sync(0xffffffffu, c.paeth, o);
            }
            f = pick_filter(c);
            if (DEBUG && rowact && stp == 0 && dbg.row_costs) {
                uint32_t* rc = dbg.row_costs + 5 * y;
                rc[0] = c.none; rc[1] = c.sub; rc[2] = c.up; rc[3] = c.avg; rc[4] = c.paeth;
            }
        }
        if (DEBUG && rowact && stp == 0 && dbg.filter_id) dbg.filter_id[y] = (uint8_t)f;
        {
            uint32_t left = left0, upleft = upleft0;
            for (int i = 0; i < sw; ++i) {
                const int x = xs + i;
                const bool valid = rowact && x < xe;
                uint32_t cur = valid ? masked(__ldg(rowp + x), k) : 0u;
                uint32_t up = __shfl_up_sync(0xffffffffu, cur, S);
                if (sr == 0) up = (valid && has_up) ? masked(__ldg(upp + x), k) : 0u;
                if (valid) {
                    uint32_t r = residual4(f, cur, left, up, upleft);
                    if (r) {
                        uint32_t b0 = r & 255u, b1 = (r >> 8) & 255u, b2 = (r >> 16) & 255u, b3 = r >> 24;
                        if (b0) atomicAdd(&hist[b0], 1u);
                        if (b1) atomicAdd(&hist[256 + b1], 1u);
                        if (b2) atomicAdd(&hist[512 + b2], 1u);
                        if (b3) atomicAdd(&hist[768 + b3], 1u);
                    }
                }
                left = cur; upleft = up;
            }
        }
    }
    __syncwarp();

    // Rule H: H_c = -sum p log2 p = log2(n) - (sum c log2 c)/n ; lane owns 32 bins of channel lane/8
    const uint32_t n = (uint32_t)w * (uint32_t)h;
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
    const uint32_t zeros = n - nz;
    if (zeros) acc += clog2_of(zeros, clog2);
    if (DEBUG && dbg.hist) { __syncwarp(); if ((lane & 7) == 0) dbg.hist[(lane >> 3) * 256] = zeros; }
    const double Hc = log2((double)n) - acc / (double)n;
    const double H0 = __shfl_sync(0xffffffffu, Hc, 0), H1 = __shfl_sync(0xffffffffu, Hc, 8);
    const double H2 = __shfl_sync(0xffffffffu, Hc, 16), H3 = __shfl_sync(0xffffffffu, Hc, 24);
    if (lane == 0) {
        const double Hbar = ((H0 + H2) + (H1 + H3)) / 4.0;     // tree of mean_channels_kernel (:314-328)
        const double sz = size_formula(Hbar, 0, 0, 0, 4ull * n, h, p);
        *out = sz;
        if (DEBUG) {
            if (dbg.match) { dbg.match[0] = 0; dbg.match[1] = 0; dbg.match[2] = 0; }
            if (dbg.size) *dbg.size = sz;
        }
    }
}

__global__ void __launch_bounds__(EST_WARPS * 32) estimate_kernel(Tables t, EstParams p) {
    __shared__ __align__(16) uint32_t hist_all[EST_WARPS][1024];
    const int v = blockIdx.y;
    const int warp = threadIdx.x >> 5;
    uint32_t* hist = hist_all[warp];
    for (int i = threadIdx.x & 31; i < 1024; i += 32) hist[i] = 0;
    __syncwarp();
    if (t.status[v] != ST_OK) return;
    const int K = t.K[v];
    const uint2* map = t.pixlab + (int64_t)v * t.H * t.W;
    const int4* bbs = t.bboxes + (int64_t)v * t.Kmax;
    double* sizes = t.seg_sizes + (int64_t)v * t.Kmax;
    DebugOut none{};
    for (int k = blockIdx.x * EST_WARPS + warp; k < K; k += gridDim.x * EST_WARPS) {
        const int4 bb = __ldg(bbs + k);
        if (needs_match_stats(bb.z - bb.x + 1, bb.w - bb.y + 1, p.L_min)) continue;   // big-segment kernel
        warp_segment<false>(map, t.W, (uint32_t)k, bb, hist, p, t.clog2, sizes + k, none);
    }
}

__global__ void __launch_bounds__(32) estimate_debug_kernel(Tables t, EstParams p, DebugOut dbg) {
    __shared__ __align__(16) uint32_t hist[1024];
    for (int i = threadIdx.x; i < 1024; i += 32) hist[i] = 0;
    __syncwarp();
    const uint2* map = t.pixlab + (int64_t)dbg.v * t.H * t.W;
    const int4 bb = t.bboxes[(int64_t)dbg.v * t.Kmax + dbg.k];
    double scratch;
    warp_segment<true>(map, t.W, (uint32_t)dbg.k, bb, hist, p, t.clog2, dbg.size ? dbg.size : &scratch, dbg);
}

cudaError_t launch_estimate(const Tables& t, const EstParams& p, cudaStream_t st) {
    // enough CTAs to fill 148 SMs several times over, never more than the segments can use
    int gx = (2368 + t.V - 1) / t.V;
    int need = (t.Kmax + EST_WARPS - 1) / EST_WARPS;
    if (gx > need) gx = need;
    if (gx < 1) gx = 1;
    dim3 grid(gx, t.V);
    estimate_kernel<<<grid, EST_WARPS * 32, 0, st>>>(t, p);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_estimate_debug(const Tables& t, const EstParams& p, int v, int k, uint32_t* row_costs,
                                  uint8_t* filter_id, uint32_t* hist, unsigned long long* match, double* size_out,
                                  cudaStream_t st) {
    DebugOut d{v, k, row_costs, filter_id, hist, match, size_out};
    estimate_debug_kernel<<<1, 32, 0, st>>>(t, p, d);
    count_launch();
    return cudaGetLastError();
}

}  // namespace icr
