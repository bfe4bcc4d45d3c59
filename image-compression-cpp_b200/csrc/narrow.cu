// narrow: Rules C, P, F, R, H, Z for every segment whose bbox is at most NARROW_MAX_W pixels wide, one warp per segment,
// the whole batch in ONE launch.  Replaces the reference's 13 stream operations per segment
// (src/png_size_estimator.cu:505-590), its two nested host loops (src/compute_rewards.cu:125,159) and its RGBA staging
// kernel (:23-53) for the bulk of the pixels.
//
// Register path: no shared-memory tile, no second walk.  A lane owns one COLUMN of the bbox (two with 32 < w <= 64) and
// the warp walks down the rows.  The three f32 planes (and the raw label unless the segment fills its bbox) arrive by
// plain coalesced loads two rows ahead; conversion to a masked RGBA word (Rules C, P) happens in registers.  The
// neighbours of a pixel never touch memory: `up` and `upleft` are the lane's own pixel and left neighbour of the row
// before, `left` is one shuffle.  Segments no wider than 16 pixels put two rows side by side in the warp; the row
// above then comes from the lanes w to the left by one more shuffle.
// Per row: five filter costs per lane (packed bytes: VABSDIFF4 for |cur - pred|, a signed VABSDIFF4 against zero for the
// sum of |int8|; Paeth on half2 pairs entirely on the FMA pipe), five warp sums (REDUX), arg-min with ties to the lowest
// id (Rule F), and right away the residual of the chosen filter from the operands still in registers -> four
// unconditional shared-memory atomics into the warp's 4x256 histogram (Rules R, H).  Lanes outside the bbox are steered
// to bin 0 by their byte-permute selector and subtracted again in the epilogue, so no atomic is predicated.
// Epilogue: c*log2(c) table entropy, Rule Z, one f64 store; the warp that stores the last size of a label map also
// computes its reward (Rule W).  Match statistics are zero here by construction: Rule M segments go to the big-segment
// kernel, bboxes wider than NARROW_MAX_W to the shared-memory path (estimate.cu).
#include "icr_internal.h"
#include "icr_pixel.cuh"
#include "icr_reward.cuh"

namespace icr {

constexpr int NRW_WARPS = 4;              // warps per CTA; warps are independent
constexpr int NRW_DEPTH = 2;              // rows in flight ahead of the row being filtered
constexpr uint32_t ZH = 0x64006400u;      // half2 {1024, 1024}: the biased form of two zero bytes

__device__ __noinline__ double nrw_clog2_slow(uint32_t c) { return (double)c * log2((double)c); }
__device__ __forceinline__ double nrw_clog2(uint32_t c, const double* __restrict__ tab) {
    return c < (uint32_t)CLOG2_N ? __ldg(tab + c) : nrw_clog2_slow(c);
}

template <typename LT, int C>
struct RawRow { float r[C], g[C], b[C]; LT l[C]; };

// One segment, one warp.  img / lab point at the planes of the segment's label map; hist is the warp's zeroed 4x256 table.
template <typename LT, int C, bool MULTI, bool FULL, bool DEBUG>
__device__ __forceinline__ void narrow_rows(const float* __restrict__ img, const LT* __restrict__ lab, uint32_t HW, int W, LT key,
                                            int x0, int y0, int w, int h, uint32_t* hist, bool adaptive, int lane,
                                            const DebugOut& dbg, uint32_t& slots) {
    static_assert(!(MULTI && C != 1), "rows side by side only for single-chunk widths");
    // ---- lane geometry
    // MULTI (w <= 16): two rows side by side, row r of a step in lanes [r * w, (r + 1) * w).  (More rows would fit the
    // narrowest bboxes, but a masked warp reduction runs once per group and lane: two rows keep the five row sums of both
    // groups in the 16-bit halves of five plain REDUX.)
    constexpr int R = MULTI ? 2 : 1;
    int r = 0, x = lane;
    if (MULTI) {
        r = lane >= w ? 1 : 0;
        x = lane - r * w;
    }
    const bool active0 = MULTI ? lane < R * w : lane < w;
    const bool active1 = C > 1 && lane + 32 < w;
    const int from = MULTI ? (active0 ? (r > 0 ? lane - w : lane + (R - 1) * w) : lane) : lane;
    const bool give_prev = MULTI && r == R - 1;
    const uint32_t field_mul = !MULTI ? 1u : (active0 ? (r ? 0x10000u : 1u) : 0u);     // a group's row sums live in its 16-bit half
    const uint32_t field_sel = (MULTI && r) ? 0x4432u : 0x4410u;
    const uint32_t m0 = (active0 && x > 0) ? 1u : 0u, m1 = active1 ? 1u : 0u;      // left neighbour exists and counts
    // upper half2 word of a pixel with blue = 0: alpha 255 inside the bbox, zero outside (those lanes never load: r = g = b = 0)
    const uint32_t zh_alpha0 = active0 ? 0x64ff6400u : ZH, zh_alpha1 = active1 ? 0x64ff6400u : ZH;
    const uint32_t key_add = lane == 0 ? 1u : 0u;        // the lane that adds the filter ids to the cost keys
    // byte selectors of the histogram indices: byte 0 / byte 2 of a residual word for a lane inside the bbox, a zero byte else
    const uint32_t sel_a0 = active0 ? 0x4440u : 0x4444u, sel_b0 = active0 ? 0x4442u : 0x4444u;
    const uint32_t sel_a1 = active1 ? 0x4440u : 0x4444u, sel_b1 = active1 ? 0x4442u : 0x4444u;
    const int RW = R * W;
    const int nsteps = (h + R - 1) / R;
    slots = (uint32_t)nsteps * 32u * (uint32_t)C;

    // the pixel the next issued load fetches (chunk 0), one running pointer per plane
    const uint32_t off0 = (uint32_t)((y0 + r) * W + x0 + x);
    const float* pr = img + off0;
    const float* pg = pr + HW;
    const float* pb = pg + HW;
    const LT* pl = lab + off0;
    int row_pf = r;

    RawRow<LT, C> raw[NRW_DEPTH];
#pragma unroll
    for (int d = 0; d < NRW_DEPTH; ++d)
#pragma unroll
        for (int c = 0; c < C; ++c) { raw[d].r[c] = 0.f; raw[d].g[c] = 0.f; raw[d].b[c] = 0.f; raw[d].l[c] = key; }

    auto issue = [&](RawRow<LT, C>& q) {
        const bool rowok = row_pf < h;
        if (active0 && rowok) {
            q.r[0] = __ldg(pr); q.g[0] = __ldg(pg); q.b[0] = __ldg(pb);
            if (!FULL) q.l[0] = __ldg(pl);
        }
        if (C > 1 && active1 && rowok) {
            q.r[C - 1] = __ldg(pr + 32); q.g[C - 1] = __ldg(pg + 32); q.b[C - 1] = __ldg(pb + 32);
            if (!FULL) q.l[C - 1] = __ldg(pl + 32);
        }
        pr += RW; pg += RW; pb += RW;
        if (!FULL) pl += RW;
        row_pf += R;
    };
#pragma unroll
    for (int d = 0; d < NRW_DEPTH; ++d) issue(raw[d]);

    // ---- state carried from row to row: the pixels of the row above and their left neighbours
    uint32_t up_p[C], up_lo[C], up_hi[C], ul_lo[C], ul_hi[C], ulp[C];
#pragma unroll
    for (int c = 0; c < C; ++c) { up_p[c] = 0u; up_lo[c] = ZH; up_hi[c] = ZH; ul_lo[c] = ZH; ul_hi[c] = ZH; ulp[c] = 0u; }

    auto step = [&](RawRow<LT, C>& q, int yrow, bool refill) {
        const int y = yrow + r;
        uint32_t cur_p[C], cur_lo[C], cur_hi[C], lf_p[C], lf_lo[C], lf_hi[C], av_p[C], pa_lo[C], pa_hi[C];
        uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0;
        // Rule C + Rule P: masked RGBA words of this row, straight in the three forms the filters want: the two biased
        // half2 words by multiply-adds on the converted bytes (FMA pipe), the packed word by one byte permute of those
#pragma unroll
        for (int c = 0; c < C; ++c) {
            uint32_t lo = byte_of(q.g[c]) * 0x10000u + (byte_of(q.r[c]) + ZH);
            uint32_t hi = byte_of(q.b[c]) + (c == 0 ? zh_alpha0 : zh_alpha1);
            if (!FULL) { const bool in = q.l[c] == key; lo = in ? lo : ZH; hi = in ? hi : ZH; }
            cur_lo[c] = lo; cur_hi[c] = hi;
            cur_p[c] = __byte_perm(lo, hi, 0x6420);
        }
        // the buffer is free again: its refill (the row NRW_DEPTH steps ahead) flies under the arithmetic of this row
        if (refill) issue(q);
        // left neighbours: one shuffle; the first column and the lanes outside the bbox see zero
#pragma unroll
        for (int c = 0; c < C; ++c) {
            uint32_t l;
            if (c == 0) l = __shfl_up_sync(0xffffffffu, cur_p[0], 1);
            else l = __shfl_sync(0xffffffffu, lane == 31 ? cur_p[0] : cur_p[c], (lane + 31) & 31);
            l *= c == 0 ? m0 : m1;
            lf_p[c] = l; lf_lo[c] = unpack_lo(l); lf_hi[c] = unpack_hi(l);
        }
        if (MULTI) {
            // rows side by side: the row above sits w lanes to the left, the first row's in the last group of the step before
            const uint32_t u = __shfl_sync(0xffffffffu, give_prev ? up_p[0] : cur_p[0], from);
            const uint32_t ul = __shfl_sync(0xffffffffu, give_prev ? ulp[0] : lf_p[0], from);
            up_p[0] = u; up_lo[0] = unpack_lo(u); up_hi[0] = unpack_hi(u);
            ul_lo[0] = unpack_lo(ul); ul_hi[0] = unpack_hi(ul);
        }
        // Rule F: five costs of this lane's pixels
#pragma unroll
        for (int c = 0; c < C; ++c) {
            av_p[c] = avg4(lf_p[c], up_p[c]);
            pa_lo[c] = paeth_h2_fma(lf_lo[c], up_lo[c], ul_lo[c]);
            pa_hi[c] = paeth_h2_fma(lf_hi[c], up_hi[c], ul_hi[c]);
            const uint32_t pa_p = __byte_perm(pa_lo[c], pa_hi[c], 0x6420);
            c0 = cost_acc(cur_p[c], c0);
            c1 = cost_acc(absdiff4(cur_p[c], lf_p[c]), c1);
            c2 = cost_acc(absdiff4(cur_p[c], up_p[c]), c2);
            c3 = cost_acc(absdiff4(cur_p[c], av_p[c]), c3);
            c4 = cost_acc(absdiff4(cur_p[c], pa_p), c4);
        }
        int f = 4;
        if (adaptive || DEBUG) {
            // Rule F arg-min with ties to the lowest id = the minimum over keys cost * 8 + id.  With one row per step the
            // keys come out of the warp sums directly: every lane contributes cost * 8 and lane 0 adds the ids
            // (multiply-adds on the FMA pipe).  With two rows per step the row sums of both rows share a REDUX in its two
            // 16-bit halves (per-lane costs stay below 2^10 and a row has at most 16 lanes), and the keys are built after.
            uint32_t k0, k1, k2, k3, k4;
            if (MULTI) {
                k0 = __byte_perm(__reduce_add_sync(0xffffffffu, c0 * field_mul), 0u, field_sel) * 8u;
                k1 = __byte_perm(__reduce_add_sync(0xffffffffu, c1 * field_mul), 0u, field_sel) * 8u + 1u;
                k2 = __byte_perm(__reduce_add_sync(0xffffffffu, c2 * field_mul), 0u, field_sel) * 8u + 2u;
                k3 = __byte_perm(__reduce_add_sync(0xffffffffu, c3 * field_mul), 0u, field_sel) * 8u + 3u;
                k4 = __byte_perm(__reduce_add_sync(0xffffffffu, c4 * field_mul), 0u, field_sel) * 8u + 4u;
            } else {
                k0 = __reduce_add_sync(0xffffffffu, c0 * 8u);
                k1 = __reduce_add_sync(0xffffffffu, c1 * 8u + key_add);
                k2 = __reduce_add_sync(0xffffffffu, c2 * 8u + 2u * key_add);
                k3 = __reduce_add_sync(0xffffffffu, c3 * 8u + 3u * key_add);
                k4 = __reduce_add_sync(0xffffffffu, c4 * 8u + 4u * key_add);
            }
            if (adaptive) f = (int)(min(min(min(k0, k1), min(k2, k3)), k4) & 7u);
            c0 = k0 >> 3; c1 = k1 >> 3; c2 = k2 >> 3; c3 = k3 >> 3; c4 = k4 >> 3;       // read by the debug tables only
        }
        const bool rowok = !MULTI || y < h;
        if (DEBUG && x == 0 && active0 && rowok) {
            if (dbg.row_costs && adaptive) {
                uint32_t* rc = dbg.row_costs + 5 * y;
                rc[0] = c0; rc[1] = c1; rc[2] = c2; rc[3] = c3; rc[4] = c4;
            }
            if (dbg.filter_id) dbg.filter_id[y] = (uint8_t)f;
        }
        // Rules R, H: residual bytes of the chosen filter, as the low bytes of two half2 words, into the histogram.
        // Pixels outside the bbox (and rows past its end) select a zero byte: they land in bin 0 and are subtracted later.
        uint32_t rlo[C], rhi[C];
        switch (f) {
            case 0:
#pragma unroll
                for (int c = 0; c < C; ++c) { rlo[c] = cur_lo[c]; rhi[c] = cur_hi[c]; }
                break;
            case 1:
#pragma unroll
                for (int c = 0; c < C; ++c) { rlo[c] = residual_h2(cur_lo[c], lf_lo[c]); rhi[c] = residual_h2(cur_hi[c], lf_hi[c]); }
                break;
            case 2:
#pragma unroll
                for (int c = 0; c < C; ++c) { rlo[c] = residual_h2(cur_lo[c], up_lo[c]); rhi[c] = residual_h2(cur_hi[c], up_hi[c]); }
                break;
            case 3:
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    rlo[c] = residual_h2(cur_lo[c], unpack_lo(av_p[c])); rhi[c] = residual_h2(cur_hi[c], unpack_hi(av_p[c]));
                }
                break;
            default:
#pragma unroll
                for (int c = 0; c < C; ++c) { rlo[c] = residual_h2(cur_lo[c], pa_lo[c]); rhi[c] = residual_h2(cur_hi[c], pa_hi[c]); }
                break;
        }
#pragma unroll
        for (int c = 0; c < C; ++c) {
            uint32_t sa = c == 0 ? sel_a0 : sel_a1, sb = c == 0 ? sel_b0 : sel_b1;
            if (MULTI && !rowok) { sa = 0x4444u; sb = 0x4444u; }
            atomicAdd(hist + __byte_perm(rlo[c], 0u, sa), 1u);
            atomicAdd(hist + 256 + __byte_perm(rlo[c], 0u, sb), 1u);
            atomicAdd(hist + 512 + __byte_perm(rhi[c], 0u, sa), 1u);
            atomicAdd(hist + 768 + __byte_perm(rhi[c], 0u, sb), 1u);
        }
        // this row becomes the row above
#pragma unroll
        for (int c = 0; c < C; ++c) {
            if (MULTI) { up_p[c] = cur_p[c]; ulp[c] = lf_p[c]; }        // handed on by the shuffles of the next step
            else { up_p[c] = cur_p[c]; up_lo[c] = cur_lo[c]; up_hi[c] = cur_hi[c]; ul_lo[c] = lf_lo[c]; ul_hi[c] = lf_hi[c]; }
        }
    };

    // two steps per trip, so that the rows in flight alternate between two register sets without moves; an odd last step
    int ys = 0;
    static_assert(NRW_DEPTH == 2, "the loop below alternates two row buffers");
    for (; ys + R < h; ys += 2 * R) {
        step(raw[0], ys, true);
        step(raw[1], ys + R, true);
    }
    if (ys < h) step(raw[0], ys, false);
}

// Rule H entropy + Rule Z from the warp's histogram; leaves the histogram zeroed for the next segment.
template <bool DEBUG>
__device__ __forceinline__ double narrow_epilogue(uint32_t* hist, int w, int h, uint32_t slots, const EstParams& p,
                                                  const double* __restrict__ clog2, int lane, const DebugOut& dbg) {
    // H_c = -sum p log2 p = log2(n) - (sum c log2 c)/n ; lane owns 32 bins of channel lane/8, visited as eight quads
    // rotated by the lane index (unrotated, the eight lanes of a quarter-warp would share four banks)
    const uint32_t npx = (uint32_t)w * (uint32_t)h;
    const uint32_t surplus = slots - npx;                // lane slots outside the bbox: all of them were counted in bin 0
    double acc = 0.0;
    uint4* h4 = reinterpret_cast<uint4*>(hist) + lane * 8;
    const bool small = npx < (uint32_t)CLOG2_N;          // no bin can lie beyond the c*log2(c) table: branch-free lookups
#pragma unroll 4
    for (int j0 = 0; j0 < 8; ++j0) {
        const int j = (j0 + lane) & 7;
        uint4 q = h4[j];
        h4[j] = make_uint4(0, 0, 0, 0);
        if (j == 0 && (lane & 7) == 0) q.x -= surplus;
        if (DEBUG && dbg.hist) { uint32_t* d = dbg.hist + lane * 32 + j * 4; d[0] = q.x; d[1] = q.y; d[2] = q.z; d[3] = q.w; }
        if (small) acc += (__ldg(clog2 + q.x) + __ldg(clog2 + q.y)) + (__ldg(clog2 + q.z) + __ldg(clog2 + q.w));
        else acc += (nrw_clog2(q.x, clog2) + nrw_clog2(q.y, clog2)) + (nrw_clog2(q.z, clog2) + nrw_clog2(q.w, clog2));
    }
    for (int o = 1; o < 8; o <<= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    const double Hc = (nrw_clog2(npx, clog2) - acc) / (double)npx;                    // log2(n) = (n log2 n)/n
    const double H0 = __shfl_sync(0xffffffffu, Hc, 0), H1 = __shfl_sync(0xffffffffu, Hc, 8);
    const double H2 = __shfl_sync(0xffffffffu, Hc, 16), H3 = __shfl_sync(0xffffffffu, Hc, 24);
    const double Hbar = ((H0 + H2) + (H1 + H3)) / 4.0;     // tree of mean_channels_kernel (:314-328)
    // Rule Z with no matches (f = 0): (1-0)*b_lit + 0*b_match is b_lit exactly, so the match term's division is not evaluated
    return (p.overhead + (double)h) + ((double)(4ull * npx) * (Hbar + p.beta)) / 8.0;
}

template <typename LT, bool DEBUG>
__device__ __forceinline__ double narrow_segment(const float* __restrict__ img, const LT* __restrict__ lab, uint32_t HW, int W, LT key,
                                                 int count, int4 bb, uint32_t* hist, const EstParams& p,
                                                 const double* __restrict__ clog2, int lane, const DebugOut& dbg) {
    const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
    const bool full = count == w * h;                    // the segment fills its bbox: no label has to be read
    const bool adaptive = p.adaptive != 0;
    uint32_t slots = 0;
    if (w > 32) {
        if (full) narrow_rows<LT, 2, false, true, DEBUG>(img, lab, HW, W, key, bb.x, bb.y, w, h, hist, adaptive, lane, dbg, slots);
        else narrow_rows<LT, 2, false, false, DEBUG>(img, lab, HW, W, key, bb.x, bb.y, w, h, hist, adaptive, lane, dbg, slots);
    } else if (w > 16) {
        if (full) narrow_rows<LT, 1, false, true, DEBUG>(img, lab, HW, W, key, bb.x, bb.y, w, h, hist, adaptive, lane, dbg, slots);
        else narrow_rows<LT, 1, false, false, DEBUG>(img, lab, HW, W, key, bb.x, bb.y, w, h, hist, adaptive, lane, dbg, slots);
    } else {
        if (full) narrow_rows<LT, 1, true, true, DEBUG>(img, lab, HW, W, key, bb.x, bb.y, w, h, hist, adaptive, lane, dbg, slots);
        else narrow_rows<LT, 1, true, false, DEBUG>(img, lab, HW, W, key, bb.x, bb.y, w, h, hist, adaptive, lane, dbg, slots);
    }
    __syncwarp();
    return narrow_epilogue<DEBUG>(hist, w, h, slots, p, clog2, lane, dbg);
}

__device__ __forceinline__ int fetch_add(int* counter) {
    int r;
    asm volatile("atom.global.add.s32 %0, [%1], 1;" : "=r"(r) : "l"(counter) : "memory");
    return r;
}

template <typename LT>
__global__ void __launch_bounds__(NRW_WARPS * 32) narrow_kernel(Tables t, Inputs in, EstParams p, RewardArgs rw) {
    __shared__ __align__(16) uint32_t s_hist[NRW_WARPS][1024];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t* hist = s_hist[warp];
    for (int i = lane; i < 1024; i += 32) hist[i] = 0;
    __syncwarp();
    const uint32_t HW = (uint32_t)t.H * (uint32_t)t.W;
    const int n_large = t.sched[SC_N_LARGE], total = n_large + t.sched[SC_N_SMALL];
    const LT* labels = static_cast<const LT*>(in.labels);
    DebugOut none{};
    // One segment per request, the larger ones first; the request for the next segment is issued before the current one
    // is worked on, so its latency is hidden.
    int flat = 0;
    if (lane == 0) flat = fetch_add(&t.sched[SC_CUR_NARROW]);
    flat = __shfl_sync(0xffffffffu, flat, 0);
    while (flat < total) {
        int fut = 0;
        if (lane == 0) fut = fetch_add(&t.sched[SC_CUR_NARROW]);
        const uint2 e = __ldg(t.list_narrow + (flat < n_large ? (int64_t)flat : t.list_cap - 1 - (int64_t)(flat - n_large)));
        const int v = (int)e.x, k = (int)e.y;
        const int64_t row = (int64_t)v * t.Kmax + k;
        const int4 bb = __ldg(t.bboxes + row);
        const LT key = (LT)__ldg(t.keys + row);
        const int count = __ldg(t.counts + row);
        const int K = __ldg(t.K + v);
        const double sz = narrow_segment<LT, false>(in.images + (int64_t)(v / in.S) * 3 * HW, labels + (int64_t)v * HW, HW, t.W, key, count,
                                                    bb, hist, p, t.clog2, lane, none);
        int last = 0;
        if (lane == 0) {
            t.seg_sizes[row] = sz;
            last = segment_done(t, v, K);
        }
        if (__shfl_sync(0xffffffffu, last, 0)) {         // every size of this map is final: Rule W
            __threadfence();
            reward_of_map<true>(t, p, rw.image_sizes, rw.S, rw.rewards, v, lane);
        }
        flat = __shfl_sync(0xffffffffu, fut, 0);
    }
}

template <typename LT>
__global__ void __launch_bounds__(32) narrow_debug_kernel(Tables t, Inputs in, EstParams p, DebugOut dbg) {
    __shared__ __align__(16) uint32_t hist[1024];
    const int lane = threadIdx.x;
    for (int i = lane; i < 1024; i += 32) hist[i] = 0;
    __syncwarp();
    const uint32_t HW = (uint32_t)t.H * (uint32_t)t.W;
    const int64_t row = (int64_t)dbg.v * t.Kmax + dbg.k;
    const int4 bb = t.bboxes[row];
    const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
    if (w <= 0 || h <= 0 || segment_class(w, h, p.L_min) != CLS_NARROW) return;     // another debug kernel acts
    const double sz = narrow_segment<LT, true>(in.images + (int64_t)(dbg.v / in.S) * 3 * HW,
                                               static_cast<const LT*>(in.labels) + (int64_t)dbg.v * HW, HW, t.W, (LT)t.keys[row],
                                               t.counts[row], bb, hist, p, t.clog2, lane, dbg);
    if (lane == 0) {
        if (dbg.match) { dbg.match[0] = 0; dbg.match[1] = 0; dbg.match[2] = 0; }
        if (dbg.size) *dbg.size = sz;
    }
}

template <typename LT>
static cudaError_t launch_narrow_t(const Tables& t, const Inputs& in, const EstParams& p, const RewardArgs& rw, cudaStream_t st) {
    static int per_sm[16] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    auto kern = narrow_kernel<LT>;
    int per = (dev >= 0 && dev < 16) ? per_sm[dev] : 0;
    if (!per) {
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, kern, NRW_WARPS * 32, 0);
        if (e != cudaSuccess) return e;
        if (per < 1) per = 1;
        if (dev >= 0 && dev < 16) per_sm[dev] = per;
    }
    kern<<<sm_count() * per, NRW_WARPS * 32, 0, st>>>(t, in, p, rw);       // persistent: warps pull segments until the list is empty
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_narrow(const Tables& t, const Inputs& in, const EstParams& p, const RewardArgs& rw, cudaStream_t st) {
    if (in.label_dtype == ICR_LABELS_I64) return launch_narrow_t<int64_t>(t, in, p, rw, st);
    return launch_narrow_t<int32_t>(t, in, p, rw, st);
}

cudaError_t launch_narrow_debug(const Tables& t, const Inputs& in, const EstParams& p, int v, int k, uint32_t* row_costs,
                                uint8_t* filter_id, uint32_t* hist, unsigned long long* match, double* size_out, cudaStream_t st) {
    DebugOut d{v, k, row_costs, filter_id, hist, match, size_out};
    if (in.label_dtype == ICR_LABELS_I64) narrow_debug_kernel<int64_t><<<1, 32, 0, st>>>(t, in, p, d);
    else narrow_debug_kernel<int32_t><<<1, 32, 0, st>>>(t, in, p, d);
    count_launch();
    return cudaGetLastError();
}

}  // namespace icr
