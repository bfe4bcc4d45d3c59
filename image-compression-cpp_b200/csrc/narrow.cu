// narrow: Rules C, P, F, R, H, Z for every segment whose bbox is at most NARROW_MAX_W pixels wide and that needs no match
// statistics, one warp per segment, the whole batch in ONE launch.  Replaces the reference's 13 stream operations per
// segment (src/png_size_estimator.cu:505-590), its two nested host loops (src/compute_rewards.cu:125,159) and its RGBA
// staging kernel (:23-53).
//
// Register path (bboxes up to 64 wide, the bulk of the pixels): no shared-memory tile, no second walk.  A lane owns one
// COLUMN of every 32-pixel chunk of the bbox (one or two chunks per row) and the warp walks down the rows.  The three f32 planes (and the raw label unless the segment
// fills its bbox) arrive by plain coalesced loads one or two rows ahead; conversion to a masked RGBA word (Rules C, P)
// happens in registers.  The neighbours of a pixel never touch memory: `up` and `upleft` are the lane's own pixel and
// left neighbour of the row before, `left` is one shuffle.  Segments no wider than 16 pixels put two rows side by side
// in the warp; the row above then comes from the lanes w to the left by one more shuffle.
// Per row: five filter costs per lane (packed bytes: VABSDIFF4 for |cur - pred|, a signed VABSDIFF4 against zero for the
// sum of |int8|; Paeth on half2 pairs entirely on the FMA pipe), five warp sums (REDUX) that deliver the arg-min keys
// cost * 8 + id directly (Rule F, ties to the lowest id), and right away the residual of the chosen filter from the
// operands still in registers -> four unconditional shared-memory atomics into the warp's 4x256 histogram (Rules R, H).
// Lanes outside the bbox are steered to bin 0 by their byte-permute selector and subtracted again in the epilogue, so no
// atomic is predicated.
// Wider bboxes (65..NARROW_MAX_W) take the same arithmetic chunk by chunk with the rows held in three small
// shared-memory row buffers instead of registers (wide_rows): any width, one code path, few registers.
// Epilogue: c*log2(c) table entropy, Rule Z, one f64 store (Rule W is the tail kernel's, big_segment.cu).  Match statistics are zero here by construction: Rule M segments and bboxes wider than
// NARROW_MAX_W go to the big-segment kernel.
#include "icr_internal.h"
#include "icr_pixel.cuh"
#include "icr_reward.cuh"

namespace icr {

#ifndef NRW_DEPTH
#define NRW_DEPTH 2                       // rows in flight ahead of the row being filtered (register path)
#endif
#ifndef NRW_MIN_CTAS
#define NRW_MIN_CTAS 6                    // resident CTAs per SM the register allocation aims at
#endif
constexpr int NRW_WARPS = 4;              // warps per CTA; warps are independent
constexpr uint32_t ZH = 0x64006400u;      // half2 {1024, 1024}: the biased form of two zero bytes

__device__ __noinline__ double nrw_clog2_slow(uint32_t c) { return (double)c * log2((double)c); }
__device__ __forceinline__ double nrw_clog2(uint32_t c, const double* __restrict__ tab) {
    return c < (uint32_t)CLOG2_N ? __ldg(tab + c) : nrw_clog2_slow(c);
}

template <typename LT, int C>
struct RawRow { float r[C], g[C], b[C]; LT l[C]; };

// One segment, one warp.  img / lab point at the planes of the segment's label map; hist is the warp's zeroed 4x256 table.
//   C       32-pixel chunks per row
//   MULTI   two rows per step (w <= 16, C = 1)
//   FULL    the segment fills its bbox: no label is read
//   ADAPT   adaptive filter choice (else every row is Paeth, :545)
// With C <= 2 the row above is carried in all three forms a pixel is used in (packed bytes, two biased half2 words);
// with more chunks only the packed words are carried and unpacked again, which keeps the register count of the whole
// kernel where the narrow rows want it.
template <typename LT, int C, bool MULTI, bool FULL, bool ADAPT, bool DEBUG>
__device__ __forceinline__ void narrow_rows(const float* __restrict__ img, const LT* __restrict__ lab, uint32_t HW, int W, LT key,
                                            int x0, int y0, int w, int h, uint32_t* hist, int lane, const DebugOut& dbg,
                                            uint32_t& slots) {
    static_assert(!(MULTI && C != 1), "rows side by side only for single-chunk widths");
    constexpr bool CARRY_FORMS = !MULTI && C <= 2;
    constexpr int DEPTH = NRW_DEPTH;
    constexpr int R = MULTI ? 2 : 1;
    // ---- lane geometry.  MULTI: row r of a step in lanes [r * w, (r + 1) * w).  (More rows would fit the narrowest
    // bboxes, but a masked warp reduction runs once per group: two rows keep the five row sums of both rows in the
    // 16-bit halves of five plain REDUX.)
    int r = 0, x = lane;
    if (MULTI) {
        r = lane >= w ? 1 : 0;
        x = lane - r * w;
    }
    bool active[C];
#pragma unroll
    for (int c = 0; c < C; ++c) active[c] = MULTI ? lane < 2 * w : lane + 32 * c < w;
    const int from = MULTI ? (active[0] ? (r ? lane - w : lane + w) : lane) : lane;
    const bool give_prev = MULTI && r == 1;
    const uint32_t field_mul = !MULTI ? 1u : (active[0] ? (r ? 0x10000u : 1u) : 0u);     // a row's sums live in its 16-bit half
    const uint32_t field_sel = (MULTI && r) ? 0x4432u : 0x4410u;
    const uint32_t key_add = lane == 0 ? 1u : 0u;        // the lane that adds the filter ids to the cost keys
    // Per-lane constants of a chunk, all functions of "this lane's column lies inside the bbox" (act = 1 / 0):
    //   lmul      the left neighbour exists and counts (act, and not the bbox's first column)
    //   zh_alpha  upper half2 word of a pixel with blue = 0: alpha 255 inside the bbox, 0 outside
    //   sel_a/b   byte 0 / byte 2 of a residual word for a lane inside the bbox, a zero byte (bin 0) else
    // Rows of one or two chunks keep them in registers; wider rows derive them from act by a multiply-add each time.
    constexpr bool KEEP_CONSTS = C <= 2;
    uint32_t act[C], lmul_[C], zh_alpha_[C], sel_a_[C], sel_b_[C];
#pragma unroll
    for (int c = 0; c < C; ++c) {
        act[c] = active[c] ? 1u : 0u;
        lmul_[c] = (active[c] && (c > 0 || x > 0)) ? 1u : 0u;
        zh_alpha_[c] = active[c] ? 0x64ff6400u : ZH;
        sel_a_[c] = active[c] ? 0x4440u : 0x4444u;
        sel_b_[c] = active[c] ? 0x4442u : 0x4444u;
    }
    const uint32_t lmul0 = lmul_[0];
    auto lmul = [&](int c) { return KEEP_CONSTS ? lmul_[c] : (c == 0 ? lmul0 : act[c]); };
    auto zh_alpha = [&](int c) { return KEEP_CONSTS ? zh_alpha_[c] : act[c] * 0x00ff0000u + ZH; };
    auto sel_a = [&](int c) { return KEEP_CONSTS ? sel_a_[c] : 0x4444u - 4u * act[c]; };
    auto sel_b = [&](int c) { return KEEP_CONSTS ? sel_b_[c] : 0x4444u - 2u * act[c]; };
    const int nsteps = (h + R - 1) / R;
    slots = (uint32_t)nsteps * 32u * (uint32_t)C;

    // ---- loads: 32-bit pixel offsets from the map's planes, one wide multiply-add per load
    const float* imgg = img + HW;
    const float* imgb = imgg + HW;
    uint32_t off_pf = (uint32_t)((y0 + r) * W + x0 + x);         // chunk 0 of the next row to be requested
    int row_pf = r;
    const uint32_t RW = (uint32_t)(R * W);

    RawRow<LT, C> raw[DEPTH];
#pragma unroll
    for (int d = 0; d < DEPTH; ++d)
#pragma unroll
        for (int c = 0; c < C; ++c) { raw[d].r[c] = 0.f; raw[d].g[c] = 0.f; raw[d].b[c] = 0.f; raw[d].l[c] = key; }

    auto issue = [&](RawRow<LT, C>& q) {
        const bool rowok = row_pf < h;
#pragma unroll
        for (int c = 0; c < C; ++c)
            if (active[c] && rowok) {
                q.r[c] = __ldg(img + off_pf + 32 * c); q.g[c] = __ldg(imgg + off_pf + 32 * c); q.b[c] = __ldg(imgb + off_pf + 32 * c);
                if (!FULL) q.l[c] = __ldg(lab + off_pf + 32 * c);
            }
        off_pf += RW; row_pf += R;
    };
#pragma unroll
    for (int d = 0; d < DEPTH; ++d) issue(raw[d]);

    // ---- state carried from row to row: the pixels of the row above and their left neighbours
    uint32_t up_p[C], ulp[C], up_lo[C], up_hi[C], ul_lo[C], ul_hi[C];
#pragma unroll
    for (int c = 0; c < C; ++c) { up_p[c] = 0u; ulp[c] = 0u; up_lo[c] = ZH; up_hi[c] = ZH; ul_lo[c] = ZH; ul_hi[c] = ZH; }

    auto step = [&](RawRow<LT, C>& q, int yrow, bool refill) {
        const int y = yrow + r;
        uint32_t cur_p[C], cur_lo[C], cur_hi[C], lf_p[C], av_p[C], pa_p[C];
        uint32_t lf_lo[C], lf_hi[C], pa_lo[C], pa_hi[C];             // kept across the filter choice only with C <= 2
        uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0;
        // Rule C + Rule P: masked RGBA words of this row, straight in the three forms the filters want: the two biased
        // half2 words by multiply-adds on the converted bytes (FMA pipe), the packed word by one byte permute of those
#pragma unroll
        for (int c = 0; c < C; ++c) {
            uint32_t lo = byte_of(q.g[c]) * 0x10000u + (byte_of(q.r[c]) + ZH);
            uint32_t hi = byte_of(q.b[c]) + zh_alpha(c);
            if (!FULL) { const bool in = q.l[c] == key; lo = in ? lo : ZH; hi = in ? hi : ZH; }
            cur_lo[c] = lo; cur_hi[c] = hi;
            cur_p[c] = __byte_perm(lo, hi, 0x6420);
        }
        // the buffer is free again: its refill (the row DEPTH steps ahead) flies under the arithmetic of this row
        if (refill) issue(q);
#pragma unroll
        for (int c = 0; c < C; ++c) {
            // left neighbour: one shuffle; the first column and the lanes outside the bbox see zero
            uint32_t l;
            if (c == 0) l = __shfl_up_sync(0xffffffffu, cur_p[0], 1);
            else l = __shfl_sync(0xffffffffu, lane == 31 ? cur_p[c > 0 ? c - 1 : 0] : cur_p[c], (lane + 31) & 31);
            l *= lmul(c);
            lf_p[c] = l;
            const uint32_t l_lo = unpack_lo(l), l_hi = unpack_hi(l);
            uint32_t u_lo, u_hi, q_lo, q_hi;                 // row above and its left neighbour as half2 words
            if (MULTI) {
                // rows side by side: the row above sits w lanes to the left, the first row's in the second half of the step before
                const uint32_t u = __shfl_sync(0xffffffffu, give_prev ? up_p[0] : cur_p[0], from);
                const uint32_t ul = __shfl_sync(0xffffffffu, give_prev ? ulp[0] : l, from);
                up_p[0] = u;
                u_lo = unpack_lo(u); u_hi = unpack_hi(u); q_lo = unpack_lo(ul); q_hi = unpack_hi(ul);
            } else if (CARRY_FORMS) {
                u_lo = up_lo[c]; u_hi = up_hi[c]; q_lo = ul_lo[c]; q_hi = ul_hi[c];
            } else {
                u_lo = unpack_lo(up_p[c]); u_hi = unpack_hi(up_p[c]); q_lo = unpack_lo(ulp[c]); q_hi = unpack_hi(ulp[c]);
            }
            // Rule F: the five costs of this lane's pixel
            av_p[c] = avg4(l, up_p[c]);
            const uint32_t p_lo = paeth_h2_fma(l_lo, u_lo, q_lo), p_hi = paeth_h2_fma(l_hi, u_hi, q_hi);
            pa_p[c] = __byte_perm(p_lo, p_hi, 0x6420);
            if (ADAPT || DEBUG) {
                c0 = cost_acc(cur_p[c], c0);
                c1 = cost_acc(absdiff4(cur_p[c], l), c1);
                c2 = cost_acc(absdiff4(cur_p[c], up_p[c]), c2);
                c3 = cost_acc(absdiff4(cur_p[c], av_p[c]), c3);
                c4 = cost_acc(absdiff4(cur_p[c], pa_p[c]), c4);
            }
            lf_lo[c] = l_lo; lf_hi[c] = l_hi; pa_lo[c] = p_lo; pa_hi[c] = p_hi;
            if (MULTI) { up_lo[0] = u_lo; up_hi[0] = u_hi; }            // read by the Up residual below
        }
        int f = 4;
        if (ADAPT || DEBUG) {
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
            if (ADAPT) f = (int)(min(min(min(k0, k1), min(k2, k3)), k4) & 7u);
            c0 = k0 >> 3; c1 = k1 >> 3; c2 = k2 >> 3; c3 = k3 >> 3; c4 = k4 >> 3;       // read by the debug tables only
        }
        const bool rowok = !MULTI || y < h;
        if (DEBUG && x == 0 && active[0] && rowok) {
            if (dbg.row_costs && ADAPT) {
                uint32_t* rc = dbg.row_costs + 5 * y;
                rc[0] = c0; rc[1] = c1; rc[2] = c2; rc[3] = c3; rc[4] = c4;
            }
            if (dbg.filter_id) dbg.filter_id[y] = (uint8_t)f;
        }
        // Rules R, H: residual bytes of the chosen filter, as the low bytes of two half2 words, into the histogram.
        // Pixels outside the bbox (and rows past its end) select a zero byte: they land in bin 0 and are subtracted later.
        // Tested in the order the filters win on natural images.
        constexpr bool KEPT = C <= 2;          // the predictors' half2 forms are still in registers
        uint32_t rlo[C], rhi[C];
        if (f == 2) {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const uint32_t a = (KEPT) ? up_lo[c] : unpack_lo(up_p[c]), b = (KEPT) ? up_hi[c] : unpack_hi(up_p[c]);
                rlo[c] = residual_h2(cur_lo[c], a); rhi[c] = residual_h2(cur_hi[c], b);
            }
        } else if (f == 4) {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const uint32_t a = (KEPT) ? pa_lo[c] : unpack_lo(pa_p[c]), b = (KEPT) ? pa_hi[c] : unpack_hi(pa_p[c]);
                rlo[c] = residual_h2(cur_lo[c], a); rhi[c] = residual_h2(cur_hi[c], b);
            }
        } else if (f == 1) {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const uint32_t a = (KEPT) ? lf_lo[c] : unpack_lo(lf_p[c]), b = (KEPT) ? lf_hi[c] : unpack_hi(lf_p[c]);
                rlo[c] = residual_h2(cur_lo[c], a); rhi[c] = residual_h2(cur_hi[c], b);
            }
        } else if (f == 0) {
#pragma unroll
            for (int c = 0; c < C; ++c) { rlo[c] = cur_lo[c]; rhi[c] = cur_hi[c]; }
        } else {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                rlo[c] = residual_h2(cur_lo[c], unpack_lo(av_p[c])); rhi[c] = residual_h2(cur_hi[c], unpack_hi(av_p[c]));
            }
        }
#pragma unroll
        for (int c = 0; c < C; ++c) {
            uint32_t sa = sel_a(c), sb = sel_b(c);
            if (MULTI && !rowok) { sa = 0x4444u; sb = 0x4444u; }
            atomicAdd(hist + __byte_perm(rlo[c], 0u, sa), 1u);
            atomicAdd(hist + 256 + __byte_perm(rlo[c], 0u, sb), 1u);
            atomicAdd(hist + 512 + __byte_perm(rhi[c], 0u, sa), 1u);
            atomicAdd(hist + 768 + __byte_perm(rhi[c], 0u, sb), 1u);
        }
        // this row becomes the row above
#pragma unroll
        for (int c = 0; c < C; ++c) {
            up_p[c] = cur_p[c]; ulp[c] = lf_p[c];
            if (CARRY_FORMS) { up_lo[c] = cur_lo[c]; up_hi[c] = cur_hi[c]; ul_lo[c] = lf_lo[c]; ul_hi[c] = lf_hi[c]; }
        }
    };

    // DEPTH steps per trip, so that the rows in flight rotate through DEPTH register sets without moves; the trip of the
    // last step leaves in the middle (no extra copy of the step: the code of all variants has to stay cache-sized)
    for (int ys = 0; ys < h; ys += DEPTH * R) {
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) {
            if (d > 0 && ys + d * R >= h) break;
            step(raw[d], ys + d * R, true);
        }
    }
}

// Bboxes wider than 64 pixels: the same arithmetic, a 32-pixel chunk at a time, with the rows in shared memory instead of
// registers, so that one loop serves any width.  rows = three buffers of NARROW_MAX_W + 1 words: the row above, this row,
// and this row's Paeth predictors; word 0 of the first two is the zero left of the bbox.  Per row: pass 1 converts the
// chunks (requested one chunk ahead), writes the row, and adds up the five costs (left neighbour by shuffle, the row above
// from its buffer); REDUX + arg-min; pass 2 reads pixel and predictor of the chosen filter back and feeds the histogram.
template <typename LT, bool ADAPT, bool DEBUG>
__device__ __forceinline__ void wide_rows(const float* __restrict__ img, const LT* __restrict__ lab, uint32_t HW, int W, LT key, bool full,
                                          int x0, int y0, int w, int h, uint32_t* hist, uint32_t* rows, int lane, const DebugOut& dbg,
                                          uint32_t& slots) {
    constexpr int RB = NARROW_MAX_W + 1;
    const int C = (w + 31) >> 5;
    slots = (uint32_t)h * 32u * (uint32_t)C;
    uint32_t* above = rows + 1;              // [-1] is the zero column
    uint32_t* here = rows + RB + 1;
    uint32_t* paeth = rows + 2 * RB;
    for (int i = lane; i < 2 * RB; i += 32) rows[i] = 0u;          // zeros above the first row and left of the bbox
    __syncwarp();
    const float* imgg = img + HW;
    const float* imgb = imgg + HW;
    const uint32_t key_add = lane == 0 ? 1u : 0u;
    // chunks are requested one ahead, in row-major chunk order
    uint32_t off = (uint32_t)(y0 * W + x0 + lane);
    float nr = 0.f, ng = 0.f, nb = 0.f;
    LT nl = key;
    int pc = 0;                              // chunk index of the request in flight within its row
    auto request = [&]() {
        const bool in = lane + 32 * pc < w;
        if (in) {
            nr = __ldg(img + off + 32 * pc); ng = __ldg(imgg + off + 32 * pc); nb = __ldg(imgb + off + 32 * pc);
            if (!full) nl = __ldg(lab + off + 32 * pc);
        }
        if (++pc == C) { pc = 0; off += (uint32_t)W; }
    };
    request();
    for (int y = 0; y < h; ++y) {
        uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0;
        uint32_t carry = 0u;                 // the pixel left of this chunk's first column
        for (int c = 0; c < C; ++c) {
            const int x = lane + 32 * c;
            const bool in = x < w;
            uint32_t cur = 0u;
            if (in) {
                cur = __byte_perm(byte_of(ng) * 0x10000u + (byte_of(nr) + ZH), byte_of(nb) + 0x64ff6400u, 0x6420);
                if (!full && nl != key) cur = 0u;
            }
            if (y + 1 < h || c + 1 < C) request();
            uint32_t l = __shfl_up_sync(0xffffffffu, cur, 1);
            if (lane == 0) l = carry;
            carry = __shfl_sync(0xffffffffu, cur, 31);
            if (!in) l = 0u;                 // lanes right of the bbox add nothing to the costs
            const int xi = in ? x : w;       // they read and write the spare word behind the row
            const uint32_t up = in ? above[xi] : 0u, ul = in ? above[xi - 1] : 0u;
            const uint32_t av = avg4(l, up);
            const uint32_t pa = __byte_perm(paeth_h2_fma(unpack_lo(l), unpack_lo(up), unpack_lo(ul)),
                                            paeth_h2_fma(unpack_hi(l), unpack_hi(up), unpack_hi(ul)), 0x6420);
            if (in) { here[x] = cur; paeth[x] = pa; }
            if (ADAPT || DEBUG) {
                c0 = cost_acc(cur, c0);
                c1 = cost_acc(absdiff4(cur, l), c1);
                c2 = cost_acc(absdiff4(cur, up), c2);
                c3 = cost_acc(absdiff4(cur, av), c3);
                c4 = cost_acc(absdiff4(cur, pa), c4);
            }
        }
        int f = 4;
        if (ADAPT || DEBUG) {
            const uint32_t k0 = __reduce_add_sync(0xffffffffu, c0 * 8u), k1 = __reduce_add_sync(0xffffffffu, c1 * 8u + key_add),
                           k2 = __reduce_add_sync(0xffffffffu, c2 * 8u + 2u * key_add),
                           k3 = __reduce_add_sync(0xffffffffu, c3 * 8u + 3u * key_add),
                           k4 = __reduce_add_sync(0xffffffffu, c4 * 8u + 4u * key_add);
            if (ADAPT) f = (int)(min(min(min(k0, k1), min(k2, k3)), k4) & 7u);
            if (DEBUG && lane == 0) {
                if (dbg.row_costs && ADAPT) {
                    uint32_t* rc = dbg.row_costs + 5 * y;
                    rc[0] = k0 >> 3; rc[1] = k1 >> 3; rc[2] = k2 >> 3; rc[3] = k3 >> 3; rc[4] = k4 >> 3;
                }
                if (dbg.filter_id) dbg.filter_id[y] = (uint8_t)f;
            }
        }
        __syncwarp();                        // the row is complete in shared memory
        for (int c = 0; c < C; ++c) {
            const int x = lane + 32 * c;
            const bool in = x < w;
            const int xi = in ? x : 0;
            const uint32_t cur = here[xi];
            uint32_t pred;
            if (f == 2) pred = above[xi];
            else if (f == 4) pred = paeth[xi];
            else if (f == 1) pred = here[xi - 1];
            else if (f == 0) pred = 0u;
            else pred = avg4(here[xi - 1], above[xi]);
            const uint32_t rlo = residual_h2(unpack_lo(cur), unpack_lo(pred)), rhi = residual_h2(unpack_hi(cur), unpack_hi(pred));
            const uint32_t sa = in ? 0x4440u : 0x4444u, sb = in ? 0x4442u : 0x4444u;
            atomicAdd(hist + __byte_perm(rlo, 0u, sa), 1u);
            atomicAdd(hist + 256 + __byte_perm(rlo, 0u, sb), 1u);
            atomicAdd(hist + 512 + __byte_perm(rhi, 0u, sa), 1u);
            atomicAdd(hist + 768 + __byte_perm(rhi, 0u, sb), 1u);
        }
        __syncwarp();                        // pass 2 is done with both rows before the next row overwrites one of them
        uint32_t* t = above; above = here; here = t;
    }
    // (the buffers' zero columns were never written: word -1 of both rows stays zero)
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

template <typename LT, bool ADAPT, bool DEBUG>
__device__ __forceinline__ double narrow_segment(const float* __restrict__ img, const LT* __restrict__ lab, uint32_t HW, int W, LT key,
                                                 int count, int4 bb, uint32_t* hist, uint32_t* rows, const EstParams& p,
                                                 const double* __restrict__ clog2, int lane, const DebugOut& dbg) {
    const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
    const bool full = count == w * h;                    // the segment fills its bbox: no label has to be read
    uint32_t slots = 0;
#define ICR_ROWS(C, MULTI)                                                                                                        \
    do {                                                                                                                          \
        if (full) narrow_rows<LT, C, MULTI, true, ADAPT, DEBUG>(img, lab, HW, W, key, bb.x, bb.y, w, h, hist, lane, dbg, slots);  \
        else narrow_rows<LT, C, MULTI, false, ADAPT, DEBUG>(img, lab, HW, W, key, bb.x, bb.y, w, h, hist, lane, dbg, slots);      \
    } while (0)
    if (w <= 16) ICR_ROWS(1, true);
    else if (w <= 32) ICR_ROWS(1, false);
    else if (w <= 64) ICR_ROWS(2, false);
    else wide_rows<LT, ADAPT, DEBUG>(img, lab, HW, W, key, full, bb.x, bb.y, w, h, hist, rows, lane, dbg, slots);
#undef ICR_ROWS
    __syncwarp();
    return narrow_epilogue<DEBUG>(hist, w, h, slots, p, clog2, lane, dbg);
}

__device__ __forceinline__ int fetch_add(int* counter) {
    int r;
    asm volatile("atom.global.add.s32 %0, [%1], 1;" : "=r"(r) : "l"(counter) : "memory");
    return r;
}

template <typename LT, bool ADAPT>
__global__ void __launch_bounds__(NRW_WARPS * 32, NRW_MIN_CTAS) narrow_kernel(Tables t, Inputs in, EstParams p, RewardArgs rw) {
    __shared__ __align__(16) uint32_t s_hist[NRW_WARPS][1024];
    __shared__ uint32_t s_rows[NRW_WARPS][3 * (NARROW_MAX_W + 1)];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t* hist = s_hist[warp];
    for (int i = lane; i < 1024; i += 32) hist[i] = 0;
    __syncwarp();
    ICR_GRID_DEP_LAUNCH();                // the tail kernel's CTAs may take the places of this kernel's CTAs as they leave
    ICR_GRID_DEP_WAIT();                  // prep's tables and lists are complete
    const uint32_t HW = (uint32_t)t.H * (uint32_t)t.W;
    // the four work lists back to back, widest bboxes first: list_a front, list_b front, list_b back, list_a back
    const int e0 = t.sched[SC_N_W128], e1 = e0 + t.sched[SC_N_W64], e2 = e1 + t.sched[SC_N_W32], total = e2 + t.sched[SC_N_W16];
    const LT* labels = static_cast<const LT*>(in.labels);
    DebugOut none{};
    // One segment per request; the request for the next segment is issued before the current one is worked on, so its
    // latency is hidden.
    int flat = 0;
    if (lane == 0) flat = fetch_add(&t.sched[SC_CUR_NARROW]);
    flat = __shfl_sync(0xffffffffu, flat, 0);
    while (flat < total) {
        int fut = 0;
        if (lane == 0) fut = fetch_add(&t.sched[SC_CUR_NARROW]);
        const uint2* slot = flat < e0 ? t.list_a + flat
                          : flat < e1 ? t.list_b + (flat - e0)
                          : flat < e2 ? t.list_b + (t.list_cap - 1 - (int64_t)(flat - e1))
                                      : t.list_a + (t.list_cap - 1 - (int64_t)(flat - e2));
        const uint2 e = __ldg(slot);
        const int v = (int)e.x, k = (int)e.y;
        const int64_t row = (int64_t)v * t.Kmax + k;
        const int4 bb = __ldg(t.bboxes + row);
        const LT key = (LT)__ldg(t.keys + row);
        const int count = __ldg(t.counts + row);
        const double sz = narrow_segment<LT, ADAPT, false>(in.images + (int64_t)(v / in.S) * 3 * HW, labels + (int64_t)v * HW, HW, t.W, key,
                                                           count, bb, hist, s_rows[warp], p, t.clog2, lane, none);
        if (lane == 0) t.seg_sizes[row] = sz;
        flat = __shfl_sync(0xffffffffu, fut, 0);
    }
}

template <typename LT>
__global__ void __launch_bounds__(32) narrow_debug_kernel(Tables t, Inputs in, EstParams p, DebugOut dbg) {
    __shared__ __align__(16) uint32_t hist[1024];
    __shared__ uint32_t rows[3 * (NARROW_MAX_W + 1)];
    const int lane = threadIdx.x;
    for (int i = lane; i < 1024; i += 32) hist[i] = 0;
    __syncwarp();
    const uint32_t HW = (uint32_t)t.H * (uint32_t)t.W;
    const int64_t row = (int64_t)dbg.v * t.Kmax + dbg.k;
    const int4 bb = t.bboxes[row];
    const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
    if (w <= 0 || h <= 0 || segment_class(w, h, p.L_min) != CLS_NARROW) return;     // the big-segment debug kernel acts
    const float* img = in.images + (int64_t)(dbg.v / in.S) * 3 * HW;
    const LT* lab = static_cast<const LT*>(in.labels) + (int64_t)dbg.v * HW;
    double sz;
    if (p.adaptive) sz = narrow_segment<LT, true, true>(img, lab, HW, t.W, (LT)t.keys[row], t.counts[row], bb, hist, rows, p, t.clog2, lane, dbg);
    else sz = narrow_segment<LT, false, true>(img, lab, HW, t.W, (LT)t.keys[row], t.counts[row], bb, hist, rows, p, t.clog2, lane, dbg);
    if (lane == 0) {
        if (dbg.match) { dbg.match[0] = 0; dbg.match[1] = 0; dbg.match[2] = 0; }
        if (dbg.size) *dbg.size = sz;
    }
}

template <typename LT, bool ADAPT>
static cudaError_t launch_narrow_t(const Tables& t, const Inputs& in, const EstParams& p, const RewardArgs& rw, cudaStream_t st) {
    static int per_sm[16] = {0};          // resident CTAs per SM, per device
    int dev = 0;
    cudaGetDevice(&dev);
    auto kern = narrow_kernel<LT, ADAPT>;
    int per = (dev >= 0 && dev < 16) ? per_sm[dev] : 0;
    if (!per) {
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, kern, NRW_WARPS * 32, 0);
        if (e != cudaSuccess) return e;
        if (per < 1) per = 1;
        if (dev >= 0 && dev < 16) per_sm[dev] = per;
    }
    const cudaError_t le = launch_dependent(kern, dim3(sm_count() * per), dim3(NRW_WARPS * 32), 0, st, t, in, p, rw);   // persistent: warps pull segments until the list is empty
    if (le != cudaSuccess) return le;
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_narrow(const Tables& t, const Inputs& in, const EstParams& p, const RewardArgs& rw, cudaStream_t st) {
    if (in.label_dtype == ICR_LABELS_I64)
        return p.adaptive ? launch_narrow_t<int64_t, true>(t, in, p, rw, st) : launch_narrow_t<int64_t, false>(t, in, p, rw, st);
    return p.adaptive ? launch_narrow_t<int32_t, true>(t, in, p, rw, st) : launch_narrow_t<int32_t, false>(t, in, p, rw, st);
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
