// big-segment path: one CTA per segment (several CTAs per segment when a small batch has fewer segments than the grid has
// CTAs: SPLIT), any size, any channel count 1..4.
// Used (a) by the batch for the few segments large enough for Rule M match statistics
// (chunk = ceil(N/65536) >= L_min, src/png_size_estimator.cu:416-451), and (b) as the
// pointer-level drop-in for estimate_png_size_masked_segment_to_output (:472-591).
// Banded path (RGBA, chunks of whole pixels, bbox up to BAND_MAX_W wide): a band of rows is fetched once into shared
// memory (aligned quads) and serves the filter choice, the residuals + histogram and the run statistics of the band; chunks
// of one pixel take their run statistics from an 8-entry table inside the residual walk.
// General path: phase 1, a warp per bbox row, lanes along x, packed-byte costs, arg-min -> filter id per row;
// phase 2, the filtered stream goes through shared memory a tile at a time: a lane per pixel computes the residual
// bytes and feeds the histogram, then a thread per reference chunk walks its bytes for the run statistics.
// Entropy and mean follow the reference's reduction trees (:281-328) exactly.
#include "icr_internal.h"
#include <mutex>
#include "icr_pixel.cuh"
#include "icr_reward.cuh"

namespace icr {

constexpr int BIG_T = 256;
constexpr int BIG_TILE_PX = 4096;       // residual words kept in shared memory per tile of the stream
constexpr int BAND_MAX_W = 1024;        // widest bbox of the banded path (its halo row lives in shared memory)
constexpr int BAND_MAX_R = 256;         // most rows of a band

struct BigDebug { uint32_t* row_costs; uint8_t* filter_id; uint32_t* hist; unsigned long long* match; double* size; };

template <typename LT>
struct SrcPlanar {          // the batch layout: the caller's f32 planes + raw labels, masked by the segment's key
    const float *r, *g, *b; const LT* lab; int W; LT key;
    __device__ SrcPlanar(const float* img, const LT* lab_, int64_t HW, int W_, LT key_)
        : r(img), g(img + HW), b(img + 2 * HW), lab(lab_), W(W_), key(key_) {}
    __device__ __forceinline__ uint32_t px(int gy, int gx) const {
        // label and planes are asked for together (no branch on the label): the neighbours of a pixel then cost one memory
        // round trip instead of two dependent ones each.  H*W < 2^31 (api.cu), so the pixel index is a 32-bit number.
        const uint32_t i = (uint32_t)gy * (uint32_t)W + (uint32_t)gx;
        const LT l = __ldg(lab + i);
        const float fr = __ldg(r + i), fg = __ldg(g + i), fb = __ldg(b + i);
        return l == key ? rgba_of(fr, fg, fb) : 0u;                                                  // Rule C, Rule P
    }
    // rows widened to 16-byte boundaries can be fetched as float4 / int4 quads: W % 4 == 0 and aligned planes
    __device__ __forceinline__ bool quads_ok() const {
        return (W & 3) == 0 && (((uintptr_t)r | (uintptr_t)g | (uintptr_t)b | (uintptr_t)lab) & 15) == 0;
    }
    // masked pixels x0 .. x0+w-1 of row gy -> dst[0..w): a lane takes aligned quads, four pixels per load
    __device__ __forceinline__ void row_quads(int gy, int x0, int w, uint32_t* dst, int lane) const {
        const int x0a = x0 & ~3, dx = x0 - x0a, nq = (dx + w + 3) >> 2;
        const uint32_t base = (uint32_t)gy * (uint32_t)W + (uint32_t)x0a;
#pragma unroll 2
        for (int q = lane; q < nq; q += 32) {
            const uint32_t i = base + 4u * (uint32_t)q;
            const float4 fr = __ldg(reinterpret_cast<const float4*>(r + i)), fg = __ldg(reinterpret_cast<const float4*>(g + i)),
                         fb = __ldg(reinterpret_cast<const float4*>(b + i));
            LT l[4];
            if constexpr (sizeof(LT) == 4) {
                const int4 v = __ldg(reinterpret_cast<const int4*>(lab + i));
                l[0] = v.x; l[1] = v.y; l[2] = v.z; l[3] = v.w;
            } else {
                const longlong2 u = __ldg(reinterpret_cast<const longlong2*>(lab + i)), v = __ldg(reinterpret_cast<const longlong2*>(lab + i + 2));
                l[0] = u.x; l[1] = u.y; l[2] = v.x; l[3] = v.y;
            }
            const uint32_t px4[4] = {rgba_of(fr.x, fg.x, fb.x), rgba_of(fr.y, fg.y, fb.y), rgba_of(fr.z, fg.z, fb.z), rgba_of(fr.w, fg.w, fb.w)};
            const int xb = 4 * q - dx;
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if ((unsigned)(xb + e) < (unsigned)w) dst[xb + e] = l[e] == key ? px4[e] : 0u;
        }
    }
};

struct SrcRaw {             // the reference's layout: u8 [H,W,C] + compact int64 labels
    const uint8_t* img; const int64_t* lab; int W, C; int64_t k;
    __device__ __forceinline__ uint32_t px(int gy, int gx) const {
        const int64_t i = (int64_t)gy * W + gx;
        const int64_t l = lab[i];
        uint32_t v = 0;
        for (int c = 0; c < C; ++c) v |= (uint32_t)img[i * C + c] << (8 * c);
        return l == k ? v : 0u;
    }
    __device__ __forceinline__ bool quads_ok() const { return false; }
    __device__ __forceinline__ void row_quads(int, int, int, uint32_t*, int) const {}
};

__device__ __forceinline__ double size_formula_big(double Hbar, unsigned long long ms, unsigned long long mc,
                                                   unsigned long long mls, unsigned long long N, int h, const EstParams& p) {
    double f = (N > 0 && ms > 0) ? (double)ms / (double)N : 0.0;
    double Lbar = mc > 0 ? (double)mls / (double)mc : (double)p.L_min;
    double b_lit = Hbar + p.beta;
    double b_match = p.token / Lbar + p.gamma;
    double b_data = (1.0 - f) * b_lit + f * b_match;
    return (p.overhead + (double)h) + ((double)N * b_data) / 8.0;
}

// shared memory of one CTA working on a big segment, declared once per kernel so that the code variants of block_segment
// share it
struct BigSmem {
    uint32_t hist[4 * 256];
    uint32_t tile[BIG_TILE_PX];         // residual words of a band / tile (chunks of more than one pixel)
    uint32_t pix[BIG_TILE_PX];          // masked pixels of a band
    uint32_t halo[BAND_MAX_W];          // the row above the band
    double red[256];
    double Hc[4];
    unsigned long long ms, mc, mls;
    uint32_t runs[8];
    int ticket;
    uint8_t filt[BAND_MAX_R];
};

// whole CTA cooperates on one segment; filt = >= h bytes of scratch (global).
// part / nparts / mg: the bands of the banded path can be shared out among up to nparts CTAs (a batch with fewer big
// segments than resident CTAs would otherwise leave most of the machine idle for the 270 us one CTA needs for a 224x224
// bbox).  Every part adds its histogram and match statistics to the segment's merge slot; the part that arrives last reads
// them back (clearing the slot) and finishes the segment.  Integer sums only, so the result does not depend on the split.
// Returns true in the CTA that wrote *out.
// SPLIT = false is the one-CTA code of before, compiled without any of the sharing (it is the path of large batches).
template <class Src, bool SPLIT = false>
__device__ bool block_segment(const Src& src, int x0, int y0, int w, int h, int C, const EstParams& p,
                              uint8_t* filt, double* out, const BigDebug& dbg, BigSmem& sh, int part = 0, int nparts = 1, BigMerge* mg = nullptr) {
    uint32_t* const s_hist = sh.hist;
    uint32_t* const s_tile = sh.tile;
    double* const s_red = sh.red;
    double* const s_Hc = sh.Hc;
    unsigned long long &s_ms = sh.ms, &s_mc = sh.mc, &s_mls = sh.mls;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    for (int i = tid; i < 4 * 256; i += BIG_T) s_hist[i] = 0;
    if (tid == 0) { s_ms = 0; s_mc = 0; s_mls = 0; }

    const unsigned long long N = (unsigned long long)w * h * C;
    const unsigned long long chunk = (N + 65535ull) / 65536ull;
    const unsigned long long nchunks = (N + chunk - 1) / chunk;
    unsigned long long ms = 0, mc = 0;

    // ---- banded path (RGBA, chunks of whole pixels, bbox no wider than BAND_MAX_W): the masked pixels of a band of rows are
    // fetched ONCE into shared memory and serve the filter choice and the residuals of the band; a band is a whole number
    // of reference chunks, so no chunk straddles two bands.  The general path below reads every pixel four times.
    int R = 0;
    if (C == 4 && (chunk & 3ull) == 0 && w <= BAND_MAX_W && chunk <= (unsigned long long)BIG_TILE_PX) {
        const int wpc = (int)(chunk >> 2);                     // pixels per chunk
        int a = w, b = wpc;
        while (b) { const int t = a % b; a = b; b = t; }       // gcd(w, wpc)
        const int g = wpc / a;                                 // rows per band must be a multiple of g
        R = min(min(h, BIG_TILE_PX / w), BAND_MAX_R);
        if (R < h) {                                           // the last band may end anywhere: it ends the stream
            R = (R / g) * g;
            // the row phases give a row to a warp: a band whose rows are a multiple of the CTA's warps keeps all of them busy
            // (224-wide bbox: 16 rows in two full rounds instead of 18 in three)
            int a2 = g, b2 = BIG_T / 32;
            while (b2) { const int t = a2 % b2; a2 = b2; b2 = t; }
            const int l = g / a2 * (BIG_T / 32);               // lcm(g, warps)
            if (R >= l) R = (R / l) * l;
        }
    }
    int parts = 1;                                             // CTAs that really share this segment
    if (SPLIT) {
        if (R > 0) parts = min(nparts, (h + R - 1) / R);
        if (part >= parts) return false;                       // nothing left for this CTA (uniform over the CTA)
    }
    if (R > 0) {
        int yb_begin = 0, yb_end = h;
        if (SPLIT) {
            const int nb = (h + R - 1) / R;
            yb_begin = (int)((long long)part * nb / parts) * R;
            yb_end = min(h, (int)((long long)(part + 1) * nb / parts) * R);
        }
        uint32_t* const s_pix = sh.pix;
        uint32_t* const s_halo = sh.halo;
        uint8_t* const s_filt = sh.filt;
        const uint32_t wpc = (uint32_t)(chunk >> 2);
        const bool quads = src.quads_ok();
        // Rule M for chunks of ONE pixel (every bbox of 49 153 .. 65 536 pixels, e.g. a 224x224 image): a chunk's runs follow from
        // which of its three byte pairs are equal, so {matched bytes, runs} of the eight cases sit in a table (low / high half
        // of a word) and the residual walk adds its pixel's entry on the spot: no second pass, no residual tile
        uint32_t* const s_runs = sh.runs;
        if (tid < 8) {
            int run = 1, m = 0, n = 0;
            for (int c = 0; c < 3; ++c) {
                if (tid >> c & 1) ++run;
                else { if (run >= p.L_min) { m += run; ++n; } run = 1; }
            }
            if (run >= p.L_min) { m += run; ++n; }
            s_runs[tid] = (uint32_t)m | (uint32_t)n << 16;
        }
        for (int yb = yb_begin; yb < yb_end; yb += R) {
            const int nr = min(R, h - yb);
            const int here = nr * w;
            // the row above the band: zeros above the bbox, else the last row of the band before (always a full band), which
            // the first band of a later part fetches itself
            if (SPLIT && yb == yb_begin && yb > 0) { for (int x = tid; x < w; x += BIG_T) s_halo[x] = src.px(y0 + yb - 1, x0 + x); }
            else { for (int x = tid; x < w; x += BIG_T) s_halo[x] = yb ? s_pix[(R - 1) * w + x] : 0u; }
            __syncthreads();
            // a warp per row, lanes along x (these bboxes are wide: no row/column split per pixel)
            for (int r = warp; r < nr; r += BIG_T / 32) {
                uint32_t* dst = s_pix + r * w;
                ICR_DCHECK(r * w + w <= BIG_TILE_PX && w <= BAND_MAX_W && nr <= BAND_MAX_R);
                if (quads) src.row_quads(y0 + yb + r, x0, w, dst, lane);
                else {
#pragma unroll 4
                    for (int x = lane; x < w; x += 32) dst[x] = src.px(y0 + yb + r, x0 + x);
                }
            }
            __syncthreads();
            // filter per row (Rule F): a warp per row
            for (int r = warp; r < nr; r += BIG_T / 32) {
                int f = 4;
                if (p.adaptive) {
                    Costs5 c = {0, 0, 0, 0, 0};
                    const uint32_t* cr = s_pix + r * w;
                    const uint32_t* ur = r ? cr - w : s_halo;
                    // the band lies in shared memory: the left neighbours are two more loads, cheaper than four shuffles and
                    // the fix-up of lane 0
#pragma unroll 2
                    for (int x = lane; x < w; x += 32) {
                        const uint32_t cur = cr[x], up = ur[x];
                        const uint32_t left = x ? cr[x - 1] : 0u, ul = x ? ur[x - 1] : 0u;
                        add_costs(c, cur, left, up, ul);
                    }
                    for (int o = 16; o; o >>= 1) {
                        c.none += __shfl_xor_sync(0xffffffffu, c.none, o);
                        c.sub += __shfl_xor_sync(0xffffffffu, c.sub, o);
                        c.up += __shfl_xor_sync(0xffffffffu, c.up, o);
                        c.avg += __shfl_xor_sync(0xffffffffu, c.avg, o);
                        c.paeth += __shfl_xor_sync(0xffffffffu, c.paeth, o);
                    }
                    f = pick_filter(c);
                    if (lane == 0 && dbg.row_costs) {
                        uint32_t* rc = dbg.row_costs + 5 * (int64_t)(yb + r);
                        rc[0] = c.none; rc[1] = c.sub; rc[2] = c.up; rc[3] = c.avg; rc[4] = c.paeth;
                    }
                }
                if (lane == 0) { s_filt[r] = (uint8_t)f; if (dbg.filter_id) dbg.filter_id[yb + r] = (uint8_t)f; }
            }
            __syncthreads();
            // residuals of the band in stream order (Rules R, H)
            for (int r = warp; r < nr; r += BIG_T / 32) {
                const uint32_t* cr = s_pix + r * w;
                const uint32_t* ur = r ? cr - w : s_halo;
                uint32_t* rr = s_tile + r * w;
                const int f = s_filt[r];
                uint32_t acc = 0;                              // one-pixel chunks: matched bytes | runs << 16 of this lane's pixels (at most 4 and 2 each)
                for (int x = lane; x < w; x += 32) {
                    const uint32_t cur = cr[x], up = ur[x];
                    const uint32_t left = x ? cr[x - 1] : 0u, ul = x ? ur[x - 1] : 0u;
                    const uint32_t r4 = residual4(f, cur, left, up, ul);
                    if (wpc == 1) {
                        const uint32_t eq = r4 ^ (r4 >> 8);    // byte c is zero when residual bytes c and c+1 are equal
                        acc += s_runs[((eq & 0xffu) ? 0u : 1u) | ((eq & 0xff00u) ? 0u : 2u) | ((eq & 0xff0000u) ? 0u : 4u)];
                    } else rr[x] = r4;
                    atomicAdd(&s_hist[r4 & 255u], 1u);
                    atomicAdd(&s_hist[256 + ((r4 >> 8) & 255u)], 1u);
                    atomicAdd(&s_hist[512 + ((r4 >> 16) & 255u)], 1u);
                    atomicAdd(&s_hist[768 + (r4 >> 24)], 1u);
                }
                ms += acc & 0xffffu; mc += acc >> 16;
            }
            __syncthreads();
            if (wpc == 1) continue;                            // statistics already taken in the residual walk
            // run statistics (Rule M): a thread per reference chunk of wpc whole pixels
            for (uint32_t w0 = (uint32_t)tid * wpc; w0 < (uint32_t)here; w0 += BIG_T * wpc) {
                const uint32_t w1 = min(w0 + wpc, (uint32_t)here);
                int prev = -1, run = 0;
                for (uint32_t wi = w0; wi < w1; ++wi) {
                    const uint32_t r4 = s_tile[wi];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const int bv = (int)((r4 >> (8 * c)) & 255u);
                        if (bv == prev) { ++run; }
                        else {
                            if (prev >= 0 && run >= p.L_min) { ms += (unsigned long long)run; mc += 1; }
                            prev = bv; run = 1;
                        }
                    }
                }
                if (run >= p.L_min) { ms += (unsigned long long)run; mc += 1; }
            }
            __syncthreads();
        }
    } else {
    // ---- phase 1: filter per row (Rule F)
    for (int y = warp; y < h; y += BIG_T / 32) {
        int f = 4;
        if (p.adaptive) {
            Costs5 c = {0, 0, 0, 0, 0};
            uint32_t last_cur = 0u, last_up = 0u;        // lane 31's pixels of the step before: lane 0's neighbours (zero left of the bbox)
#pragma unroll 2
            for (int xb = 0; xb < w; xb += 32) {         // the lane to the left holds this pixel's left and upper-left neighbours
                const int x = xb + lane;
                const uint32_t cur = x < w ? src.px(y0 + y, x0 + x) : 0u;
                const uint32_t up = (x < w && y > 0) ? src.px(y0 + y - 1, x0 + x) : 0u;
                uint32_t left = __shfl_up_sync(0xffffffffu, cur, 1), ul = __shfl_up_sync(0xffffffffu, up, 1);
                if (lane == 0) { left = last_cur; ul = last_up; }
                last_cur = __shfl_sync(0xffffffffu, cur, 31);
                last_up = __shfl_sync(0xffffffffu, up, 31);
                if (x < w) add_costs(c, cur, left, up, ul);
            }
            for (int o = 16; o; o >>= 1) {
                c.none += __shfl_xor_sync(0xffffffffu, c.none, o);
                c.sub += __shfl_xor_sync(0xffffffffu, c.sub, o);
                c.up += __shfl_xor_sync(0xffffffffu, c.up, o);
                c.avg += __shfl_xor_sync(0xffffffffu, c.avg, o);
                c.paeth += __shfl_xor_sync(0xffffffffu, c.paeth, o);
            }
            f = pick_filter(c);
            if (lane == 0 && dbg.row_costs) {
                uint32_t* rc = dbg.row_costs + 5 * (int64_t)y;
                rc[0] = c.none; rc[1] = c.sub; rc[2] = c.up; rc[3] = c.avg; rc[4] = c.paeth;
            }
        }
        if (lane == 0) { filt[y] = (uint8_t)f; if (dbg.filter_id) dbg.filter_id[y] = (uint8_t)f; }
    }
    __syncthreads();

    // ---- phase 2: the filtered stream, tile by tile (Rules R, H, M).  A tile is a whole number of reference chunks and of
    // pixels.  Step A: a lane per pixel in stream order -- coalesced loads, left neighbours by shuffle -- leaves the pixel's
    // residual bytes in shared memory and feeds the histogram.  Step B: a thread per reference chunk walks its bytes there.
    if (chunk <= (unsigned long long)BIG_TILE_PX) {
        const uint32_t n_px = (uint32_t)w * (uint32_t)h, ck = (uint32_t)chunk;
        const uint32_t tile_px = ck * ((uint32_t)BIG_TILE_PX / ck);               // tile bytes = tile_px * C: chunks * C whole chunks
        for (uint32_t p0 = 0; p0 < n_px; p0 += tile_px) {
            const uint32_t here = min(tile_px, n_px - p0);
            // every warp takes a contiguous run of the tile's pixels, 32 per step, so that lane 0's neighbours are lane 31's
            // pixels of the step before; only a warp's first step fetches them
            const uint32_t span = (((here + BIG_T / 32 - 1) / (BIG_T / 32)) + 31u) & ~31u;
            const uint32_t wbeg = (uint32_t)warp * span, wend = min(wbeg + span, here);
            uint32_t last_cur = 0u, last_up = 0u;
#pragma unroll 2
            for (uint32_t base = wbeg; base < wend; base += 32) {
                const uint32_t q = base + lane;
                const bool in = q < here;
                const uint32_t pi = p0 + (in ? q : 0u);
                const int y = (int)(pi / (uint32_t)w), x = (int)(pi - (uint32_t)y * (uint32_t)w);
                const uint32_t cur = in ? src.px(y0 + y, x0 + x) : 0u;
                const uint32_t up = (in && y > 0) ? src.px(y0 + y - 1, x0 + x) : 0u;
                uint32_t left = __shfl_up_sync(0xffffffffu, cur, 1), ul = __shfl_up_sync(0xffffffffu, up, 1);
                if (lane == 0) {
                    if (base != wbeg) { left = last_cur; ul = last_up; }
                    else if (in && x > 0) {
                        left = src.px(y0 + y, x0 + x - 1);
                        ul = y > 0 ? src.px(y0 + y - 1, x0 + x - 1) : 0u;
                    }
                }
                last_cur = __shfl_sync(0xffffffffu, cur, 31);
                last_up = __shfl_sync(0xffffffffu, up, 31);
                if (x == 0) { left = 0u; ul = 0u; }
                if (in) {
                    const uint32_t r4 = residual4(filt[y], cur, left, up, ul);
                    s_tile[q] = r4;
                    for (int c = 0; c < C; ++c) atomicAdd(&s_hist[c * 256 + ((r4 >> (8 * c)) & 255u)], 1u);
                }
            }
            __syncthreads();
            const uint32_t bytes = here * (uint32_t)C;
            if (C == 4 && (ck & 3u) == 0) {               // chunks of whole pixels: four bytes per word, no channel counter
                const uint32_t wpc = ck >> 2;
                for (uint32_t w0 = (uint32_t)tid * wpc; w0 < here; w0 += BIG_T * wpc) {
                    const uint32_t w1 = min(w0 + wpc, here);
                    int prev = -1, run = 0;
                    for (uint32_t wi = w0; wi < w1; ++wi) {
                        const uint32_t r4 = s_tile[wi];
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            const int bv = (int)((r4 >> (8 * c)) & 255u);
                            if (bv == prev) { ++run; }
                            else {
                                if (prev >= 0 && run >= p.L_min) { ms += (unsigned long long)run; mc += 1; }
                                prev = bv; run = 1;
                            }
                        }
                    }
                    if (run >= p.L_min) { ms += (unsigned long long)run; mc += 1; }
                }
            } else
            for (uint32_t start = (uint32_t)tid * ck; start < bytes; start += BIG_T * ck) {
                const uint32_t end = min(start + ck, bytes);
                uint32_t px_i = start / (uint32_t)C;
                int c = (int)(start - px_i * (uint32_t)C);
                uint32_t r4 = s_tile[px_i];
                int prev = -1, run = 0;
                for (uint32_t i = start; i < end; ++i) {
                    const int bv = (int)((r4 >> (8 * c)) & 255u);
                    if (bv == prev) { ++run; }
                    else {
                        if (prev >= 0 && run >= p.L_min) { ms += (unsigned long long)run; mc += 1; }
                        prev = bv; run = 1;
                    }
                    if (++c == C) { c = 0; if (i + 1 < end) r4 = s_tile[++px_i]; }
                }
                if (run >= p.L_min) { ms += (unsigned long long)run; mc += 1; }
            }
            __syncthreads();
        }
    } else {
        // chunks longer than a tile (segments beyond 2^28 bytes): a thread per chunk reads its pixels directly
        for (unsigned long long t = tid; t < nchunks; t += BIG_T) {
            const unsigned long long start = t * chunk;
            const unsigned long long end = min(start + chunk, N);
            const unsigned long long pi0 = start / (unsigned)C;
            int c = (int)(start - pi0 * (unsigned)C);
            int y = (int)(pi0 / (unsigned)w), x = (int)(pi0 - (unsigned long long)y * (unsigned)w);
            bool fresh = true;
            uint32_t r4 = 0;
            int prev = -1, run = 0;
            for (unsigned long long i = start; i < end; ++i) {
                if (fresh) {
                    fresh = false;
                    uint32_t cur = src.px(y0 + y, x0 + x);
                    uint32_t left = x > 0 ? src.px(y0 + y, x0 + x - 1) : 0u;
                    uint32_t up = y > 0 ? src.px(y0 + y - 1, x0 + x) : 0u;
                    uint32_t ul = (x > 0 && y > 0) ? src.px(y0 + y - 1, x0 + x - 1) : 0u;
                    r4 = residual4(filt[y], cur, left, up, ul);
                }
                const int bv = (int)((r4 >> (8 * c)) & 255u);
                atomicAdd(&s_hist[c * 256 + bv], 1u);
                if (bv == prev) { ++run; }
                else {
                    if (prev >= 0 && run >= p.L_min) { ms += (unsigned long long)run; mc += 1; }
                    prev = bv; run = 1;
                }
                if (++c == C) { c = 0; fresh = true; if (++x == w) { x = 0; ++y; } }
            }
            if (run >= p.L_min) { ms += (unsigned long long)run; mc += 1; }
        }
    }
    }   // general path
    if (ms) { atomicAdd(&s_ms, ms); atomicAdd(&s_mls, ms); }    // match_symbols == match_len_sum (:440-442)
    if (mc) atomicAdd(&s_mc, mc);
    __syncthreads();
    if (SPLIT && parts > 1) {
        int& s_ticket = sh.ticket;
        for (int i = tid; i < 4 * 256; i += BIG_T) { const uint32_t u = s_hist[i]; if (u) atomicAdd(&mg->hist[i], u); }
        if (tid == 0) { if (s_ms) atomicAdd(&mg->ms, s_ms); if (s_mc) atomicAdd(&mg->mc, s_mc); }
        __threadfence();
        __syncthreads();
        if (tid == 0) s_ticket = (int)atomicAdd(&mg->done, 1u);
        __syncthreads();
        if (s_ticket != parts - 1) return false;               // another part will finish the segment
        __threadfence();
        for (int i = tid; i < 4 * 256; i += BIG_T) s_hist[i] = atomicExch(&mg->hist[i], 0u);     // read past L1 and leave the slot clear
        if (tid == 0) { s_ms = atomicExch(&mg->ms, 0ull); s_mls = s_ms; s_mc = atomicExch(&mg->mc, 0ull); mg->done = 0u; }
        __syncthreads();
    }

    // ---- entropy per channel with the reference's 256-leaf tree (:281-309)
    const uint32_t n = (uint32_t)((unsigned long long)w * (unsigned long long)h);
    for (int c = 0; c < C; ++c) {
        double term = 0.0;
        uint32_t hv = s_hist[c * 256 + tid];
        if (dbg.hist) dbg.hist[c * 256 + tid] = hv;
        if (hv) { double q = (double)hv / (double)n; term = -q * log2(q); }
        s_red[tid] = term;
        __syncthreads();
        for (int st = 128; st > 0; st >>= 1) { if (tid < st) s_red[tid] += s_red[tid + st]; __syncthreads(); }
        if (tid == 0) s_Hc[c] = s_red[0];
        __syncthreads();
    }
    if (tid == 0) {
        double H[4] = {0, 0, 0, 0};
        for (int c = 0; c < C; ++c) H[c] = s_Hc[c];
        const double Hbar = ((H[0] + H[2]) + (H[1] + H[3])) / (double)C;   // tree of :314-328
        const double sz = size_formula_big(Hbar, s_ms, s_mc, s_mls, N, h, p);
        *out = sz;
        if (dbg.size) *dbg.size = sz;
        if (dbg.match) { dbg.match[0] = s_ms; dbg.match[1] = s_mc; dbg.match[2] = s_mls; }
    }
    __syncthreads();
    return true;
}

// Persistent CTAs pull the segments prep put on list_big (Rule M can fire, or the bbox is wider or larger than the
// register path takes), one segment per CTA at a time.  This is always the last kernel of a batch call: whichever
// CTA finishes last resets the work lists for the next call and, in a sharded run, tells the peers that this rank's
// rewards have landed.
// SPLIT: the launch has so few label maps that big segments may be shared out among CTAs (decided here from the list's
// length); large batches run the kernel compiled without that code.
template <typename LT, bool SPLIT>
__global__ void __launch_bounds__(BIG_T, 4) big_kernel(Tables t, Inputs in, EstParams p, RewardArgs rw) {
    __shared__ int s_next, s_last;
    __shared__ BigSmem sh;
    ICR_GRID_DEP_WAIT();                  // every segment size of the estimator is final
    const int64_t HW = (int64_t)t.H * t.W;
    BigDebug none{};
    const int total = t.sched[SC_N_BIG];
    // fewer big segments than half the grid: P CTAs per segment (work item = segment * P + part)
    int P = 1;
    if (SPLIT && total > 0 && total <= t.merge_cap) P = max(1, min((int)gridDim.x / total, BIG_MAX_PARTS));
    const int items = total * P;
    uint8_t* filt = t.big_filter + (int64_t)blockIdx.x * t.H;
    // Rule W for the maps whose sizes all came from the kernel before: a warp per map
    for (int v = blockIdx.x * (BIG_T / 32) + (threadIdx.x >> 5); v < t.V; v += gridDim.x * (BIG_T / 32))
        if (t.status[v] == ST_OK && t.nbig[v] == 0) reward_of_map<false>(t, p, rw.image_sizes, rw.S, rw.rewards, v, threadIdx.x & 31);
    for (;;) {
        if (threadIdx.x == 0) s_next = total > 0 ? atomicAdd(&t.sched[SC_CUR_BIG], 1) : 0;
        __syncthreads();
        if (s_next >= items) break;
        const int i = s_next / P, part = s_next - i * P;
        const uint2 e = t.list_big[i];
        const int v = (int)e.x, k = (int)e.y;
        const int64_t row = (int64_t)v * t.Kmax + k;
        const int4 bb = t.bboxes[row];
        SrcPlanar<LT> src(in.images + (int64_t)(v / in.S) * 3 * HW, static_cast<const LT*>(in.labels) + (int64_t)v * HW, HW, t.W,
                          (LT)t.keys[row]);
        const bool wrote = block_segment<SrcPlanar<LT>, SPLIT>(src, bb.x, bb.y, bb.z - bb.x + 1, bb.w - bb.y + 1, 4, p, filt, t.seg_sizes + row,
                                                               none, sh, part, P, SPLIT ? t.big_merge + i : nullptr);
        if (threadIdx.x == 0) s_last = wrote ? segment_done(t, v, t.nbig[v]) : 0;     // the thread that stored the size
        __syncthreads();
        if (s_last && threadIdx.x < 32) {                 // every size of this map is final: Rule W
            __threadfence();
            reward_of_map<true>(t, p, rw.image_sizes, rw.S, rw.rewards, v, threadIdx.x);
        }
        __syncthreads();                                  // s_next / s_last are reused
    }
    // ---- end of the batch call
    __syncthreads();
    if (threadIdx.x != 0) return;
    if (t.peer_flags) __threadfence_system();             // this CTA's peer stores, if it wrote rewards
    if (atomicAdd(&t.sched[SC_TAIL_DONE], 1) != (int)gridDim.x - 1) return;
    for (int c = 0; c < SCHED_N; ++c) t.sched[c] = 0;
    if (t.peer_flags) {
        __threadfence_system();
        for (int g = 0; g < t.peer_count; ++g)
            asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(t.peer_flags[g] + t.peer_rank), "l"(t.peer_step) : "memory");
    }
}

__device__ __forceinline__ bool flagged(int4 bb, int L_min) { return segment_class(bb.z - bb.x + 1, bb.w - bb.y + 1, L_min) == CLS_BIG; }

template <typename LT>
__global__ void __launch_bounds__(BIG_T) big_debug_kernel(Tables t, Inputs in, EstParams p, int v, int k, BigDebug dbg) {
    const int4 bb = t.bboxes[(int64_t)v * t.Kmax + k];
    if (!flagged(bb, p.L_min)) return;          // the warp path's debug kernel handles it
    __shared__ double scratch;
    const int64_t HW = (int64_t)t.H * t.W;
    SrcPlanar<LT> src(in.images + (int64_t)(v / in.S) * 3 * HW, static_cast<const LT*>(in.labels) + (int64_t)v * HW, HW, t.W,
                      (LT)t.keys[(int64_t)v * t.Kmax + k]);
    __shared__ BigSmem sh;
    block_segment(src, bb.x, bb.y, bb.z - bb.x + 1, bb.w - bb.y + 1, 4, p, t.big_filter, &scratch, dbg, sh);
}

__global__ void __launch_bounds__(BIG_T) single_segment_kernel(SrcRaw src, int x0, int y0, int w, int h, EstParams p,
                                                                uint8_t* filt, double* out, BigDebug dbg) {
    __shared__ BigSmem sh;
    block_segment(src, x0, y0, w, h, src.C, p, filt, out, dbg, sh);
}

cudaError_t launch_big(const Tables& t, const Inputs& in, const EstParams& p, const RewardArgs& rw, cudaStream_t st) {
    int grid = sm_count() * 4;
    if (grid > BIG_GRID_MAX) grid = BIG_GRID_MAX;
    // a big segment can only get a second CTA when the grid holds two per segment, and every map may bring one
    const bool split = 2 * t.V <= grid;
    if (in.label_dtype == ICR_LABELS_I64) {
        if (split) launch_dependent(big_kernel<int64_t, true>, dim3(grid), dim3(BIG_T), 0, st, t, in, p, rw);
        else launch_dependent(big_kernel<int64_t, false>, dim3(grid), dim3(BIG_T), 0, st, t, in, p, rw);
    } else {
        if (split) launch_dependent(big_kernel<int32_t, true>, dim3(grid), dim3(BIG_T), 0, st, t, in, p, rw);
        else launch_dependent(big_kernel<int32_t, false>, dim3(grid), dim3(BIG_T), 0, st, t, in, p, rw);
    }
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_big_debug(const Tables& t, const Inputs& in, const EstParams& p, int v, int k, uint32_t* row_costs,
                             uint8_t* filter_id, uint32_t* hist, unsigned long long* match, double* size_out, cudaStream_t st) {
    BigDebug d{row_costs, filter_id, hist, match, size_out};
    if (in.label_dtype == ICR_LABELS_I64) big_debug_kernel<int64_t><<<1, BIG_T, 0, st>>>(t, in, p, v, k, d);
    else big_debug_kernel<int32_t><<<1, BIG_T, 0, st>>>(t, in, p, v, k, d);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_single_segment(const uint8_t* img, int W, int H, int channels, const int64_t* labels, int64_t k,
                                  int x0, int y0, int w, int h, const EstParams& p, double* out, uint32_t* row_costs,
                                  uint8_t* filter_id, uint32_t* hist, unsigned long long* match, uint8_t* filter_scratch,
                                  cudaStream_t st) {
    (void)H;
    SrcRaw src{img, labels, W, channels, k};
    BigDebug d{row_costs, filter_id, hist, match, nullptr};
    single_segment_kernel<<<1, BIG_T, 0, st>>>(src, x0, y0, w, h, p, filter_scratch, out, d);
    count_launch();
    return cudaGetLastError();
}

// ---- the same launch for callers that may not allocate (CUDA graph capture, serving loops): the h bytes of filter-id
// scratch live in a file-scope device array that every GPU loading the library holds until exit.  All calls on a device
// share it: block_segment writes filt[y] of every row before it reads it, so nothing has to be cleared between calls,
// but two calls that overlap on the GPU (different streams) race on it.
static __device__ __align__(256) uint8_t g_noalloc_filter[NOALLOC_MAX_ROWS];

// device address of the array on the current device: in host code the symbol is only a handle, the address comes
// from the runtime, once per device
static cudaError_t noalloc_filter_ptr(uint8_t** out) {
    static std::mutex mu;
    static uint8_t* ptr[16] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 16) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        uint8_t* q = nullptr;
        e = cudaGetSymbolAddress((void**)&q, g_noalloc_filter);
        if (e != cudaSuccess) return e;
        ptr[dev] = q;
    }
    *out = ptr[dev];
    return cudaSuccess;
}

cudaError_t launch_single_segment_noalloc(const uint8_t* img, int W, int H, int channels, const int64_t* labels, int64_t k,
                                          int x0, int y0, int w, int h, const EstParams& p, double* out, uint32_t* row_costs,
                                          uint8_t* filter_id, uint32_t* hist, unsigned long long* match, cudaStream_t st) {
    (void)H;
    if (h > NOALLOC_MAX_ROWS) return cudaErrorInvalidValue;
    uint8_t* filt = nullptr;
    cudaError_t e = noalloc_filter_ptr(&filt);
    if (e != cudaSuccess) return e;
    SrcRaw src{img, labels, W, channels, k};
    BigDebug d{row_costs, filter_id, hist, match, nullptr};
    single_segment_kernel<<<1, BIG_T, 0, st>>>(src, x0, y0, w, h, p, filt, out, d);
    count_launch();
    return cudaGetLastError();
}

}  // namespace icr
