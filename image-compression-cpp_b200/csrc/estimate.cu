// estimate (the shared-memory tile path): Rules C, P, F, R, H, Z for every segment whose bbox is at most NARROW_MAX_W
// pixels wide and that needs no match statistics, one warp per segment, the whole batch in ONE launch.
// Replaces the reference's 13 stream operations per segment (src/png_size_estimator.cu:505-590),
// its two nested host loops (src/compute_rewards.cu:125,159) and its RGBA staging kernel (:23-53).
//
// Persistent warps pull segments from the work list prep built (in map order for this kernel), one per request with the next
// request in flight.  A warp owns one segment and walks its bbox in groups of R rows:
//   stage    every lane takes pixels of the group as aligned quads in row-major order (coalesced, all lanes busy
//            whatever the width).  The quads of group g+1 are requested with cp.async into the warp's raw window
//            while group g is walked, so their latency is not waited for; on arrival: three f32 planes ->
//            saturating round to bytes (Rule C), raw label compared with the segment's key (Rule P) -> one masked
//            RGBA word in a shared-memory tile.  The row above the group stays in a rolling slot; a zero guard
//            column stands left of x = 0.
//   costs    lane = (row of the group, strip of the row).  It walks its strip left to right with the left
//            and upper-left neighbours carried in registers, accumulates the five filter costs with
//            packed-byte arithmetic and parks the Avg and Paeth predictors in two small planes.
//   choose   strips of a row are combined by a butterfly; arg-min with ties to the lowest id (Rule F).
//   residual second walk: predictor = one shared-memory load whose base/step depend on the row's filter
//            (zero word, left, up, Avg plane, Paeth plane), the four residual bytes of every pixel of the bbox
//            (zeros included) go to the warp's 4x256 histogram with unconditional ATOMS.POPC.INC.
//   entropy  branch-free c*log2(c) table look-ups, Rule Z, one f64 store per segment (Rule W waits for the tail kernel,
//            big_segment.cu: signalling every segment cost more than its warp-per-map pass).
// No residual stream, RGBA plane or compact-id plane is ever written to global memory.  Segments that
// need Rule M match statistics (chunk >= L_min) or are wider than NARROW_MAX_W go to the big-segment kernel.
// narrow.cu holds the other estimator for the same work lists (rows in registers, no tile); api.cu picks one.
#include "icr_internal.h"
#include "icr_pixel.cuh"
#include "icr_reward.cuh"

namespace icr {

constexpr int EST_WARPS = 4;          // warps per CTA
constexpr int EST_MAX_W = NARROW_MAX_W;
constexpr int MAXSW = 16;             // widest strip one lane walks

// shared-memory budget per warp.  choose_geometry() only returns geometries whose (R+1) row slots of w+2 words fit TILE_CAP;
// one row per group (S = 32) always fits: 2 * (EST_MAX_W + 2) <= TILE_CAP.
constexpr int TILE_CAP = 456;
constexpr int TILE_WORDS = TILE_CAP + 8;
constexpr int RAW_BYTES = 4096;                   // async staging window per warp
template <typename LT> struct RawWindow {
    static constexpr int parts = 3 + (int)sizeof(LT) / 4;          // 16-byte pieces per quad: r, g, b, labels (two for int64)
    static constexpr int quads = RAW_BYTES / (16 * parts);         // 64 quads with int32 labels, 51 with int64
    static constexpr int per_lane = (quads + 31) / 32;
};
constexpr int PLANE_WORDS = TILE_CAP;             // R rows x (w|1) words < (R+1) * (w+2)
static_assert(2 * (EST_MAX_W + 2) <= TILE_CAP, "a single-row group of the widest bbox must fit the tile");

struct __align__(16) WarpSmem {
    uint32_t hist[1024];              // 4 x 256 residual histogram
    uint32_t tile[TILE_WORDS];        // masked RGBA words, (R+1) rolling row slots of (w+1)|1 words
    uint32_t pavg[PLANE_WORDS];       // Avg predictor of the group's pixels
    uint32_t ppae[PLANE_WORDS];       // Paeth predictor
    uint4 raw[RAW_BYTES / 16];        // the next row group's first quads as they lie in HBM (f32 planes, labels), filled by cp.async
};

// c*log2(c): table for the counts that occur in practice, one out-of-line copy of the slow path
__device__ __noinline__ double clog2_slow(uint32_t c) { return (double)c * log2((double)c); }
__device__ __forceinline__ double clog2_of(uint32_t c, const double* __restrict__ tab) {
    return c < (uint32_t)CLOG2_N ? __ldg(tab + c) : clog2_slow(c);
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

// strips per row S = 2^lg (R = 32/S rows per group) with strip width <= MAXSW, minimising warp steps.  Lanes 0..5 cost
// one candidate each and a warp minimum over (cost, lg) picks the cheapest, the smallest lg among equals.
__device__ __forceinline__ int choose_geometry(int w, int h, int lane) {
    const int l = min(lane, 5);
    const int sw = (w + (1 << l) - 1) >> l;
    const int groups = ((h << l) + 31) >> 5;
    const uint32_t cost = (uint32_t)groups * (uint32_t)(20 * sw + 50 + 3 * l);    // per group: sw pixel steps (~70 instructions each) + ~180 for request, conversion, butterfly, choice
    const bool ok = lane <= 5 && sw <= MAXSW && ((32 >> l) + 1) * (w + 2) <= TILE_CAP;
    // groups <= 2^20 (area cap of this path) and the bracket is below 400: cost * 8 fits 32 bits
    const uint32_t key = __reduce_min_sync(0xffffffffu, ok ? cost * 8u + (uint32_t)l : 0xffffffffu);
    return key == 0xffffffffu ? 5 : (int)(key & 7u);
}

// One request to the work list.  Plain PTX: through atomicAdd the compiler aggregates the (single-lane) atomic over the
// warp and consumes its result with a shuffle right behind it, which makes the warp wait out the atomic's latency at
// the top of every segment instead of at the bottom, where the result is needed.
__device__ __forceinline__ int next_segment(int* counter) {
    int r;
    asm volatile("atom.global.add.s32 %0, [%1], 1;" : "=r"(r) : "l"(counter) : "memory");
    return r;
}

// ---- stage rows yb .. yb+nr-1 of the bbox into the tile: lanes take the nr*w pixels in row-major order
template <typename LT, bool FULL>
__device__ __forceinline__ void stage_rows(const float* __restrict__ img, const LT* __restrict__ lab, int64_t HW, int W, LT key,
                                           int w, float rw, int nr, int s0, int R, int tst, uint32_t* tile, int lane) {
    const int total = nr * w;
    constexpr int U = 4;                              // pixels in flight per lane
    for (int q0 = lane; q0 < total; q0 += 32 * U) {
        float r[U], g[U], b[U];
        LT l[U];
        int dst[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int q = q0 + 32 * u;
            // row = q / w by float reciprocal: q < 32*256 and row <= 32, so (q+0.5)/w is at least 0.5/256 away from an
            // integer while the rounding error stays below 1e-5
            const int rr = __float2int_rz(((float)q + 0.5f) * rw);
            const int x = q - rr * w;
            int slot = s0 + rr; if (slot > R) slot -= R + 1;
            dst[u] = q < total ? slot * tst + 1 + x : -1;
            const int off = rr * W + x;
            if (q < total) {
                r[u] = __ldg(img + off); g[u] = __ldg(img + HW + off); b[u] = __ldg(img + 2 * HW + off);
                if (!FULL) l[u] = __ldg(lab + off);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (dst[u] >= 0) {
                const uint32_t px = rgba_of(r[u], g[u], b[u]);
                ICR_DCHECK(dst[u] < TILE_WORDS);
                tile[dst[u]] = (FULL || l[u] == key) ? px : 0u;
            }
        }
    }
}

// Same, 4 pixels per lane and load: rows are widened to 16-byte boundaries [x0a, x0a + 4*nq) and fetched as float4 / int4
// quads (needs W % 4 == 0 and 16-byte aligned planes); pixels of a quad that fall outside the bbox are dropped.
__device__ __forceinline__ void load_quad(const int32_t* p, int32_t (&l)[4]) {
    const int4 q = __ldg(reinterpret_cast<const int4*>(p));
    l[0] = q.x; l[1] = q.y; l[2] = q.z; l[3] = q.w;
}
__device__ __forceinline__ void load_quad(const int64_t* p, int64_t (&l)[4]) {
    const longlong2 a = __ldg(reinterpret_cast<const longlong2*>(p)), b = __ldg(reinterpret_cast<const longlong2*>(p + 2));
    l[0] = a.x; l[1] = a.y; l[2] = b.x; l[3] = b.y;
}

template <typename LT, bool FULL>
__device__ __forceinline__ void stage_rows_vec(const float* __restrict__ img, const LT* __restrict__ lab, uint32_t HW, uint32_t base, int W, LT key,
                                               int w, int dx, int nq, float rnq, int nr, int s0, int R, int tst, uint32_t* tile, int lane,
                                               int qfirst = 0) {
    const int total = nr * nq;
    constexpr int U = 2;                              // quads in flight per lane
    for (int q0 = qfirst + lane; q0 < total; q0 += 32 * U) {
        float4 r[U], g[U], b[U];
        LT l[U][4];
        int dst[U], xb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int q = q0 + 32 * u;
            const int rr = __float2int_rz(((float)q + 0.5f) * rnq);       // q / nq, exact for q < 32*64 (see stage_rows)
            const int xq = q - rr * nq;
            int slot = s0 + rr; if (slot > R) slot -= R + 1;
            xb[u] = q < total ? 4 * xq - dx : w;      // bbox column of the quad's first pixel; w = nothing to store
            dst[u] = slot * tst + 1 + xb[u];
            const uint32_t off = base + (uint32_t)(rr * W + 4 * xq);      // 32-bit offsets from the map's planes: one wide multiply-add per load
            if (q < total) {
                r[u] = __ldg(reinterpret_cast<const float4*>(img + off));
                g[u] = __ldg(reinterpret_cast<const float4*>(img + (off + HW)));
                b[u] = __ldg(reinterpret_cast<const float4*>(img + (off + 2u * HW)));
                if (!FULL) load_quad(lab + off, l[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (xb[u] < w) {
                const uint32_t px[4] = {rgba_of(r[u].x, g[u].x, b[u].x), rgba_of(r[u].y, g[u].y, b[u].y),
                                        rgba_of(r[u].z, g[u].z, b[u].z), rgba_of(r[u].w, g[u].w, b[u].w)};
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if ((unsigned)(xb[u] + e) < (unsigned)w) { ICR_DCHECK(dst[u] + e >= 0 && dst[u] + e < TILE_WORDS); tile[dst[u] + e] = (FULL || l[u][e] == key) ? px[e] : 0u; }
            }
        }
    }
}

// ---- the same staging split in two, so that the memory latency of a row group is spent under the walks of the group
// before it: raw_request asks the first RawWindow::quads quads of a group into the warp's raw window with cp.async (no
// registers are held while they fly), raw_convert turns them into tile words once they have landed.  Every lane converts
// exactly the quads it requested, so the only ordering needed is the lane's own cp.async.wait_group.
__device__ __forceinline__ void cp_async16(uint32_t saddr, const void* g) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(saddr), "l"(g) : "memory");
}

template <typename LT>
__device__ __forceinline__ void raw_request(const float* __restrict__ img, const LT* __restrict__ lab, uint32_t HW, uint32_t base, int W,
                                            int nq, float rnq, int nr, bool full, uint4* raw, int lane) {
    using RW = RawWindow<LT>;
    const int total = min(nr * nq, RW::quads);
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(raw);
#pragma unroll
    for (int u = 0; u < RW::per_lane; ++u) {
        const int q = lane + 32 * u;
        if (q < total) {
            const int rr = __float2int_rz(((float)q + 0.5f) * rnq);
            const uint32_t off = base + (uint32_t)(rr * W + 4 * (q - rr * nq));
            const uint32_t sa = sbase + 16u * q;
            cp_async16(sa, img + off);
            cp_async16(sa + 16u * RW::quads, img + (off + HW));
            cp_async16(sa + 32u * RW::quads, img + (off + 2u * HW));
            if (!full) {
                cp_async16(sa + 48u * RW::quads, lab + off);
                if (sizeof(LT) == 8) cp_async16(sa + 64u * RW::quads, lab + off + 2);
            }
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

__device__ __forceinline__ void raw_labels(const uint4* raw, int quads, int q, int32_t (&l)[4]) {
    const uint4 a = raw[3 * quads + q];
    l[0] = (int32_t)a.x; l[1] = (int32_t)a.y; l[2] = (int32_t)a.z; l[3] = (int32_t)a.w;
}
__device__ __forceinline__ void raw_labels(const uint4* raw, int quads, int q, int64_t (&l)[4]) {
    const uint4 a = raw[3 * quads + q], b = raw[4 * quads + q];
    l[0] = (int64_t)((uint64_t)a.y << 32 | a.x); l[1] = (int64_t)((uint64_t)a.w << 32 | a.z);
    l[2] = (int64_t)((uint64_t)b.y << 32 | b.x); l[3] = (int64_t)((uint64_t)b.w << 32 | b.z);
}

template <typename LT>
__device__ __forceinline__ void raw_convert(LT key, int w, int dx, int nq, float rnq, int nr, int s0, int R, int tst, bool full,
                                            const uint4* raw, uint32_t* tile, int lane) {
    using RW = RawWindow<LT>;
    const int total = min(nr * nq, RW::quads);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
    for (int u = 0; u < RW::per_lane; ++u) {
        const int q = lane + 32 * u;
        if (q < total) {
            const int rr = __float2int_rz(((float)q + 0.5f) * rnq);
            const int xb = 4 * (q - rr * nq) - dx;
            int slot = s0 + rr; if (slot > R) slot -= R + 1;
            const int dst = slot * tst + 1 + xb;
            ICR_DCHECK(q < RW::quads && RW::parts * RW::quads <= RAW_BYTES / 16);
            const uint4 ur = raw[q], ug = raw[RW::quads + q], ub = raw[2 * RW::quads + q];
            LT l[4] = {key, key, key, key};
            if (!full) raw_labels(raw, RW::quads, q, l);
            const uint32_t px[4] = {rgba_of(__uint_as_float(ur.x), __uint_as_float(ug.x), __uint_as_float(ub.x)),
                                    rgba_of(__uint_as_float(ur.y), __uint_as_float(ug.y), __uint_as_float(ub.y)),
                                    rgba_of(__uint_as_float(ur.z), __uint_as_float(ug.z), __uint_as_float(ub.z)),
                                    rgba_of(__uint_as_float(ur.w), __uint_as_float(ug.w), __uint_as_float(ub.w))};
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if ((unsigned)(xb + e) < (unsigned)w) { ICR_DCHECK(dst + e >= 0 && dst + e < TILE_WORDS); tile[dst + e] = l[e] == key ? px[e] : 0u; }
        }
    }
}

// One segment, one warp.  img/lab point at the image planes / label plane of the segment's map.
template <typename LT, bool DEBUG, bool VEC>
__device__ void warp_segment(const float* __restrict__ img, const LT* __restrict__ lab, int64_t HW, int W, LT key, int count,
                             int4 bb, WarpSmem& sm, const EstParams& p, const double* __restrict__ clog2, double* out,
                             const DebugOut& dbg) {
    const int lane = threadIdx.x & 31;
    const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
    if (w <= 0 || h <= 0) { if (lane == 0) *out = 0.0; return; }   // src/compute_rewards.cu:164
    uint32_t* hist = sm.hist;
    uint32_t* tile = sm.tile;

    const int lg = choose_geometry(w, h, lane);
    const int S = 1 << lg, R = 32 >> lg;
    const int sr = lane >> lg, stp = lane & (S - 1);
    const int sw = (w + S - 1) >> lg;
    const int xs = stp * sw;
    const int nstrip = max(0, min(sw, w - xs));
    const int tst = (w + 1) | 1;                         // guard column + row, odd so rows land on different banks
    const int pst = w | 1;
    const bool full = count == w * h;                    // the segment fills its bbox: no label has to be read
    const float rw = 1.0f / (float)w;
    const float* imgp = img + (int64_t)bb.y * W + bb.x;
    const LT* labp = lab + (int64_t)bb.y * W + bb.x;
    // quad staging: rows widened to 16-byte boundaries
    const int x0a = bb.x & ~3, dx = bb.x - x0a, nq = (dx + w + 3) >> 2;
    const float rnq = 1.0f / (float)nq;

    // zero the guard column of every slot and the slot of row -1
    for (int s = lane; s <= R; s += 32) tile[s * tst] = 0u;
    for (int x = lane; x < tst; x += 32) tile[x] = 0u;
    int s0 = 1;                                          // slot of the group's first row; row y lives in slot (y+1) mod (R+1)

    // the map's plane pointers as opaque register pairs: every staging load is then one wide multiply-add away instead of
    // being re-derived from the kernel arguments
    const float* imgv = img;
    const LT* labv = lab;
    if (VEC) { asm volatile("" : "+l"(imgv)); asm volatile("" : "+l"(labv)); }

    if (VEC) raw_request<LT>(imgv, labv, (uint32_t)HW, (uint32_t)(bb.y * W + x0a), W, nq, rnq, min(R, h), full, sm.raw, lane);

    for (int yb = 0; yb < h; yb += R) {
        __syncwarp();                                    // the previous group's walks are done with the slots about to be rewritten
        const int nr = min(R, h - yb);
        if (VEC) {
            const uint32_t base = (uint32_t)((bb.y + yb) * W + x0a);
            raw_convert<LT>(key, w, dx, nq, rnq, nr, s0, R, tst, full, sm.raw, tile, lane);
            if (nr * nq > RawWindow<LT>::quads) {        // the part of a large group that does not fit the window: straight from HBM
                if (full) stage_rows_vec<LT, true>(imgv, labv, (uint32_t)HW, base, W, key, w, dx, nq, rnq, nr, s0, R, tst, tile, lane, RawWindow<LT>::quads);
                else stage_rows_vec<LT, false>(imgv, labv, (uint32_t)HW, base, W, key, w, dx, nq, rnq, nr, s0, R, tst, tile, lane, RawWindow<LT>::quads);
            }
            // the next group's quads fly while this group is walked twice
            if (yb + R < h) raw_request<LT>(imgv, labv, (uint32_t)HW, base + (uint32_t)(R * W), W, nq, rnq, min(R, h - yb - R), full, sm.raw, lane);
        } else {
            if (full) stage_rows<LT, true>(imgp + (int64_t)yb * W, labp + (int64_t)yb * W, HW, W, key, w, rw, nr, s0, R, tst, tile, lane);
            else stage_rows<LT, false>(imgp + (int64_t)yb * W, labp + (int64_t)yb * W, HW, W, key, w, rw, nr, s0, R, tst, tile, lane);
        }
        __syncwarp();

        // ---- lane = (row sr of the group, strip stp)
        const int y = yb + sr;
        const int n = y < h ? nstrip : 0;
        int slot = s0 + sr; if (slot > R) slot -= R + 1;
        const int uslot = slot == 0 ? R : slot - 1;
        const uint32_t* trow = tile + slot * tst + 1 + xs;
        const uint32_t* urow = tile + uslot * tst + 1 + xs;
        uint32_t* par = sm.pavg + sr * pst + xs;
        uint32_t* ppr = sm.ppae + sr * pst + xs;

        Costs5 c = {0, 0, 0, 0, 0};
        ICR_DCHECK((R + 1) * tst <= TILE_WORDS && R * pst <= PLANE_WORDS && n <= MAXSW);
        ICR_DCHECK(n == 0 || (trow - 1 >= tile && trow + n < tile + TILE_WORDS && urow - 1 >= tile && urow + n < tile + TILE_WORDS));
        ICR_DCHECK(n == 0 || (par + n <= sm.pavg + PLANE_WORDS && ppr + n <= sm.ppae + PLANE_WORDS));
        if (n > 0) {
            uint32_t left = trow[-1], ul = urow[-1];
            uint32_t l_lo = unpack_lo(left), l_hi = unpack_hi(left), ul_lo = unpack_lo(ul), ul_hi = unpack_hi(ul);
            uint32_t cur = trow[0], up = urow[0];
#pragma unroll 4
            for (int i = 0; i < n; ++i) {
                // next pixel's loads first: they do not depend on this pixel's stores (one word past the strip is readable)
                const uint32_t ncur = trow[i + 1], nup = urow[i + 1];
                const uint32_t u_lo = unpack_lo(up), u_hi = unpack_hi(up);
                const uint32_t pa = avg4(left, up);
                #ifdef EST_PAETH_FMA
                const uint32_t pp = __byte_perm(paeth_h2_fma(l_lo, u_lo, ul_lo), paeth_h2_fma(l_hi, u_hi, ul_hi), 0x6420);
#else
                const uint32_t pp = __byte_perm(paeth_h2(l_lo, u_lo, ul_lo), paeth_h2(l_hi, u_hi, ul_hi), 0x6420);
#endif
                c.none = cost_acc(cur, c.none);
                c.sub = cost_acc(absdiff4(cur, left), c.sub);
                c.up = cost_acc(absdiff4(cur, up), c.up);
                c.avg = cost_acc(absdiff4(cur, pa), c.avg);
                c.paeth = cost_acc(absdiff4(cur, pp), c.paeth);
                par[i] = pa; ppr[i] = pp;
                left = cur; l_lo = unpack_lo(cur); l_hi = unpack_hi(cur);
                ul_lo = u_lo; ul_hi = u_hi;
                cur = ncur; up = nup;
            }
        }
        for (int o = 1; o < S; o <<= 1) {
            c.none += __shfl_xor_sync(0xffffffffu, c.none, o);
            c.sub += __shfl_xor_sync(0xffffffffu, c.sub, o);
            c.up += __shfl_xor_sync(0xffffffffu, c.up, o);
            c.avg += __shfl_xor_sync(0xffffffffu, c.avg, o);
            c.paeth += __shfl_xor_sync(0xffffffffu, c.paeth, o);
        }
        const int f = p.adaptive ? pick_filter(c) : 4;
        if (DEBUG && y < h && stp == 0) {
            if (dbg.row_costs && p.adaptive) {
                uint32_t* rc = dbg.row_costs + 5 * y;
                rc[0] = c.none; rc[1] = c.sub; rc[2] = c.up; rc[3] = c.avg; rc[4] = c.paeth;
            }
            if (dbg.filter_id) dbg.filter_id[y] = (uint8_t)f;
        }

        // ---- residual bytes of the chosen filter -> histogram (Rules R, H)
        if (n > 0) {
            const uint32_t* pred = tile;                 // f = 0: the guard word of slot 0, always zero
            int step = 0;
            if (f == 1) { pred = trow - 1; step = 1; }
            if (f == 2) { pred = urow; step = 1; }
            if (f == 3) { pred = par; step = 1; }
            if (f == 4) { pred = ppr; step = 1; }
            uint32_t cur = trow[0], pr = *pred;
#pragma unroll 4
            for (int i = 0; i < n; ++i) {
                pred += step;
                const uint32_t ncur = trow[i + 1], npr = *pred;      // loads of the next pixel ahead of this pixel's reductions
                const uint32_t r = __vsub4(cur, pr);
                atomicAdd(hist + __byte_perm(r, 0, 0x4440), 1u);
                atomicAdd(hist + 256 + __byte_perm(r, 0, 0x4441), 1u);
                atomicAdd(hist + 512 + __byte_perm(r, 0, 0x4442), 1u);
                atomicAdd(hist + 768 + (r >> 24), 1u);
                ICR_DCHECK(pred >= tile - 1 && (pred < tile + TILE_WORDS || (pred >= sm.pavg && pred <= sm.ppae + PLANE_WORDS)));
                cur = ncur; pr = npr;
            }
        }
        s0 = s0 == 0 ? R : s0 - 1;                       // (yb + R + 1) mod (R+1)
    }
    __syncwarp();

    // Rule H: H_c = -sum p log2 p = log2(n) - (sum c log2 c)/n ; lane owns 32 bins of channel lane/8.
    // Every pixel of the bbox was counted (bin 0 included), so the table is complete; it is left zeroed for the next segment.
    const uint32_t npx = (uint32_t)w * (uint32_t)h;
    double acc = 0.0;
    uint4* h4 = reinterpret_cast<uint4*>(hist) + lane * 8;
    if (npx < (uint32_t)CLOG2_N) {               // no bin can lie beyond the c*log2(c) table: branch-free lookups (tab[0] = 0)
#pragma unroll 4
        for (int j0 = 0; j0 < 8; ++j0) {
            const int j = (j0 + lane) & 7;       // rotated by lane: the eight lanes of a quarter-warp are 128 B apart, unrotated they would share four banks
            const uint4 q = h4[j];
            h4[j] = make_uint4(0, 0, 0, 0);
            if (DEBUG && dbg.hist) { uint32_t* d = dbg.hist + lane * 32 + j * 4; d[0] = q.x; d[1] = q.y; d[2] = q.z; d[3] = q.w; }
            acc += (__ldg(clog2 + q.x) + __ldg(clog2 + q.y)) + (__ldg(clog2 + q.z) + __ldg(clog2 + q.w));
        }
    } else {
#pragma unroll 1
        for (int j0 = 0; j0 < 8; ++j0) {
            const int j = (j0 + lane) & 7;       // rotated by lane: the eight lanes of a quarter-warp are 128 B apart, unrotated they would share four banks
            const uint4 q = h4[j];
            h4[j] = make_uint4(0, 0, 0, 0);
            if (DEBUG && dbg.hist) { uint32_t* d = dbg.hist + lane * 32 + j * 4; d[0] = q.x; d[1] = q.y; d[2] = q.z; d[3] = q.w; }
            if (q.x) acc += clog2_of(q.x, clog2);
            if (q.y) acc += clog2_of(q.y, clog2);
            if (q.z) acc += clog2_of(q.z, clog2);
            if (q.w) acc += clog2_of(q.w, clog2);
        }
    }
    for (int o = 1; o < 8; o <<= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    const double Hc = (clog2_of(npx, clog2) - acc) / (double)npx;                    // log2(n) = (n log2 n)/n
    const double H0 = __shfl_sync(0xffffffffu, Hc, 0), H1 = __shfl_sync(0xffffffffu, Hc, 8);
    const double H2 = __shfl_sync(0xffffffffu, Hc, 16), H3 = __shfl_sync(0xffffffffu, Hc, 24);
    if (lane == 0) {
        const double Hbar = ((H0 + H2) + (H1 + H3)) / 4.0;     // tree of mean_channels_kernel (:314-328)
        // Rule Z with no matches (f = 0): (1-0)*b_lit + 0*b_match is b_lit exactly, so the match term's division is not evaluated
        const double sz = (p.overhead + (double)h) + ((double)(4ull * npx) * (Hbar + p.beta)) / 8.0;
        *out = sz;
        if (DEBUG) {
            if (dbg.match) { dbg.match[0] = 0; dbg.match[1] = 0; dbg.match[2] = 0; }
            if (dbg.size) *dbg.size = sz;
        }
    }
}

template <typename LT, bool VEC>
__global__ void __launch_bounds__(EST_WARPS * 32, 4) estimate_kernel(Tables t, Inputs in, EstParams p, RewardArgs rw) {
    extern __shared__ __align__(16) unsigned char est_smem[];
    WarpSmem* wsm = reinterpret_cast<WarpSmem*>(est_smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    WarpSmem& sm = wsm[warp];
    for (int i = lane; i < 1024; i += 32) sm.hist[i] = 0;
    __syncwarp();
    ICR_GRID_DEP_LAUNCH();                // the tail kernel's CTAs may take the places of this kernel's CTAs as they leave
    ICR_GRID_DEP_WAIT();                  // prep's tables and lists are complete
    const int64_t HW = (int64_t)t.H * t.W;
    // the four work lists back to back (prep fills only the first for this kernel; by width class for narrow.cu):
    // list_a front, list_b front, list_b back, list_a back
    const int e0 = t.sched[SC_N_W128], e1 = e0 + t.sched[SC_N_W64], e2 = e1 + t.sched[SC_N_W32], total = e2 + t.sched[SC_N_W16];
    const LT* labels = static_cast<const LT*>(in.labels);
    DebugOut none{};
    // One segment per request; the request for the next segment is issued before the current one is worked on.
    int flat = 0;
    if (lane == 0) flat = next_segment(&t.sched[SC_CUR_NARROW]);
    flat = __shfl_sync(0xffffffffu, flat, 0);
    while (flat < total) {
        int fut = 0;
        if (lane == 0) fut = next_segment(&t.sched[SC_CUR_NARROW]);
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
        warp_segment<LT, false, VEC>(in.images + (int64_t)(v / in.S) * 3 * HW, labels + (int64_t)v * HW, HW, t.W, key, count, bb, sm, p,
                                     t.clog2, t.seg_sizes + row, none);
        flat = __shfl_sync(0xffffffffu, fut, 0);
    }
}

template <typename LT, bool VEC>
__global__ void __launch_bounds__(32) estimate_debug_kernel(Tables t, Inputs in, EstParams p, DebugOut dbg) {
    extern __shared__ __align__(16) unsigned char est_smem[];
    WarpSmem& sm = *reinterpret_cast<WarpSmem*>(est_smem);
    for (int i = threadIdx.x; i < 1024; i += 32) sm.hist[i] = 0;
    __syncwarp();
    const int64_t HW = (int64_t)t.H * t.W;
    const int4 bb = t.bboxes[(int64_t)dbg.v * t.Kmax + dbg.k];
    if (segment_class(bb.z - bb.x + 1, bb.w - bb.y + 1, p.L_min) != CLS_NARROW) return;     // the big-segment debug kernel acts
    const LT key = (LT)t.keys[(int64_t)dbg.v * t.Kmax + dbg.k];
    double scratch;
    warp_segment<LT, true, VEC>(in.images + (int64_t)(dbg.v / in.S) * 3 * HW, static_cast<const LT*>(in.labels) + (int64_t)dbg.v * HW,
                           HW, t.W, key, t.counts[(int64_t)dbg.v * t.Kmax + dbg.k], bb, sm, p, t.clog2,
                           dbg.size ? dbg.size : &scratch, dbg);
}

// quad staging needs 16-byte aligned rows in every plane
static bool vec_ok(const Tables& t, const Inputs& in) {
    return t.W % 4 == 0 && (uintptr_t)in.images % 16 == 0 && (uintptr_t)in.labels % 16 == 0;
}

template <typename LT, bool VEC>
static cudaError_t launch_est_t(const Tables& t, const Inputs& in, const EstParams& p, const RewardArgs& rw, cudaStream_t st) {
    static int per_sm[16] = {0};          // resident CTAs per SM, per device (function attributes are per device too)
    const size_t smem = sizeof(WarpSmem) * EST_WARPS;
    auto kern = estimate_kernel<LT, VEC>;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per = (dev >= 0 && dev < 16) ? per_sm[dev] : 0;
    if (!per) {
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, kern, EST_WARPS * 32, smem);
        if (e != cudaSuccess) return e;
        if (per < 1) per = 1;
        if (dev >= 0 && dev < 16) per_sm[dev] = per;
    }
    e = launch_dependent(kern, dim3(sm_count() * per), dim3(EST_WARPS * 32), smem, st, t, in, p, rw);     // persistent: warps pull work until the list is empty
    if (e != cudaSuccess) return e;
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_tile(const Tables& t, const Inputs& in, const EstParams& p, const RewardArgs& rw, cudaStream_t st) {
    const bool vec = vec_ok(t, in);
    if (in.label_dtype == ICR_LABELS_I64) return vec ? launch_est_t<int64_t, true>(t, in, p, rw, st) : launch_est_t<int64_t, false>(t, in, p, rw, st);
    return vec ? launch_est_t<int32_t, true>(t, in, p, rw, st) : launch_est_t<int32_t, false>(t, in, p, rw, st);
}

template <typename LT, bool VEC>
static cudaError_t launch_dbg_t(const Tables& t, const Inputs& in, const EstParams& p, const DebugOut& d, cudaStream_t st) {
    auto kern = estimate_debug_kernel<LT, VEC>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(WarpSmem));
    if (e != cudaSuccess) return e;
    kern<<<1, 32, sizeof(WarpSmem), st>>>(t, in, p, d);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_estimate_debug(const Tables& t, const Inputs& in, const EstParams& p, int v, int k, uint32_t* row_costs,
                                  uint8_t* filter_id, uint32_t* hist, unsigned long long* match, double* size_out,
                                  cudaStream_t st) {
    DebugOut d{v, k, row_costs, filter_id, hist, match, size_out};
    const bool vec = vec_ok(t, in);
    if (in.label_dtype == ICR_LABELS_I64) return vec ? launch_dbg_t<int64_t, true>(t, in, p, d, st) : launch_dbg_t<int64_t, false>(t, in, p, d, st);
    return vec ? launch_dbg_t<int32_t, true>(t, in, p, d, st) : launch_dbg_t<int32_t, false>(t, in, p, d, st);
}

}  // namespace icr
