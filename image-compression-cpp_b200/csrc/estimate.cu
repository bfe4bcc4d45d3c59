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

// c*log2(c): table for the counts that occur in practice, one out-of-line copy of the slow path (log2 in double is
// ~100 instructions; inlined at every use it blew the kernel up to 96 KB and the warps stalled on instruction fetch)
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

// ---- asynchronous row staging: one bulk copy (TMA, no tensor map) per bbox row and plane, completion on an mbarrier
constexpr int RAW_PX = 384;           // max over geometries of R * (16-byte aligned row width)

template <typename LT>
struct __align__(16) WarpSmem {
    uint32_t hist[1024];              // 4 x 256 residual histogram
    float raw[3 * RAW_PX];            // R rows x wa pixels of the three f32 planes, as delivered
    LT rawlab[RAW_PX];                // and of the label plane
    uint32_t tile[TILE_WORDS];        // masked RGBA words, (R+1) rolling row slots
    unsigned long long mbar;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* m, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(m)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* m, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(m)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* m, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_u32(m)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* m) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(m)) : "memory");
}

// BULK: rows arrive through bulk copies (needs W % 4 == 0 and 16-byte aligned planes); otherwise plain loads.
template <typename LT, bool DEBUG, bool BULK>
__device__ void warp_segment(const float* __restrict__ img, const LT* __restrict__ lab, int64_t HW, int W, LT key, int count,
                             int4 bb, WarpSmem<LT>& sm, uint32_t& phase, const EstParams& p,
                             const double* __restrict__ clog2, double* out, const DebugOut& dbg) {
    const int lane = threadIdx.x & 31;
    const int w = bb.z - bb.x + 1, h = bb.w - bb.y + 1;
    if (w <= 0 || h <= 0) { if (lane == 0) *out = 0.0; return; }   // src/compute_rewards.cu:164
    uint32_t* hist = sm.hist;
    uint32_t* tile = sm.tile;

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
    // conversion geometry: lanes along x, tw = power of two covering min(w, 32)
    int twl = 0;
    while ((1 << twl) < w && twl < 5) ++twl;
    const int tx = lane & ((1 << twl) - 1), ty = lane >> twl, rpp = 32 >> twl;
    const bool full = count == w * h;                    // the segment fills its bbox: no label has to be read
    // bulk-copy geometry: rows widened to 16-byte boundaries
    const int x0a = bb.x & ~3, wa = ((bb.x + w + 3) & ~3) - x0a, dx = bb.x - x0a;
    const float* imgp = img + (int64_t)bb.y * W + (BULK ? x0a : bb.x);
    const LT* labp = lab + (int64_t)bb.y * W + (BULK ? x0a : bb.x);

    auto issue = [&](int yb) {
        const int nrows = min(R, h - yb);
        if (lane == 0) mbar_expect_tx(&sm.mbar, (uint32_t)nrows * (uint32_t)wa * (12u + (full ? 0u : (uint32_t)sizeof(LT))));
        __syncwarp();
        if (lane < nrows) {
            const int64_t off = (int64_t)(yb + lane) * W;
            bulk_g2s(sm.raw + (0 * R + lane) * wa, imgp + off, wa * 4, &sm.mbar);
            bulk_g2s(sm.raw + (1 * R + lane) * wa, imgp + HW + off, wa * 4, &sm.mbar);
            bulk_g2s(sm.raw + (2 * R + lane) * wa, imgp + 2 * HW + off, wa * 4, &sm.mbar);
            if (!full) bulk_g2s(sm.rawlab + lane * wa, labp + off, wa * (int)sizeof(LT), &sm.mbar);
        }
    };

    // zero the guard column of every slot and the slot of row -1
    for (int s = lane; s <= R; s += 32) tile[s * stride] = 0u;
    for (int x = lane; x < stride + 8; x += 32) tile[x] = 0u;
    int s0 = 1;                                          // slot of the group's first row; row y lives in slot (y+1) mod (R+1)
    if (BULK) issue(0);

    for (int yb = 0; yb < h; yb += R) {
        __syncwarp();
        // ---- rows yb .. yb+R-1 -> masked, packed RGBA words in the tile (lanes along x)
        if (BULK) { mbar_wait(&sm.mbar, phase); phase ^= 1u; }
        for (int r0 = 0; r0 < R; r0 += rpp) {
            const int rr = r0 + ty, yy = yb + rr;
            if (rr < R && yy < h) {
                int slot = s0 + rr; if (slot > R) slot -= R + 1;
                uint32_t* trow = tile + slot * stride + 1;
                if (BULK) {
                    const float* r0p = sm.raw + rr * wa + dx;
                    const LT* l0p = sm.rawlab + rr * wa + dx;
                    for (int x = tx; x < w; x += (1 << twl)) {
                        const uint32_t px = to_byte(r0p[x]) | to_byte(r0p[R * wa + x]) << 8 | to_byte(r0p[2 * R * wa + x]) << 16 | 0xff000000u;
                        trow[x] = (full || l0p[x] == key) ? px : 0u;
                    }
                } else {
                    const int64_t rowoff = (int64_t)yy * W;
                    if (full) { for (int x = tx; x < w; x += (1 << twl)) trow[x] = masked_rgba<LT, true>(imgp, labp, HW, rowoff + x, key); }
                    else { for (int x = tx; x < w; x += (1 << twl)) trow[x] = masked_rgba<LT, false>(imgp, labp, HW, rowoff + x, key); }
                }
            }
        }
        __syncwarp();
        if (BULK && yb + R < h) {                        // the raw rows are consumed: fetch the next group while this one is computed
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(yb + R);
        }
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
#pragma unroll 1
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
    const double Hc = clog2_of(npx, clog2) / (double)npx - acc / (double)npx;       // log2(n) = (n log2 n)/n
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

template <typename LT, bool BULK>
__global__ void __launch_bounds__(EST_WARPS * 32) estimate_kernel(Tables t, Inputs in, EstParams p) {
    extern __shared__ __align__(16) unsigned char est_smem[];
    WarpSmem<LT>* wsm = reinterpret_cast<WarpSmem<LT>*>(est_smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    WarpSmem<LT>& sm = wsm[warp];
    for (int i = lane; i < 1024; i += 32) sm.hist[i] = 0;
    if (lane == 0) { mbar_init(&sm.mbar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncwarp();
    uint32_t phase = 0;
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
            warp_segment<LT, false, BULK>(in.images + (int64_t)(v / in.S) * 3 * HW, labels + (int64_t)v * HW, HW, t.W, key,
                                          __ldg(t.counts + (int64_t)v * t.Kmax + k), bb, sm, phase, p, t.clog2,
                                          t.seg_sizes + (int64_t)v * t.Kmax + k, none);
        }
    }
}

template <typename LT, bool BULK>
__global__ void __launch_bounds__(32) estimate_debug_kernel(Tables t, Inputs in, EstParams p, DebugOut dbg) {
    extern __shared__ __align__(16) unsigned char est_smem[];
    WarpSmem<LT>& sm = *reinterpret_cast<WarpSmem<LT>*>(est_smem);
    for (int i = threadIdx.x; i < 1024; i += 32) sm.hist[i] = 0;
    if (threadIdx.x == 0) { mbar_init(&sm.mbar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncwarp();
    uint32_t phase = 0;
    const int64_t HW = (int64_t)t.H * t.W;
    const int4 bb = t.bboxes[(int64_t)dbg.v * t.Kmax + dbg.k];
    if (for_big_kernel(bb.z - bb.x + 1, bb.w - bb.y + 1, p.L_min)) return;     // the big-segment debug kernel acts
    const LT key = (LT)t.keys[(int64_t)dbg.v * t.Kmax + dbg.k];
    double scratch;
    warp_segment<LT, true, BULK>(in.images + (int64_t)(dbg.v / in.S) * 3 * HW, static_cast<const LT*>(in.labels) + (int64_t)dbg.v * HW,
                                 HW, t.W, key, t.counts[(int64_t)dbg.v * t.Kmax + dbg.k], bb, sm, phase, p, t.clog2,
                                 dbg.size ? dbg.size : &scratch, dbg);
}

static bool bulk_ok(const Tables& t, const Inputs& in) {
    return t.W % 4 == 0 && (uintptr_t)in.images % 16 == 0 && (uintptr_t)in.labels % 16 == 0;
}

template <typename LT, bool BULK>
static cudaError_t launch_est_t(const Tables& t, const Inputs& in, const EstParams& p, cudaStream_t st) {
    static int grid = 0;
    const size_t smem = sizeof(WarpSmem<LT>) * EST_WARPS;
    auto kern = estimate_kernel<LT, BULK>;
    if (!grid) {
        int dev = 0, sms = 148, per = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, kern, EST_WARPS * 32, smem);
        grid = sms * (per < 1 ? 1 : per);     // persistent: every CTA resident, warps pull work until the list is empty
    }
    kern<<<grid, EST_WARPS * 32, smem, st>>>(t, in, p);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_estimate(const Tables& t, const Inputs& in, const EstParams& p, cudaStream_t st) {
    const bool bulk = bulk_ok(t, in);
    if (in.label_dtype == ICR_LABELS_I64) return bulk ? launch_est_t<int64_t, true>(t, in, p, st) : launch_est_t<int64_t, false>(t, in, p, st);
    return bulk ? launch_est_t<int32_t, true>(t, in, p, st) : launch_est_t<int32_t, false>(t, in, p, st);
}

template <typename LT, bool BULK>
static cudaError_t launch_dbg_t(const Tables& t, const Inputs& in, const EstParams& p, const DebugOut& d, cudaStream_t st) {
    auto kern = estimate_debug_kernel<LT, BULK>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(WarpSmem<LT>));
    if (e != cudaSuccess) return e;
    kern<<<1, 32, sizeof(WarpSmem<LT>), st>>>(t, in, p, d);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_estimate_debug(const Tables& t, const Inputs& in, const EstParams& p, int v, int k, uint32_t* row_costs,
                                  uint8_t* filter_id, uint32_t* hist, unsigned long long* match, double* size_out,
                                  cudaStream_t st) {
    DebugOut d{v, k, row_costs, filter_id, hist, match, size_out};
    const bool bulk = bulk_ok(t, in);
    if (in.label_dtype == ICR_LABELS_I64) return bulk ? launch_dbg_t<int64_t, true>(t, in, p, d, st) : launch_dbg_t<int64_t, false>(t, in, p, d, st);
    return bulk ? launch_dbg_t<int32_t, true>(t, in, p, d, st) : launch_dbg_t<int32_t, false>(t, in, p, d, st);
}

}  // namespace icr
