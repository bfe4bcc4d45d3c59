// prep: labels only, one CTA per label map at a time.  Two reference steps and the estimator's work lists:
//   Rule L  label compaction (sorted-unique rank)   (src/compute_rewards.cu:131-141, at::_unique)
//   Rule S  per-segment count + inclusive bbox      (src/segment_stats.cu:11-42)
// The label plane is read ONCE, coalesced, and reduced to a change bitmap: one bit per pixel, set where a run of equal
// labels starts (the label differs from the pixel before, or a row starts).  Everything else works on run heads only --
// a few thousand per map instead of tens of thousands of pixels -- and with every lane busy:
//   mark     a thread per 8 labels: two 128-bit loads, eight compares, one byte of the bitmap (shared memory)
//   compact  the set bits become an ascending list of head positions (block-wide popcount scan); a run's length is the
//            distance to the next head, because every row start is a head
//   scan A   a thread per head: its label marks a circular presence bitmap indexed by (label mod capacity) and feeds
//            min / max.  While the spread max-min is below the capacity the residues are distinct, so a popcount
//            prefix scan gives rank = number of set bits walking up from the minimum's slot.  No sort, no host sync on K.
//   scan B   a thread per head: count += length, bbox min / max as unconditional shared-memory atomics on the rank's row
//            (the reference does five global atomics per PIXEL, src/segment_stats.cu:37-41)
//   publish  keys, table rows, and one entry {map, segment} per segment in the work list of the kernel that will take it
//            (register path by width class, CTA-per-segment path), appended with one atomic per warp.
// Maps whose run heads overflow the list (average run below 8 pixels) walk the change bitmap instead of the list.
// The image is not touched here: the estimator masks by comparing raw labels with the segment's key.
#include "icr_internal.h"
#include "icr_reward.cuh"
#include <climits>

namespace icr {

constexpr int PREP_T = 256;          // several small CTAs per SM: the phases of one map are latency-bound, those of different maps overlap
constexpr int PREP_W = PREP_T / 32;

// floor(p / W) for p < 2^31 by one multiply-high and a correction (M = ceil(2^32 / W) overshoots by at most one);
// W = 1 has M = 2^32, stored as 0
struct DivW {
    uint32_t W, M;
    __device__ __forceinline__ void operator()(uint32_t p, int& y, int& x) const {
        uint32_t q = M ? __umulhi(p, M) : p;
        int r = (int)(p - q * W);
        if (r < 0) { r += (int)W; --q; }
        y = (int)q; x = r;
    }
};

// block-wide exclusive scan of one value per thread; returns the exclusive prefix, *total = the block's sum
template <int PREP_W>
__device__ __forceinline__ uint32_t block_exclusive(uint32_t mine, uint32_t* warp_tot, uint32_t* total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = mine;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += n; }
    __syncthreads();                                   // warp_tot may still be read from an earlier scan
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    uint32_t woff = 0, tot = 0;
    for (int i = 0; i < PREP_W; ++i) { const uint32_t x = warp_tot[i]; if (i < wid) woff += x; tot += x; }
    *total = tot;
    return woff + inc - mine;
}

template <typename LT>
__device__ __forceinline__ void load8(const LT* p, LT (&l)[8]) {
    if constexpr (sizeof(LT) == 4) {
        const int4 a = __ldg(reinterpret_cast<const int4*>(p)), b = __ldg(reinterpret_cast<const int4*>(p) + 1);
        l[0] = a.x; l[1] = a.y; l[2] = a.z; l[3] = a.w; l[4] = b.x; l[5] = b.y; l[6] = b.z; l[7] = b.w;
    } else {
        const longlong2* q = reinterpret_cast<const longlong2*>(p);
        const longlong2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
        l[0] = a.x; l[1] = a.y; l[2] = b.x; l[3] = b.y; l[4] = c.x; l[5] = c.y; l[6] = d.x; l[7] = d.y;
    }
}

// shared-memory stats table of one map (or its rows in global memory when K > KS_SMEM): element stride bst
struct StatsPtr { int32_t *cnt, *bx0, *by0, *bx1, *by1; int bst; };

template <bool SMEM>
__device__ __forceinline__ void add_run(const StatsPtr& st, int kc, int n, int xa, int y) {
    const int k = kc * st.bst;
    ICR_DCHECK(kc >= 0 && xa >= 0 && y >= 0 && n > 0);
    atomicAdd(st.cnt + kc, n);
    if (SMEM) {
        atomicMin(st.bx0 + k, xa); atomicMax(st.bx1 + k, xa + n - 1);
        atomicMin(st.by0 + k, y); atomicMax(st.by1 + k, y);
    } else {
        // table rows in global memory (more than KS_SMEM segments): skip the atomics that cannot change anything
        if (xa < st.bx0[k]) atomicMin(st.bx0 + k, xa);
        if (xa + n - 1 > st.bx1[k]) atomicMax(st.bx1 + k, xa + n - 1);
        if (y < st.by0[k]) atomicMin(st.by0 + k, y);
        if (y > st.by1[k]) atomicMax(st.by1 + k, y);
    }
}

// append this thread's segment (or nothing) to one of the five work lists (0..3: register path by width class, 4: big-
// segment kernel): one atomic per warp and list
__device__ __forceinline__ void append_segment(const Tables& t, bool have, int cat, int v, int k) {
    const int lane = threadIdx.x & 31;
    const uint32_t ltm = (1u << lane) - 1u;
#pragma unroll
    for (int c = 0; c < 5; ++c) {
        const uint32_t m = __ballot_sync(0xffffffffu, have && cat == c);
        if (!m) continue;
        int base = 0;
        if (lane == __ffs(m) - 1) base = atomicAdd(&t.sched[SC_N_W128 + c], __popc(m));
        base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
        if (have && cat == c) {
            const int64_t pos = base + __popc(m & ltm);
            ICR_DCHECK(pos >= 0 && pos < t.list_cap);
            const uint2 e = make_uint2((uint32_t)v, (uint32_t)k);
            if (c == 0) t.list_a[pos] = e;
            else if (c == 1) t.list_b[pos] = e;
            else if (c == 2) t.list_b[t.list_cap - 1 - pos] = e;
            else if (c == 3) t.list_a[t.list_cap - 1 - pos] = e;
            else t.list_big[pos] = e;
        }
    }
}

// T threads per CTA: 256 (four CTAs per SM) when the batch fills the machine; 512 or 1024 when it has fewer maps than that
// many CTAs, so that a map's latency-bound phases get more threads instead of leaving SMs idle
template <typename LT, bool VEC, bool CHG_SMEM, int T>
__global__ void __launch_bounds__(T, 1024 / T) prep_kernel(const LT* __restrict__ labels, Tables t, int cap_words, int ks_cap, int lists,
                                                       int L_min, RewardArgs rw) {
    extern __shared__ uint32_t smem[];
    __shared__ long long red_lo[(T / 32)], red_hi[(T / 32)];
    __shared__ uint32_t warp_tot[(T / 32)];
    __shared__ long long s_min;
    __shared__ int s_bad, s_nbig;

    ICR_GRID_DEP_LAUNCH();                // the estimator's CTAs may take the places of this kernel's CTAs as they leave
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int H = t.H, W = t.W;
    const uint32_t HW = (uint32_t)H * (uint32_t)W;                       // below 2^31 (icr_workspace_create)
    const uint32_t nwords = (HW + 31u) >> 5, ngroups = (HW + 7u) >> 3;
    const uint32_t cmask = (uint32_t)cap_words * 32u - 1u;
    const DivW divw{(uint32_t)W, (uint32_t)((0x100000000ull + (uint32_t)W - 1u) / (uint32_t)W)};

    uint32_t* bitmap = smem;
    uint32_t* un = smem + cap_words;                                      // change bitmap, later the popcount prefix
    const uint32_t un_words = CHG_SMEM ? max((uint32_t)cap_words, nwords) : (uint32_t)cap_words;
    int32_t* stab = reinterpret_cast<int32_t*>(un + un_words);            // 5 * ks_cap
    uint32_t* prefix = un;

    for (int v = blockIdx.x; v < t.V; v += gridDim.x) {
        const LT* lab = labels + (int64_t)v * HW;
        uint32_t* chgw = CHG_SMEM ? un : reinterpret_cast<uint32_t*>(t.chg_global + (int64_t)v * t.chg_stride);
        uint8_t* chgb = reinterpret_cast<uint8_t*>(chgw);
        uint32_t* heads = t.heads + (int64_t)v * t.head_cap;
        uint32_t* hslot = t.hslot + (int64_t)v * t.head_cap;

        for (int i = tid; i < cap_words; i += T) bitmap[i] = 0;
        if (tid == 0) s_nbig = 0;
        for (uint32_t g = ngroups + tid; g < nwords * 4u; g += T) chgb[g] = 0;     // padding bits of the last word

        // ---- mark: one bit per pixel, set at run heads
        if (VEC) {
            // W % 8 == 0: a group of 8 labels lies in one row and is 32 (64) bytes aligned.  A thread takes NG groups per
            // trip and asks for the next trip's labels before it looks at this trip's: (2 * NG) x 32 bytes in flight
            constexpr int NG = sizeof(LT) == 4 ? 2 : 1;
            LT cur[NG][8], nxt[NG][8], pv[NG], npv[NG];
            int x[NG];
            const int stepx = (int)((T * NG * 8u) % (uint32_t)W);
#pragma unroll
            for (int j = 0; j < NG; ++j) {
                const uint32_t g = tid + j * T;
                int y0;
                divw(g * 8u, y0, x[j]);
                npv[j] = 0;
                if (g < ngroups) { load8<LT>(lab + g * 8u, nxt[j]); if (x[j]) npv[j] = __ldg(lab + g * 8u - 1); }
            }
            for (uint32_t g0 = tid; g0 < ngroups; g0 += T * NG) {
                bool rowstart[NG];
#pragma unroll
                for (int j = 0; j < NG; ++j) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) cur[j][i] = nxt[j][i];
                    pv[j] = npv[j];
                    rowstart[j] = x[j] == 0;
                    x[j] += stepx; if (x[j] >= W) x[j] -= W;
                    const uint32_t gn = g0 + (j + NG) * T;
                    if (gn < ngroups) { load8<LT>(lab + gn * 8u, nxt[j]); npv[j] = x[j] ? __ldg(lab + gn * 8u - 1) : (LT)0; }
                }
#pragma unroll
                for (int j = 0; j < NG; ++j) {
                    const uint32_t g = g0 + j * T;
                    if (g >= ngroups) break;
                    uint32_t m = (rowstart[j] || cur[j][0] != pv[j]) ? 1u : 0u;
#pragma unroll
                    for (int i = 1; i < 8; ++i) m |= (cur[j][i] != cur[j][i - 1]) ? (1u << i) : 0u;
                    chgb[g] = (uint8_t)m;
                }
            }
        } else {
            for (uint32_t g = tid; g < ngroups; g += T) {
                const uint32_t p0 = g * 8u;
                int y, x;
                divw(p0, y, x);
                LT pv = p0 ? __ldg(lab + p0 - 1) : (LT)0;
                uint32_t m = 0;
                for (int i = 0; i < 8 && p0 + i < HW; ++i) {
                    const LT l = __ldg(lab + p0 + i);
                    if (x == 0 || l != pv) m |= 1u << i;
                    pv = l;
                    if (++x == W) x = 0;
                }
                chgb[g] = (uint8_t)m;
            }
        }
        __syncthreads();

        // ---- compact: the set bits as an ascending list of pixel indices
        const uint32_t wpt = (nwords + T - 1) / T;
        const uint32_t w0 = min((uint32_t)tid * wpt, nwords), w1 = min(w0 + wpt, nwords);
        uint32_t mine = 0;
        for (uint32_t w = w0; w < w1; ++w) mine += __popc(chgw[w]);
        uint32_t nheads;
        uint32_t hb = block_exclusive<T / 32>(mine, warp_tot, &nheads);
        const bool listed = nheads <= (uint32_t)t.head_cap;
        if (listed)
            for (uint32_t w = w0; w < w1; ++w) {
                uint32_t bits = chgw[w];
                while (bits) { const int b = __ffs(bits) - 1; bits &= bits - 1; ICR_DCHECK(hb < (uint32_t)t.head_cap); heads[hb++] = w * 32u + b; }
            }
        __syncthreads();

        // ---- scan A: min / max and the circular presence bitmap (plain read first: almost every bit is already set)
        LT lo = sizeof(LT) == 8 ? (LT)LLONG_MAX : (LT)INT_MAX, hi = sizeof(LT) == 8 ? (LT)LLONG_MIN : (LT)INT_MIN;
        auto mark = [&](LT l) -> uint32_t {
            lo = min(lo, l); hi = max(hi, l);
            const uint32_t idx = (uint32_t)l & cmask, bit = 1u << (idx & 31);
            ICR_DCHECK((idx >> 5) < (uint32_t)cap_words);
            if (!(bitmap[idx >> 5] & bit)) atomicOr(&bitmap[idx >> 5], bit);
            return idx;
        };
        if (listed) {
            // four heads per thread and trip: the positions, then the labels, are asked for together (two memory round trips
            // per trip instead of two per head)
            for (uint32_t i0 = tid; i0 < nheads; i0 += 4 * T) {
                uint32_t p[4];
                LT l[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) { const uint32_t i = i0 + j * T; p[j] = i < nheads ? heads[i] : 0xffffffffu; }
#pragma unroll
                for (int j = 0; j < 4; ++j) l[j] = p[j] != 0xffffffffu ? __ldg(lab + p[j]) : (LT)0;
#pragma unroll
                for (int j = 0; j < 4; ++j) if (p[j] != 0xffffffffu) hslot[i0 + j * T] = mark(l[j]);
            }
        } else {
            for (uint32_t w = tid; w < nwords; w += T) {
                uint32_t bits = chgw[w];
                while (bits) { const int b = __ffs(bits) - 1; bits &= bits - 1; mark(__ldg(lab + w * 32u + b)); }
            }
        }
        long long lo64 = lo, hi64 = hi;
        for (int o = 16; o; o >>= 1) {
            lo64 = min(lo64, __shfl_xor_sync(0xffffffffu, lo64, o));
            hi64 = max(hi64, __shfl_xor_sync(0xffffffffu, hi64, o));
        }
        if (lane == 0) { red_lo[wid] = lo64; red_hi[wid] = hi64; }
        __syncthreads();
        if (tid == 0) {
            for (int i = 1; i < (T / 32); ++i) { lo64 = min(lo64, red_lo[i]); hi64 = max(hi64, red_hi[i]); }
            s_min = lo64;
            // the spread must fit the bitmap (unsigned arithmetic also covers overflow of hi-lo)
            s_bad = ((unsigned long long)hi64 - (unsigned long long)lo64 > (unsigned long long)cmask) ? ST_LABEL_RANGE : ST_OK;
        }
        // a map that walks the change bitmap needs it beyond this point, where the prefix array takes its place in shared
        // memory: it moves to the (unused) head list region
        const uint32_t* chg_it = chgw;
        if (CHG_SMEM && !listed) {
            uint32_t* spill = heads;                                   // head_cap >= HW/8 words, the bitmap has HW/32
            for (uint32_t w = tid; w < nwords; w += T) spill[w] = chgw[w];
            chg_it = spill;
        }
        __syncthreads();
        const long long lmin = s_min;
        const uint32_t smin = (uint32_t)lmin & cmask;
        int bad = s_bad, K = 0;

        if (!bad) {
            // ---- exclusive popcount scan over bitmap words: prefix[w] = set bits below word w
            const int per = (cap_words + T - 1) / T;
            const int b0 = min(tid * per, cap_words), b1 = min(b0 + per, cap_words);
            uint32_t cnt = 0;
            for (int w = b0; w < b1; ++w) cnt += __popc(bitmap[w]);
            uint32_t total;
            uint32_t run = block_exclusive<T / 32>(cnt, warp_tot, &total);
            for (int w = b0; w < b1; ++w) { prefix[w] = run; run += __popc(bitmap[w]); }
            K = (int)total;
            if (K > t.Kmax) bad = ST_K_OVERFLOW;
            __syncthreads();
        }

        if (!bad) {
            const uint32_t pmin = prefix[smin >> 5] + __popc(bitmap[smin >> 5] & ((1u << (smin & 31)) - 1u));
            auto rank_slot = [&](uint32_t slot) -> uint32_t {
                const uint32_t ps = prefix[slot >> 5] + __popc(bitmap[slot >> 5] & ((1u << (slot & 31)) - 1u));
                return slot >= smin ? ps - pmin : (uint32_t)K - (pmin - ps);
            };
            // ---- keys: original label of each compact id, ascending
            long long* keys = t.keys + (int64_t)v * t.Kmax;
            for (int w = tid; w < cap_words; w += T) {
                uint32_t bits = bitmap[w];
                while (bits) {
                    const int b = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const uint32_t slot = (uint32_t)w * 32u + b;
                    ICR_DCHECK(rank_slot(slot) < (uint32_t)K && K <= t.Kmax);
                    keys[rank_slot(slot)] = lmin + (long long)((slot - smin) & cmask);
                }
            }
            // ---- scan B (Rule S): stats tables in shared memory when K fits, else straight into the global rows
            int32_t* g_cnt = t.counts + (int64_t)v * t.Kmax;
            int32_t* g_bb = reinterpret_cast<int32_t*>(t.bboxes + (int64_t)v * t.Kmax);
            const bool in_smem = K <= ks_cap;
            const StatsPtr st = in_smem ? StatsPtr{stab, stab + ks_cap, stab + 2 * ks_cap, stab + 3 * ks_cap, stab + 4 * ks_cap, 1}
                                        : StatsPtr{g_cnt, g_bb, g_bb + 1, g_bb + 2, g_bb + 3, 4};
            for (int k = tid; k < K; k += T) {
                const int kk = k * st.bst;
                st.cnt[k] = 0; st.bx0[kk] = W; st.by0[kk] = H; st.bx1[kk] = -1; st.by1[kk] = -1;
            }
            __syncthreads();
            if (listed) {
                for (uint32_t i0 = tid; i0 < nheads; i0 += 4 * T) {
                    uint32_t p[4], nx[4], sl[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint32_t i = i0 + j * T;
                        p[j] = i < nheads ? heads[i] : 0xffffffffu;
                        nx[j] = i + 1 < nheads ? heads[i + 1] : HW;      // every row start is a head: the run ends in its row
                        sl[j] = i < nheads ? hslot[i] : 0u;
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        if (p[j] == 0xffffffffu) continue;
                        int y, x;
                        divw(p[j], y, x);
                        const int kc = (int)rank_slot(sl[j]);
                        if (in_smem) add_run<true>(st, kc, (int)(nx[j] - p[j]), x, y); else add_run<false>(st, kc, (int)(nx[j] - p[j]), x, y);
                    }
                }
            } else {
                for (uint32_t w = tid; w < nwords; w += T) {
                    uint32_t bits = chg_it[w];
                    while (bits) {
                        const int b = __ffs(bits) - 1;
                        bits &= bits - 1;
                        const uint32_t p = w * 32u + b;
                        uint32_t nx;
                        if (bits) nx = w * 32u + (__ffs(bits) - 1);
                        else {
                            uint32_t ww = w + 1;
                            while (ww < nwords && chg_it[ww] == 0u) ++ww;
                            nx = ww < nwords ? ww * 32u + (__ffs(chg_it[ww]) - 1) : HW;
                        }
                        int y, x;
                        divw(p, y, x);
                        const int kc = (int)rank_slot((uint32_t)__ldg(lab + p) & cmask);
                        if (in_smem) add_run<true>(st, kc, (int)(nx - p), x, y); else add_run<false>(st, kc, (int)(nx - p), x, y);
                    }
                }
            }
            __syncthreads();
            // ---- table rows out, and every segment into the work list of the kernel that takes it
            for (int k0 = 0; k0 < K; k0 += T) {
                const int k = k0 + tid;
                const bool have = k < K;
                int4 bb = make_int4(0, 0, -1, -1);
                if (have) {
                    if (in_smem) {
                        bb = make_int4(st.bx0[k], st.by0[k], st.bx1[k], st.by1[k]);
                        g_cnt[k] = st.cnt[k]; t.bboxes[(int64_t)v * t.Kmax + k] = bb;
                    } else {
                        bb = __ldcg(t.bboxes + (int64_t)v * t.Kmax + k);      // written by atomics of the whole CTA: read past L1
                    }
                }
                if (lists) {
                    const int bw = bb.z - bb.x + 1, bh = bb.w - bb.y + 1;
                    // lists = 1: one warp-path list in map order (neighbouring segments are walked at about the same time, so
                    // the sectors and label rows they share are still in the L2); lists = 2: by width class
                    const int cat = segment_class(bw, bh, L_min) == CLS_NARROW ? (lists == 2 ? width_class(bw) : 0) : 4;
                    append_segment(t, have, cat, v, k);
                    const uint32_t bigs = __ballot_sync(0xffffffffu, have && cat == 4);
                    if (lane == 0 && bigs) atomicAdd(&s_nbig, __popc(bigs));
                }
            }
        }

        // ---- publish
        __syncthreads();
        if (tid == 0) {
            t.nbig[v] = bad ? 0 : s_nbig;
            t.K[v] = bad == ST_LABEL_RANGE ? 0 : K;
            t.status[v] = bad;
            t.done[v] = 0;
            if (bad) {
                atomicExch(&t.err[0], 1); atomicMax(&t.err[1], 0x7fffffff - v);
                atomicOr(&t.err[2], bad); atomicMax(&t.err[3], K);
                if (lists) store_reward(t, rw.rewards, v, __longlong_as_double(0x7ff8000000000000ll));   // nobody else will
            }
        }
        __syncthreads();                               // shared memory is reused by the CTA's next map
    }
}

template <typename LT, bool VEC, bool CHG_SMEM, int T>
static cudaError_t launch_tt(const void* labels, const Tables& t, int lists, int L_min, const RewardArgs& rw, cudaStream_t st) {
    const int cap_words = (int)(label_capacity(t.H, t.W) / 32);
    const int64_t nwords = chg_bytes(t.H, t.W) / 4;
    const int64_t un_words = CHG_SMEM ? (nwords > cap_words ? nwords : cap_words) : cap_words;
    const int ks_cap = t.Kmax < KS_SMEM ? t.Kmax : KS_SMEM;              // rows of the shared-memory stats table
    const size_t smem = ((size_t)cap_words + (size_t)un_words) * 4 + (size_t)ks_cap * 5 * 4;
    auto k = prep_kernel<LT, VEC, CHG_SMEM, T>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k, T, smem);
    if (e != cudaSuccess) return e;
    const int resident = sm_count() * (per < 1 ? 1 : per);
    const int grid = t.V < resident ? t.V : resident;                   // persistent beyond one wave: CTAs stride over the maps
    k<<<grid, T, smem, st>>>(static_cast<const LT*>(labels), t, cap_words, ks_cap, lists, L_min, rw);
    count_launch();
    return cudaGetLastError();
}

template <typename LT, bool VEC, bool CHG_SMEM>
static cudaError_t launch_t(const void* labels, const Tables& t, int lists, int L_min, const RewardArgs& rw, cudaStream_t st) {
    if (VEC && CHG_SMEM) {                      // the common case gets the wide CTAs for small batches
        if (t.V <= sm_count()) return launch_tt<LT, VEC, CHG_SMEM, (VEC && CHG_SMEM) ? 1024 : PREP_T>(labels, t, lists, L_min, rw, st);
        if (t.V <= 2 * sm_count()) return launch_tt<LT, VEC, CHG_SMEM, (VEC && CHG_SMEM) ? 512 : PREP_T>(labels, t, lists, L_min, rw, st);
    }
    return launch_tt<LT, VEC, CHG_SMEM, PREP_T>(labels, t, lists, L_min, rw, st);
}

template <typename LT>
static cudaError_t launch_lt(const void* labels, const Tables& t, int lists, int L_min, const RewardArgs& rw, cudaStream_t st) {
    const bool vec = t.W % 8 == 0 && (uintptr_t)labels % 16 == 0;
    const bool sm = t.chg_global == nullptr;
    if (vec) return sm ? launch_t<LT, true, true>(labels, t, lists, L_min, rw, st) : launch_t<LT, true, false>(labels, t, lists, L_min, rw, st);
    return sm ? launch_t<LT, false, true>(labels, t, lists, L_min, rw, st) : launch_t<LT, false, false>(labels, t, lists, L_min, rw, st);
}

cudaError_t launch_prep(const void* labels, int label_dtype, const Tables& t, int lists, int L_min, const RewardArgs& rw,
                        cudaStream_t st) {
    if (label_dtype == ICR_LABELS_I64) return launch_lt<int64_t>(labels, t, lists, L_min, rw, st);
    return launch_lt<int32_t>(labels, t, lists, L_min, rw, st);
}

// ---- compact ids (what at::_unique's inverse holds): rank of every pixel's label among the map's keys, by binary search
// with the previous pixel's answer reused along runs.  Only icr_segment_stats asks for this plane.
template <typename LT>
__global__ void __launch_bounds__(256) compact_ids_kernel(const LT* __restrict__ labels, Tables t, int32_t* __restrict__ out) {
    const int v = blockIdx.y;
    if (t.status[v] != ST_OK) return;
    const int K = t.K[v];
    const int64_t HW = (int64_t)t.H * t.W;
    const LT* lab = labels + v * HW;
    const long long* keys = t.keys + (int64_t)v * t.Kmax;
    int32_t* o = out + v * HW;
    long long last = 0; int last_rank = -1;
    const int64_t per = (HW + (int64_t)gridDim.x * blockDim.x - 1) / ((int64_t)gridDim.x * blockDim.x);
    const int64_t a = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * per, b = a + per < HW ? a + per : HW;
    for (int64_t p = a; p < b; ++p) {
        const long long l = (long long)lab[p];
        if (last_rank < 0 || l != last) {
            int lo = 0, hi = K - 1;
            while (lo < hi) { const int mid = (lo + hi) >> 1; if (__ldg(keys + mid) < l) lo = mid + 1; else hi = mid; }
            last = l; last_rank = lo;
        }
        o[p] = last_rank;
    }
}

cudaError_t launch_compact_ids(const void* labels, int label_dtype, const Tables& t, int32_t* compact_out, cudaStream_t st) {
    const int64_t HW = (int64_t)t.H * t.W;
    int gx = (int)((HW + 256 * 8 - 1) / (256 * 8));
    if (gx < 1) gx = 1;
    if (gx > 1024) gx = 1024;
    const dim3 grid(gx, t.V);
    if (label_dtype == ICR_LABELS_I64) compact_ids_kernel<int64_t><<<grid, 256, 0, st>>>(static_cast<const int64_t*>(labels), t, compact_out);
    else compact_ids_kernel<int32_t><<<grid, 256, 0, st>>>(static_cast<const int32_t*>(labels), t, compact_out);
    count_launch();
    return cudaGetLastError();
}

// ---- flat (map, segment) numbering for the slices: exclusive scan of the segment counts of the usable maps
__global__ void __launch_bounds__(1024) seg_prefix_kernel(Tables t) {
    __shared__ uint32_t wt[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < t.V; base += 1024) {
        const int i = base + tid;
        const uint32_t mine = (i < t.V && t.status[i] == ST_OK) ? (uint32_t)t.K[i] : 0u;
        uint32_t inc = mine;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += n; }
        if (lane == 31) wt[wid] = inc;
        __syncthreads();
        uint32_t woff = 0, total = 0;
        for (int j = 0; j < 32; ++j) { const uint32_t x = wt[j]; if (j < wid) woff += x; total += x; }
        if (i < t.V) t.seg_prefix[i] = (int32_t)(s_carry + woff + inc - mine);
        __syncthreads();
        if (tid == 0) s_carry += total;
        __syncthreads();
    }
    if (tid == 0) t.seg_prefix[t.V] = (int32_t)s_carry;
}

cudaError_t launch_seg_prefix(const Tables& t, cudaStream_t st) {
    seg_prefix_kernel<<<1, 1024, 0, st>>>(t);
    count_launch();
    return cudaGetLastError();
}

}  // namespace icr
