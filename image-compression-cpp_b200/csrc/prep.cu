// prep: one CTA per label map, labels only.  Two reference steps in ONE sweep over the label plane:
//   Rule L  label compaction (sorted-unique rank)   (src/compute_rewards.cu:131-141, at::_unique)
//   Rule S  per-segment count + inclusive bbox      (src/segment_stats.cu:11-42)
// Rule L is a circular presence bitmap in shared memory indexed by (label mod capacity): while the
// spread max-min is below the capacity the residues are distinct, so the sweep sets bits and finds
// min/max at once; a popcount prefix scan then gives rank = number of set bits walking up from the
// minimum's slot.  No sort, no host sync on K.
// The sweep works on RUNS, not pixels: a warp looks at 32*VEC consecutive pixels (VEC = 8 labels per
// lane on the int32 path), marks the pixels where the label changes (or a row / the warp's span starts),
// and only those run heads touch the bitmap and append a record to the map's run list (sweep_mark).
// Rule S then walks the run list (sweep_list): one add for the count and four min/max for the bbox per
// run, unconditional shared-memory atomics.  Maps whose runs overflow the list, very wide maps and the
// compact-id output take a second pixel sweep instead (sweep_runs).  The image is not touched here: the
// estimator masks by comparing raw labels with the segment's key.
// Every CTA leaves the largest bbox of its map for the big-segment kernel's early exit; the last CTA to
// finish scans the per-map segment counts into the flat work list the estimator's scheduler walks.
#include "icr_internal.h"
#include <climits>

namespace icr {

constexpr int PREP_T = 512;

template <typename LT, int VEC>
__device__ __forceinline__ void load_labels(const LT* lab, int64_t p, bool inrange, LT (&out)[VEC]) {
    if constexpr (VEC == 4 && sizeof(LT) == 4) {
        int4 q = inrange ? __ldg(reinterpret_cast<const int4*>(lab + p)) : make_int4(0, 0, 0, 0);
        out[0] = q.x; out[1] = q.y; out[2] = q.z; out[3] = q.w;
    } else if constexpr (VEC == 8 && sizeof(LT) == 4) {
        int4 q = make_int4(0, 0, 0, 0), r = q;
        if (inrange) { q = __ldg(reinterpret_cast<const int4*>(lab + p)); r = __ldg(reinterpret_cast<const int4*>(lab + p + 4)); }
        out[0] = q.x; out[1] = q.y; out[2] = q.z; out[3] = q.w; out[4] = r.x; out[5] = r.y; out[6] = r.z; out[7] = r.w;
    } else if constexpr (VEC == 4 && sizeof(LT) == 8) {
        longlong2 a = make_longlong2(0, 0), b = a;
        if (inrange) { a = __ldg(reinterpret_cast<const longlong2*>(lab + p)); b = __ldg(reinterpret_cast<const longlong2*>(lab + p + 2)); }
        out[0] = a.x; out[1] = a.y; out[2] = b.x; out[3] = b.y;
    } else {
        out[0] = inrange ? __ldg(lab + p) : (LT)0;
    }
}

// shared-memory stats table of one map (or its rows in global memory when K > KS_SMEM): struct of arrays, element stride bst
struct StatsPtr { int32_t *cnt, *bx0, *by0, *bx1, *by1; int bst; };

// ---- the second pixel sweep (fallback and compact-id output), on runs.  SMEM selects shared-memory atomics at compile time (a generic pointer would cost an address-space
// check per atomic).
template <typename LT, int VEC, bool SMEM, bool COMPACT_OUT, class RankOf>
__device__ __forceinline__ void sweep_runs(const LT* __restrict__ lab, int64_t HW, int W, const StatsPtr& st, RankOf rank_of,
                                           int32_t* __restrict__ compact_row) {
    const int tid = threadIdx.x, lane = tid & 31;
    // position of this thread's first pixel; advanced incrementally (W % VEC == 0 on the vector path keeps a group in one row)
    int64_t p = (int64_t)tid * VEC;
    int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    const int stepx = (PREP_T * VEC) % W, stepy = (PREP_T * VEC) / W;
    LT nxt[VEC];
    load_labels<LT, VEC>(lab, p, p < HW, nxt);
    for (int64_t base = 0; base < HW; base += (int64_t)PREP_T * VEC, p += (int64_t)PREP_T * VEC) {
        const bool in = p < HW;
        LT l[VEC];
#pragma unroll
        for (int i = 0; i < VEC; ++i) l[i] = nxt[i];
        // the next step's labels are requested before this step's work: one memory latency per sweep, not per step
        load_labels<LT, VEC>(lab, p + (int64_t)PREP_T * VEC, p + (int64_t)PREP_T * VEC < HW, nxt);
        if (COMPACT_OUT && in) {
            uint32_t id[VEC];
#pragma unroll
            for (int i = 0; i < VEC; ++i) id[i] = (i > 0 && l[i] == l[i - 1]) ? id[i - 1] : rank_of(l[i]);
            int32_t* co = compact_row + p;
            if constexpr (VEC >= 4) {
#pragma unroll
                for (int j = 0; j < VEC; j += 4)
                    *reinterpret_cast<int4*>(co + j) = make_int4((int)id[j], (int)id[j + 1], (int)id[j + 2], (int)id[j + 3]);
            } else co[0] = (int)id[0];
        }
        // ---- run heads among this lane's pixels: the label differs from the pixel before, or a row / the warp's span starts
        const LT prevl = __shfl_up_sync(0xffffffffu, l[VEC - 1], 1);
        uint32_t hm = 0;
        if (in) {
            if (lane == 0 || x == 0 || l[0] != prevl) hm = 1u;
#pragma unroll
            for (int i = 1; i < VEC; ++i) if (l[i] != l[i - 1]) hm |= 1u << i;
        }
        // where the next head after this lane's pixels stands, in pixels from the span's start (the span's end if none)
        const uint32_t has = __ballot_sync(0xffffffffu, hm != 0u);
        const uint32_t inm = __ballot_sync(0xffffffffu, in);
        const uint32_t above = lane == 31 ? 0u : (has & (0xffffffffu << (lane + 1)));
        const int nl = above ? __ffs(above) - 1 : 32;                         // next lane holding a head
        const int first = __ffs(hm) - 1;                                        // -1 if none
        const int nfirst = __shfl_sync(0xffffffffu, first, nl & 31);
        const int next_after = above ? nl * VEC + nfirst : __popc(inm) * VEC;
        // ---- every lane takes its lowest unprocessed head; usually one round
        while (__any_sync(0xffffffffu, hm != 0u)) {
            if (hm) {
                const int i = __ffs(hm) - 1;
                hm &= hm - 1;
                LT lv = l[0];
#pragma unroll
                for (int j = 1; j < VEC; ++j) if (i == j) lv = l[j];
                const int endpos = hm ? lane * VEC + (__ffs(hm) - 1) : next_after;
                const int n = endpos - (lane * VEC + i);
                const int kc = (int)rank_of(lv), k = kc * st.bst;
                const int xa = x + i;
                if (SMEM) {
                    atomicAdd(st.cnt + kc, n);
                    atomicMin(st.bx0 + k, xa); atomicMax(st.bx1 + k, xa + n - 1);
                    atomicMin(st.by0 + k, y); atomicMax(st.by1 + k, y);
                } else {
                    // table rows in global memory (more than KS_SMEM segments): skip the atomics that cannot change anything
                    atomicAdd(st.cnt + kc, n);
                    if (xa < st.bx0[k]) atomicMin(st.bx0 + k, xa);
                    if (xa + n - 1 > st.bx1[k]) atomicMax(st.bx1 + k, xa + n - 1);
                    if (y < st.by0[k]) atomicMin(st.by0 + k, y);
                    if (y > st.by1[k]) atomicMax(st.by1 + k, y);
                }
            }
        }
        x += stepx; y += stepy;
        if (x >= W) { x -= W; ++y; }
    }
}

// ---- sweep 1 on runs: the same head detection as sweep_runs; a head marks its label in the presence bitmap, feeds min/max and
// appends one record {slot | (len-1) << 20, y << 16 | x} to its warp's share of the map's run list (every warp owns
// run_cap / warps records and counts in a register: no atomic, no exchange).  With the lists complete, Rule S needs no
// second pass over the pixels: sweep_list walks the records, every warp its own.  Returns the warp's record count
// (beyond `region` when the share overflowed: the caller then falls back to the pixel sweep).
template <typename LT, int VEC>
__device__ __forceinline__ int sweep_mark(const LT* __restrict__ lab, int64_t HW, int W, uint32_t* bitmap, uint32_t cmask, LT& lo, LT& hi,
                                          uint2* __restrict__ runs, int region) {
    int wcount = 0;
    const int tid = threadIdx.x, lane = tid & 31;
    int64_t p = (int64_t)tid * VEC;
    int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    const int stepx = (PREP_T * VEC) % W, stepy = (PREP_T * VEC) / W;
    const uint32_t lt_mask = (1u << lane) - 1u;
    LT nxt[VEC];
    load_labels<LT, VEC>(lab, p, p < HW, nxt);
    for (int64_t base = 0; base < HW; base += (int64_t)PREP_T * VEC, p += (int64_t)PREP_T * VEC) {
        const bool in = p < HW;
        LT l[VEC];
#pragma unroll
        for (int i = 0; i < VEC; ++i) l[i] = nxt[i];
        load_labels<LT, VEC>(lab, p + (int64_t)PREP_T * VEC, p + (int64_t)PREP_T * VEC < HW, nxt);
        const LT prevl = __shfl_up_sync(0xffffffffu, l[VEC - 1], 1);
        uint32_t hm = 0;
        if (in) {
            if (lane == 0 || x == 0 || l[0] != prevl) hm = 1u;
#pragma unroll
            for (int i = 1; i < VEC; ++i) if (l[i] != l[i - 1]) hm |= 1u << i;
        }
        const uint32_t has = __ballot_sync(0xffffffffu, hm != 0u);
        const uint32_t inm = __ballot_sync(0xffffffffu, in);
        const uint32_t above = lane == 31 ? 0u : (has & (0xffffffffu << (lane + 1)));
        const int nl = above ? __ffs(above) - 1 : 32;
        const int first = __ffs(hm) - 1;
        const int nfirst = __shfl_sync(0xffffffffu, first, nl & 31);
        const int next_after = above ? nl * VEC + nfirst : __popc(inm) * VEC;
        uint32_t act;
        while ((act = __ballot_sync(0xffffffffu, hm != 0u)) != 0u) {
            if (hm) {
                const int i = __ffs(hm) - 1;
                hm &= hm - 1;
                LT lv = l[0];
#pragma unroll
                for (int j = 1; j < VEC; ++j) if (i == j) lv = l[j];
                const int endpos = hm ? lane * VEC + (__ffs(hm) - 1) : next_after;
                const int n = endpos - (lane * VEC + i);
                lo = min(lo, lv); hi = max(hi, lv);
                const uint32_t idx = (uint32_t)lv & cmask, bit = 1u << (idx & 31);
                if (!(bitmap[idx >> 5] & bit)) atomicOr(&bitmap[idx >> 5], bit);
                const int pos = wcount + __popc(act & lt_mask);
                if (pos < region) runs[pos] = make_uint2(idx | (uint32_t)(n - 1) << 20, (uint32_t)y << 16 | (uint32_t)(x + i));
            }
            wcount += __popc(act);
        }
        x += stepx; y += stepy;
        if (x >= W) { x -= W; ++y; }
    }
    return wcount;
}

template <bool SMEM, class RankSlot>
__device__ __forceinline__ void sweep_list(const uint2* runs, int nruns, const StatsPtr& st, RankSlot rank_slot) {
    for (int r = threadIdx.x & 31; r < nruns; r += 32) {
        const uint2 rec = runs[r];                    // written by this CTA before a barrier: a plain (coherent) load
        const int n = (int)(rec.x >> 20) + 1, y = (int)(rec.y >> 16), xa = (int)(rec.y & 0xffffu);
        const int kc = (int)rank_slot(rec.x & 0xfffffu), k = kc * st.bst;
        atomicAdd(st.cnt + kc, n);
        if (SMEM) {
            atomicMin(st.bx0 + k, xa); atomicMax(st.bx1 + k, xa + n - 1);
            atomicMin(st.by0 + k, y); atomicMax(st.by1 + k, y);
        } else {
            if (xa < st.bx0[k]) atomicMin(st.bx0 + k, xa);
            if (xa + n - 1 > st.bx1[k]) atomicMax(st.bx1 + k, xa + n - 1);
            if (y < st.by0[k]) atomicMin(st.by0 + k, y);
            if (y > st.by1[k]) atomicMax(st.by1 + k, y);
        }
    }
}

template <typename LT, int VEC, bool COMPACT_OUT>
__global__ void __launch_bounds__(PREP_T, 2) prep_kernel(const LT* __restrict__ labels, Tables t, int cap_words,
                                                       int32_t* __restrict__ compact_out) {
    extern __shared__ uint32_t smem[];
    uint32_t* bitmap = smem;
    uint32_t* prefix = smem + cap_words;
    int32_t* stab = reinterpret_cast<int32_t*>(smem + 2 * cap_words);   // 5 * KS_SMEM
    __shared__ long long red_lo[PREP_T / 32], red_hi[PREP_T / 32];
    __shared__ uint32_t warp_tot[PREP_T / 32];
    __shared__ long long s_min;
    __shared__ int s_bad, s_last, s_over, s_maxarea, s_maxw;
    __shared__ int s_wruns[PREP_T / 32];

    const int v = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int H = t.H, W = t.W;
    const int64_t HW = (int64_t)H * W;
    const LT* lab = labels + (int64_t)v * HW;
    const uint32_t cmask = (uint32_t)cap_words * 32u - 1u;
    int K = 0;
    int bad = ST_OK;

    for (int i = tid; i < cap_words; i += PREP_T) bitmap[i] = 0;
    if (tid == 0) { s_over = 0; s_maxarea = 0; s_maxw = 0; }
    __syncthreads();
    // the run list serves the batch path; the compact-id output needs a pixel pass anyway
    const bool use_list = !COMPACT_OUT && t.run_cap > 0;
    const int region = t.run_cap / (PREP_T / 32);                    // records a warp may write
    uint2* runs = t.runs + (int64_t)v * t.run_cap + (int64_t)wid * region;

    // ---- sweep 1: min / max and the circular presence bitmap (plain read first: almost every bit is already set)
    LT lo = sizeof(LT) == 8 ? (LT)LLONG_MAX : (LT)INT_MAX, hi = sizeof(LT) == 8 ? (LT)LLONG_MIN : (LT)INT_MIN;
    if (use_list) {
        const int nw = sweep_mark<LT, VEC>(lab, HW, W, bitmap, cmask, lo, hi, runs, region);
        if (lane == 0) { s_wruns[wid] = nw; if (nw > region) s_over = 1; }
    }
    LT nxt[VEC];
    if (!use_list) load_labels<LT, VEC>(lab, (int64_t)tid * VEC, (int64_t)tid * VEC < HW, nxt);
    for (int64_t base = 0; base < HW && !use_list; base += (int64_t)PREP_T * VEC) {
        const int64_t p = base + (int64_t)tid * VEC;
        LT l[VEC];
        const bool in = p < HW;
#pragma unroll
        for (int i = 0; i < VEC; ++i) l[i] = nxt[i];
        load_labels<LT, VEC>(lab, p + (int64_t)PREP_T * VEC, p + (int64_t)PREP_T * VEC < HW, nxt);
        if (in) {
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                if (i > 0 && l[i] == l[i - 1]) continue;
                lo = min(lo, l[i]); hi = max(hi, l[i]);
                const uint32_t idx = (uint32_t)l[i] & cmask, bit = 1u << (idx & 31);
                if (!(bitmap[idx >> 5] & bit)) atomicOr(&bitmap[idx >> 5], bit);
            }
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
        for (int i = 1; i < PREP_T / 32; ++i) { lo64 = min(lo64, red_lo[i]); hi64 = max(hi64, red_hi[i]); }
        s_min = lo64;
        // the spread must fit the bitmap (unsigned arithmetic also covers overflow of hi-lo)
        s_bad = ((unsigned long long)hi64 - (unsigned long long)lo64 > (unsigned long long)cmask) ? ST_LABEL_RANGE : ST_OK;
    }
    __syncthreads();
    const long long lmin = s_min;
    const uint32_t smin = (uint32_t)lmin & cmask;
    bad = s_bad;

    if (!bad) {
        // ---- exclusive popcount scan over bitmap words: prefix[w] = set bits below word w
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
        K = (int)total;
        if (K > t.Kmax) bad = ST_K_OVERFLOW;
        __syncthreads();
    }

    if (!bad) {
        const uint32_t pmin = prefix[smin >> 5] + __popc(bitmap[smin >> 5] & ((1u << (smin & 31)) - 1u));
        auto rank_slot = [&](uint32_t slot) -> uint32_t {
            uint32_t ps = prefix[slot >> 5] + __popc(bitmap[slot >> 5] & ((1u << (slot & 31)) - 1u));
            return slot >= smin ? ps - pmin : (uint32_t)K - (pmin - ps);
        };
        auto rank_of = [&](LT label) -> uint32_t { return rank_slot((uint32_t)label & cmask); };
        // ---- keys: original label of each compact id, ascending
        long long* keys = t.keys + (int64_t)v * t.Kmax;
        for (int w = tid; w < cap_words; w += PREP_T) {
            uint32_t bits = bitmap[w];
            while (bits) {
                int b = __ffs(bits) - 1;
                bits &= bits - 1;
                uint32_t slot = (uint32_t)w * 32u + b;
                keys[rank_slot(slot)] = lmin + (long long)((slot - smin) & cmask);
            }
        }
        // ---- stats tables: shared memory when K fits, else straight into the global rows
        int32_t* g_cnt = t.counts + (int64_t)v * t.Kmax;
        int32_t* g_bb = reinterpret_cast<int32_t*>(t.bboxes + (int64_t)v * t.Kmax);
        int32_t* compact_row = COMPACT_OUT ? compact_out + (int64_t)v * HW : nullptr;
        if (K <= KS_SMEM) {
            const StatsPtr st{stab, stab + KS_SMEM, stab + 2 * KS_SMEM, stab + 3 * KS_SMEM, stab + 4 * KS_SMEM, 1};
            for (int k = tid; k < K; k += PREP_T) { st.cnt[k] = 0; st.bx0[k] = W; st.by0[k] = H; st.bx1[k] = -1; st.by1[k] = -1; }
            __syncthreads();
            if (use_list && !s_over) sweep_list<true>(runs, s_wruns[wid], st, rank_slot);
            else sweep_runs<LT, VEC, true, COMPACT_OUT>(lab, HW, W, st, rank_of, compact_row);
            __syncthreads();
            int marea = 0, mw = 0;
            for (int k = tid; k < K; k += PREP_T) {
                g_cnt[k] = st.cnt[k];
                t.bboxes[(int64_t)v * t.Kmax + k] = make_int4(st.bx0[k], st.by0[k], st.bx1[k], st.by1[k]);
                const int bw = st.bx1[k] - st.bx0[k] + 1, bh = st.by1[k] - st.by0[k] + 1;
                marea = max(marea, bw * bh); mw = max(mw, bw);
            }
            marea = __reduce_max_sync(0xffffffffu, marea); mw = __reduce_max_sync(0xffffffffu, mw);
            if (lane == 0 && marea > 0) { atomicMax(&s_maxarea, marea); atomicMax(&s_maxw, mw); }
            __syncthreads();
        } else {
            const StatsPtr st{g_cnt, g_bb, g_bb + 1, g_bb + 2, g_bb + 3, 4};
            for (int k = tid; k < K; k += PREP_T) { g_cnt[k] = 0; g_bb[4 * k] = W; g_bb[4 * k + 1] = H; g_bb[4 * k + 2] = -1; g_bb[4 * k + 3] = -1; }
            __syncthreads();
            if (use_list && !s_over) sweep_list<false>(runs, s_wruns[wid], st, rank_slot);
            else sweep_runs<LT, VEC, false, COMPACT_OUT>(lab, HW, W, st, rank_of, compact_row);
        }
    }

    // ---- publish, and let the last CTA build the flat work list (exclusive scan of the usable segment counts)
    if (tid == 0) {
        t.K[v] = bad == ST_LABEL_RANGE ? 0 : K;
        // the largest bbox of the map, for the big-segment kernel's early exit; not tracked when the rows live in global memory
        t.map_max[v] = (bad || K > KS_SMEM) ? make_int2(INT_MAX, INT_MAX) : make_int2(s_maxarea, s_maxw);
        t.status[v] = bad;
        if (bad) { atomicExch(&t.err[0], 1); atomicMax(&t.err[1], 0x7fffffff - v); }
        __threadfence();
        s_last = atomicAdd(&t.sched[0], 1) == t.V - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    uint32_t carry = 0;
    for (int base = 0; base < t.V; base += PREP_T) {
        const int i = base + tid;
        const volatile int32_t* vK = t.K;
        const volatile int32_t* vS = t.status;
        uint32_t mine = (i < t.V && vS[i] == ST_OK) ? (uint32_t)vK[i] : 0u;
        uint32_t inc = mine;
        for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += n; }
        __syncthreads();
        if (lane == 31) warp_tot[wid] = inc;
        __syncthreads();
        uint32_t woff = 0, total = 0;
        for (int j = 0; j < PREP_T / 32; ++j) { uint32_t x = warp_tot[j]; if (j < wid) woff += x; total += x; }
        if (i < t.V) t.seg_prefix[i] = (int32_t)(carry + woff + inc - mine);
        carry += total;
    }
    if (tid == 0) { t.seg_prefix[t.V] = (int32_t)carry; t.sched[0] = 0; t.sched[1] = 0; }
}

template <typename LT, int VEC>
static cudaError_t launch_t(const void* labels, const Tables& t, int32_t* compact_out, cudaStream_t st) {
    const int cap_words = (int)(label_capacity(t.H, t.W) / 32);
    const size_t smem = (size_t)cap_words * 8 + (size_t)KS_SMEM * 5 * 4;
    auto k1 = prep_kernel<LT, VEC, true>;
    auto k0 = prep_kernel<LT, VEC, false>;
    auto k = compact_out ? k1 : k0;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k<<<t.V, PREP_T, smem, st>>>(static_cast<const LT*>(labels), t, cap_words, compact_out);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_prep(const void* labels, int label_dtype, const Tables& t, int32_t* compact_out, cudaStream_t st) {
    const int64_t HW = (int64_t)t.H * t.W;
    const bool i64 = label_dtype == ICR_LABELS_I64;
    const bool vec = (t.W % 4 == 0) && ((uintptr_t)labels % 16 == 0) && HW % 4 == 0 && (!compact_out || (uintptr_t)compact_out % 16 == 0);
    if (i64) return vec ? launch_t<int64_t, 4>(labels, t, compact_out, st) : launch_t<int64_t, 1>(labels, t, compact_out, st);
    if (vec && t.W % 8 == 0 && HW % 8 == 0 && !compact_out) return launch_t<int32_t, 8>(labels, t, compact_out, st);
    return vec ? launch_t<int32_t, 4>(labels, t, compact_out, st) : launch_t<int32_t, 1>(labels, t, compact_out, st);
}

}  // namespace icr
