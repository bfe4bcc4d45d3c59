/* TEST INFRASTRUCTURE ONLY -- see reward_oracle.h.  CPU restatement of the reference's
 * compression-reward path.  Structure differs from the reference on purpose: the masked slice
 * (Rule P) is materialised first and then treated as an ordinary bpp=4 PNG scanline image,
 * which is what the reference's neighbour masking amounts to
 * (src/png_size_estimator.cu:88-117 vs src/image_slicer.cpp:59-66).
 */
#include "reward_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

void orc_default_params(orc_params* p) { /* include/compute_rewards.cuh:9-16 */
    p->min_pixels = 1; p->L_min = 4; p->beta = 0.012167f; p->b_match_token = 18.0f;
    p->gamma = 0.1f; p->overhead_base = 9.308622; p->adaptive_filter = 1; p->lambda = 0.5;
}

/* src/compute_rewards.cu:15-20 : clamp, round half to even (lrintf), clamp again */
static uint8_t to_byte(float v) {
    if (v < 0.f) v = 0.f; else if (v > 255.f) v = 255.f;
    long iv = lrintf(v);
    if (iv < 0) iv = 0; else if (iv > 255) iv = 255;
    return (uint8_t)iv;
}

void orc_chw3_f32_to_rgba(const float* chw, int H, int W, uint8_t* rgba) { /* :23-53 */
    const size_t plane = (size_t)H * W;
    for (size_t i = 0; i < plane; ++i) {
        rgba[4 * i + 0] = to_byte(chw[i]);
        rgba[4 * i + 1] = to_byte(chw[plane + i]);
        rgba[4 * i + 2] = to_byte(chw[2 * plane + i]);
        rgba[4 * i + 3] = 255;
    }
}

static int cmp_i64(const void* a, const void* b) {
    int64_t x = *(const int64_t*)a, y = *(const int64_t*)b;
    return (x > y) - (x < y);
}

int64_t orc_compact_labels(const int64_t* labels, int64_t n, int64_t* compact, int64_t* uniq) {
    /* src/compute_rewards.cu:131-141 : at::_unique(flat, sorted=true, return_inverse=true) */
    if (n <= 0) return 0;
    int64_t* s = (int64_t*)malloc(sizeof(int64_t) * (size_t)n);
    memcpy(s, labels, sizeof(int64_t) * (size_t)n);
    qsort(s, (size_t)n, sizeof(int64_t), cmp_i64);
    int64_t K = 0;
    for (int64_t i = 0; i < n; ++i) if (i == 0 || s[i] != s[i - 1]) s[K++] = s[i];
    for (int64_t i = 0; i < n; ++i) {           /* inverse = rank by binary search */
        int64_t lo = 0, hi = K - 1, v = labels[i];
        while (lo < hi) { int64_t mid = (lo + hi) / 2; if (s[mid] < v) lo = mid + 1; else hi = mid; }
        compact[i] = lo;
    }
    if (uniq) memcpy(uniq, s, sizeof(int64_t) * (size_t)K);
    free(s);
    return K;
}

void orc_segment_stats(const int64_t* compact, int H, int W, int K, int32_t* counts, int32_t* bboxes) {
    for (int k = 0; k < K; ++k) {               /* src/segment_stats.cu:11-20 */
        counts[k] = 0;
        bboxes[4 * k + 0] = W; bboxes[4 * k + 1] = H; bboxes[4 * k + 2] = -1; bboxes[4 * k + 3] = -1;
    }
    for (int y = 0; y < H; ++y)                 /* src/segment_stats.cu:22-42 */
        for (int x = 0; x < W; ++x) {
            int k = (int)compact[(size_t)y * W + x];
            int32_t* bb = bboxes + 4 * k;
            counts[k] += 1;
            if (x < bb[0]) bb[0] = x;
            if (y < bb[1]) bb[1] = y;
            if (x > bb[2]) bb[2] = x;
            if (y > bb[3]) bb[3] = y;
        }
}

void orc_slice(const uint8_t* rgba, const int64_t* compact, int W, int64_t k,
               int x0, int y0, int w, int h, uint8_t* out) {
    /* src/image_slicer.cpp:59-66 : crop, then masked copy onto a zero canvas */
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            size_t src = (size_t)(y0 + y) * W + (x0 + x);
            uint8_t* d = out + ((size_t)y * w + x) * 4;
            if (compact[src] == k) memcpy(d, rgba + 4 * src, 4); else memset(d, 0, 4);
        }
}

static int paeth(int a, int b, int c) {         /* src/png_size_estimator.cu:47-55 */
    int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    if (pa <= pb && pa <= pc) return a;
    return pb <= pc ? b : c;
}

/* predictor of filter f for slice byte (y, i) with i = x*4+c, bpp = 4 */
static int predict(const uint8_t* s, int w, int y, int i, int f) {
    const size_t stride = (size_t)w * 4;
    int a = i >= 4 ? s[y * stride + i - 4] : 0;                         /* left  */
    int b = y > 0 ? s[(y - 1) * stride + i] : 0;                        /* up    */
    int c = (y > 0 && i >= 4) ? s[(y - 1) * stride + i - 4] : 0;        /* up-left */
    switch (f) {
        case 0: return 0;
        case 1: return a;
        case 2: return b;
        case 3: return (a + b) >> 1;
        default: return paeth(a, b, c);
    }
}

static double size_formula(double Hbar, uint64_t ms, uint64_t mc, uint64_t mls, size_t N,
                           const orc_params* p, int h) {
    /* src/png_size_estimator.cu:10-28 ; beta/token/gamma widened from float (:588) */
    double f = (N > 0 && ms > 0) ? (double)ms / (double)N : 0.0;
    double Lbar = mc > 0 ? (double)mls / (double)mc : (double)p->L_min;
    double b_lit = Hbar + (double)p->beta;
    double b_match = (double)p->b_match_token / Lbar + (double)p->gamma;
    double b_data = (1.0 - f) * b_lit + f * b_match;
    return (p->overhead_base + (double)h) + ((double)N * b_data) / 8.0;
}

double orc_estimate_segment(const uint8_t* rgba, const int64_t* compact, int W, int H,
                            int64_t k, int x0, int y0, int w, int h,
                            const orc_params* p, orc_debug* dbg) {
    (void)H;
    if (w <= 0 || h <= 0) return 0.0;            /* src/png_size_estimator.cu:485-488 */
    const size_t N = (size_t)w * h * 4;
    const size_t stride = (size_t)w * 4;
    uint8_t* s = (uint8_t*)malloc(N);
    uint8_t* res = (uint8_t*)malloc(N);
    uint8_t* filt = (uint8_t*)malloc((size_t)h);
    orc_slice(rgba, compact, W, k, x0, y0, w, h, s);

    /* Rule F  :60-205 : cost = sum |int8(residual)|, arg-min with strict <, or all Paeth (:545) */
    for (int y = 0; y < h; ++y) {
        if (!p->adaptive_filter) { filt[y] = 4; continue; }
        uint32_t cost[5] = {0, 0, 0, 0, 0};
        for (int i = 0; i < (int)stride; ++i) {
            int cur = s[y * stride + i];
            for (int f = 0; f < 5; ++f) {
                int8_t r = (int8_t)(uint8_t)((cur - predict(s, w, y, i, f)) & 0xFF);
                cost[f] += (uint32_t)abs((int)r);
            }
        }
        int best = 0;
        for (int f = 1; f < 5; ++f) if (cost[f] < cost[best]) best = f;
        filt[y] = (uint8_t)best;
        if (dbg && dbg->row_costs) memcpy(dbg->row_costs + 5 * (size_t)y, cost, sizeof cost);
    }

    /* Rule R  :210-275 : residual stream, no filter-type byte */
    for (int y = 0; y < h; ++y)
        for (int i = 0; i < (int)stride; ++i)
            res[y * stride + i] = (uint8_t)((s[y * stride + i] - predict(s, w, y, i, filt[y])) & 0xFF);

    /* Rule H  :365-392 histogram, :281-328 entropy with the reference's tree order */
    uint32_t hist[4 * 256];
    memset(hist, 0, sizeof hist);
    for (size_t i = 0; i < N; ++i) hist[(i & 3) * 256 + res[i]] += 1;
    const uint32_t per_channel = (uint32_t)((uint64_t)w * (uint64_t)h);
    double Hc[4];
    for (int c = 0; c < 4; ++c) {
        double t[256];
        for (int v = 0; v < 256; ++v) {
            uint32_t hv = hist[c * 256 + v];
            if (hv) { double q = (double)hv / (double)per_channel; t[v] = -q * log2(q); } else t[v] = 0.0;
        }
        for (int st = 128; st > 0; st >>= 1) for (int v = 0; v < st; ++v) t[v] += t[v + st];
        Hc[c] = t[0];
    }
    double Hbar = ((Hc[0] + Hc[2]) + (Hc[1] + Hc[3])) / 4.0;   /* tree of :314-328 */

    /* Rule M  :397-463 : 65536 fixed chunks of ceil(N/65536) bytes, runs never cross a chunk */
    uint64_t ms = 0, mc = 0, mls = 0;
    const size_t chunk = (N + 65535) / 65536;
    for (size_t start = 0; start < N; start += chunk) {
        size_t end = start + chunk < N ? start + chunk : N;
        uint8_t cur = res[start];
        int run = 1;
        for (size_t i = start + 1; i <= end; ++i) {
            if (i < end && res[i] == cur) { ++run; continue; }
            if (run >= p->L_min) { ms += (uint64_t)run; mc += 1; mls += (uint64_t)run; }
            if (i < end) { cur = res[i]; run = 1; }
        }
    }

    if (dbg) {
        if (dbg->filter_id) memcpy(dbg->filter_id, filt, (size_t)h);
        if (dbg->residuals) memcpy(dbg->residuals, res, N);
        if (dbg->hist) memcpy(dbg->hist, hist, sizeof hist);
        if (dbg->match) { dbg->match[0] = ms; dbg->match[1] = mc; dbg->match[2] = mls; }
        if (dbg->Hbar) *dbg->Hbar = Hbar;
    }
    free(s); free(res); free(filt);
    /* skip_check_kernel (:4-8) writes 0 for count < min_pixels but finalize_size_kernel (:581)
     * overwrites it unconditionally, so min_pixels has no effect on the size. */
    return size_formula(Hbar, ms, mc, mls, N, p, h);   /* Rule Z */
}

void orc_compute_rewards(const float* images, const int64_t* labels, const double* image_sizes,
                         int B, int H, int W, const orc_params* p, double* rewards,
                         int32_t* K_out, int Kstride, int32_t* counts_out, int32_t* bboxes_out,
                         double* seg_sizes_out) {
    const size_t HW = (size_t)H * W;
    uint8_t* rgba = (uint8_t*)malloc(HW * 4);
    int64_t* compact = (int64_t*)malloc(HW * sizeof(int64_t));
    int32_t* counts = (int32_t*)malloc(HW * sizeof(int32_t));
    int32_t* bboxes = (int32_t*)malloc(HW * 4 * sizeof(int32_t));
    for (int b = 0; b < B; ++b) {               /* src/compute_rewards.cu:125 */
        orc_chw3_f32_to_rgba(images + (size_t)b * 3 * HW, H, W, rgba);
        int K = (int)orc_compact_labels(labels + (size_t)b * HW, (int64_t)HW, compact, NULL);
        if (K_out) K_out[b] = K;
        if (K <= 0) { rewards[b] = 0.0; continue; }    /* :136-139 */
        orc_segment_stats(compact, H, W, K, counts, bboxes);
        double total = 0.0;                      /* seg_sizes.sum() :187 */
        int n_valid = 0;
        for (int k = 0; k < K; ++k) {            /* :159 */
            const int32_t* bb = bboxes + 4 * k;
            double sz = 0.0;
            if (!(bb[2] < bb[0] || bb[3] < bb[1]))
                sz = orc_estimate_segment(rgba, compact, W, H, k, bb[0], bb[1],
                                          bb[2] - bb[0] + 1, bb[3] - bb[1] + 1, p, NULL);
            total += sz;
            if (counts[k] >= p->min_pixels) ++n_valid;   /* :183 */
            if (k < Kstride) {
                if (counts_out) counts_out[(size_t)b * Kstride + k] = counts[k];
                if (bboxes_out) memcpy(bboxes_out + ((size_t)b * Kstride + k) * 4, bb, 4 * sizeof(int32_t));
                if (seg_sizes_out) seg_sizes_out[(size_t)b * Kstride + k] = sz;
            }
        }
        double P = n_valid == 1 ? 1.0 : 0.0;     /* :184-185 */
        double G = (image_sizes[b] - total) / image_sizes[b];   /* :189 */
        rewards[b] = G - p->lambda * P;          /* :191 */
    }
    free(rgba); free(compact); free(counts); free(bboxes);
}

double orc_ema_update(double value, int* initialized, double momentum, const double* rewards, int n) {
    double m = 0.0;                              /* include/ema_baseline.hpp:13 rewards.mean() */
    for (int i = 0; i < n; ++i) m += rewards[i];
    m /= (double)n;
    if (!*initialized) { *initialized = 1; return m; }     /* :14-16 */
    return value * momentum + m * (1.0 - momentum);        /* :18 */
}
