// Packed-byte PNG filter arithmetic on one RGBA pixel per 32-bit register (sm_100a).
// Rules F and R of the reference (src/png_size_estimator.cu:47-55,119-152,261-270), restated for
// 4 channels at once: VABSDIFF4 for |a-b| per byte and for the signed-byte cost, half2 lanes for
// the Paeth three-way select (bytes are exact in fp16, so every compare is exact).
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

namespace icr {

// sum over the 4 bytes of |int8(d)| + acc, where d holds |cur-pred| per byte (0..255).
// |int8(r)| is symmetric in r -> -r, so the wrapped residual and |cur-pred| cost the same; with signed byte
// operands VABSDIFF4 against zero is exactly |int8(d)| (d = 128 gives 128), accumulated in the same instruction.
__device__ __forceinline__ uint32_t cost_acc(uint32_t d, uint32_t acc) {
    uint32_t r;
    asm("vabsdiff4.u32.s32.s32.add %0, %1, %2, %3;" : "=r"(r) : "r"(d), "r"(0u), "r"(acc));
    return r;
}

// Rule C (src/compute_rewards.cu:15-20): clamp to [0,255] and round half to even.  A float -> u8 conversion
// saturates to the destination range and maps NaN to 0, which is what fminf(fmaxf(v,0),255) + lrintf produce.
__device__ __forceinline__ uint32_t byte_of(float v) {
    uint32_t r;
    asm("cvt.rni.u8.f32 %0, %1;" : "=r"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ uint32_t rgba_of(float r, float g, float b) {
    // two multiply-adds (FMA pipe) and one OR instead of two shifts and three ORs on the ALU pipe
    return (byte_of(g) * 0x100u + byte_of(r)) | (byte_of(b) * 0x10000u + 0xff000000u);
}

__device__ __forceinline__ uint32_t absdiff4(uint32_t a, uint32_t b) { return __vabsdiffu4(a, b); }

// (a+b)>>1 per byte, no carry between bytes (Avg predictor, :140-145)
__device__ __forceinline__ uint32_t avg4(uint32_t a, uint32_t b) { return (a & b) + (((a ^ b) & 0xfefefefeu) >> 1); }

// bytes (lo, lo+1) of x as half2 {1024+b_lo, 1024+b_hi}: 0x64xx is 1024+xx in fp16, exact.
__device__ __forceinline__ uint32_t unpack_lo(uint32_t x) { return __byte_perm(x, 0x64646464u, 0x5140); }
__device__ __forceinline__ uint32_t unpack_hi(uint32_t x) { return __byte_perm(x, 0x64646464u, 0x5342); }

__device__ __forceinline__ __half2 as_h2(uint32_t x) { return *reinterpret_cast<__half2*>(&x); }

// Paeth select on two channels held as biased half2 (returns the biased half2 bits of the winner).
// Reference order (:47-55): a if pa<=pb && pa<=pc, else b if pb<=pc, else c; with
// pa=|b-c|, pb=|a-c|, pc=|a+b-2c|.  All values are integers below 2048: exact in fp16.
__device__ __forceinline__ uint32_t paeth_h2(uint32_t A, uint32_t B, uint32_t C) {
    __half2 da = __hsub2(as_h2(B), as_h2(C));      // p - a
    __half2 db = __hsub2(as_h2(A), as_h2(C));      // p - b
    __half2 dc = __hadd2(da, db);                  // p - c
    __half2 pa = __habs2(da), pb = __habs2(db), pc = __habs2(dc);
    // the three-way select as exact half arithmetic on 1.0/0.0 masks: it runs on the FMA pipe, which the walk leaves
    // idle, instead of the ALU pipe the walk saturates.  bc = mb ? B : C = C + mb*(B - C); result = bc + ma*(A - bc).
    __half2 ma = __hle2(pa, __hmin2(pb, pc));      // pa <= pb && pa <= pc
    __half2 mb = __hle2(pb, pc);
    __half2 bc = __hfma2(mb, da, as_h2(C));
    __half2 r = __hfma2(ma, __hsub2(as_h2(A), bc), bc);
    return *reinterpret_cast<uint32_t*>(&r);
}

// The same select with every step on the FMA pipe (no HSET2 / HMNMX2, which share the ALU pipe with the byte
// arithmetic): for integer-valued t, sat(t) is the indicator [t >= 1] and relu(t) = max(t, 0), so
//   nb = [pb > pc] = sat(pb - pc),   na = [pa > min(pb, pc)] = sat((pa - pb) + relu(pb - pc)),
//   result = A + na * ((nb ? C : B) - A).  Ten instructions per two channels, all HADD2 / HFMA2.
__device__ __forceinline__ uint32_t paeth_h2_fma(uint32_t A, uint32_t B, uint32_t C) {
    const __half2 one = __float2half2_rn(1.0f);
    __half2 da = __hsub2(as_h2(B), as_h2(C));      // p - a
    __half2 db = __hsub2(as_h2(A), as_h2(C));      // p - b
    __half2 dc = __hadd2(da, db);                  // p - c
    __half2 pa = __habs2(da), pb = __habs2(db), pc = __habs2(dc);
    __half2 nb = __hfma2_sat(pb, one, __hneg2(pc));
    __half2 t1 = __hfma2_relu(pb, one, __hneg2(pc));
    __half2 t2 = __hsub2(pa, pb);
    __half2 na = __hadd2_sat(t2, t1);
    __half2 bc = __hfma2(nb, __hneg2(da), as_h2(B));
    __half2 e = __hsub2(bc, as_h2(A));
    __half2 r = __hfma2(na, e, as_h2(A));
    return *reinterpret_cast<uint32_t*>(&r);
}

// residual bytes of two channels held as biased half2 {1024+cur} and {1024+pred}: (cur - pred) + 1280 lies in
// [1025, 1535], where fp16 has unit spacing, so the low byte of each half of the result is (cur - pred) mod 256
__device__ __forceinline__ uint32_t residual_h2(uint32_t cur, uint32_t pred) {
    __half2 r = __hadd2(__hsub2(as_h2(cur), as_h2(pred)), __float2half2_rn(1280.0f));
    return *reinterpret_cast<uint32_t*>(&r);
}

// Paeth predictor of 4 packed bytes.
__device__ __forceinline__ uint32_t paeth4(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t lo = paeth_h2(unpack_lo(a), unpack_lo(b), unpack_lo(c));
    uint32_t hi = paeth_h2(unpack_hi(a), unpack_hi(b), unpack_hi(c));
    return __byte_perm(lo, hi, 0x6420);
}

// scalar restatement, used by the self-test and the generic-channel path
__device__ __forceinline__ int paeth1(int a, int b, int c) {
    int p = a + b - c;
    int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

struct Costs5 { uint32_t none, sub, up, avg, paeth; };

// add the 5 filter costs of one pixel (cur) given its masked neighbours
__device__ __forceinline__ void add_costs(Costs5& c, uint32_t cur, uint32_t left, uint32_t up, uint32_t upleft) {
    c.none = cost_acc(cur, c.none);
    c.sub = cost_acc(absdiff4(cur, left), c.sub);
    c.up = cost_acc(absdiff4(cur, up), c.up);
    c.avg = cost_acc(absdiff4(cur, avg4(left, up)), c.avg);
    c.paeth = cost_acc(absdiff4(cur, paeth4(left, up, upleft)), c.paeth);
}

// arg-min with strict <, ties to the lowest id (:195-204)
__device__ __forceinline__ int pick_filter(const Costs5& c) {
    uint32_t best = c.none; int id = 0;
    if (c.sub < best) { best = c.sub; id = 1; }
    if (c.up < best) { best = c.up; id = 2; }
    if (c.avg < best) { best = c.avg; id = 3; }
    if (c.paeth < best) { best = c.paeth; id = 4; }
    return id;
}

// residual bytes of one pixel under filter f (:261-270)
__device__ __forceinline__ uint32_t residual4(int f, uint32_t cur, uint32_t left, uint32_t up, uint32_t upleft) {
    uint32_t pred;
    switch (f) {
        case 0: return cur;
        case 1: pred = left; break;
        case 2: pred = up; break;
        case 3: pred = avg4(left, up); break;
        default: pred = paeth4(left, up, upleft); break;
    }
    return __vsub4(cur, pred);
}

}  // namespace icr
