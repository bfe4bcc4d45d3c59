// Instruction-throughput microbenchmarks for the pipes the estimator leans on (sm_100a).
// Prints warp-instructions per cycle per SM for each op at 1 CTA/SM x 1024 threads.
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#define ITER 4096
template <int OP> __global__ void __launch_bounds__(1024) k(uint32_t* out, long long* cyc, uint32_t seed) {
    __shared__ uint32_t sm[8192];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = 0;
    __syncthreads();
    uint32_t a = seed * 2654435761u + threadIdx.x * 40503u, b = a ^ 0x9e3779b9u, c = b * 7u + 1u, d = c ^ (a >> 3);
    uint32_t e = a + 11u, f = b + 13u, g = c + 17u, h = d + 19u;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (OP == 0) { a = __vabsdiffu4(a, b); c = __vabsdiffu4(c, d); e = __vabsdiffu4(e, f); g = __vabsdiffu4(g, h); }
            if (OP == 1) { a = a ^ (b & c); c = c ^ (d & a); e = e ^ (f & g); g = g ^ (h & e); }
            if (OP == 2) { a = __byte_perm(a, b, 0x5140); c = __byte_perm(c, d, 0x5342); e = __byte_perm(e, f, 0x6420); g = __byte_perm(g, h, 0x3120); }
            if (OP == 3) { __half2 x = __hadd2(*(__half2*)&a, *(__half2*)&b); a = *(uint32_t*)&x; __half2 y = __hadd2(*(__half2*)&c, *(__half2*)&d); c = *(uint32_t*)&y;
                           __half2 z = __hadd2(*(__half2*)&e, *(__half2*)&f); e = *(uint32_t*)&z; __half2 w = __hadd2(*(__half2*)&g, *(__half2*)&h); g = *(uint32_t*)&w; }
            if (OP == 4) { a = __hle2_mask(*(__half2*)&a, *(__half2*)&b); c = __hle2_mask(*(__half2*)&c, *(__half2*)&d); e = __hle2_mask(*(__half2*)&e, *(__half2*)&f); g = __hle2_mask(*(__half2*)&g, *(__half2*)&h); }
            if (OP == 5) { a = __shfl_up_sync(0xffffffffu, a, 1); c = __shfl_up_sync(0xffffffffu, c, 2); e = __shfl_up_sync(0xffffffffu, e, 4); g = __shfl_up_sync(0xffffffffu, g, 8); }
            if (OP == 6) { atomicAdd(&sm[(a >> 7) & 8191], 1u); atomicAdd(&sm[(c >> 9) & 8191], 1u); atomicAdd(&sm[(e >> 5) & 8191], 1u); atomicAdd(&sm[(g >> 11) & 8191], 1u);
                           a = a * 1664525u + 1013904223u; c = c * 22695477u + 1u; e = e * 1103515245u + 12345u; g = g * 134775813u + 1u; }
            if (OP == 7) { a = a * 1664525u + 1013904223u; c = c * 22695477u + 1u; e = e * 1103515245u + 12345u; g = g * 134775813u + 1u; }
            if (OP == 8) { a += sm[(a >> 7) & 8191]; c += sm[(c >> 9) & 8191]; e += sm[(e >> 5) & 8191]; g += sm[(g >> 11) & 8191]; }
            if (OP == 9) { atomicAdd(&sm[(a >> 7) & 255], 1u); atomicAdd(&sm[(c >> 9) & 255], 1u); atomicAdd(&sm[(e >> 5) & 255], 1u); atomicAdd(&sm[(g >> 11) & 255], 1u);
                           a = a * 1664525u + 1013904223u; c = c * 22695477u + 1u; e = e * 1103515245u + 12345u; g = g * 134775813u + 1u; }
            if (OP == 10) { sm[(threadIdx.x * 33 + u) & 8191] = a; a += 3; c += 5; e += 7; g += 9; }
            if (OP == 11) { a = __vsub4(a, b); c = __vsub4(c, d); e = __vsub4(e, f); g = __vsub4(g, h); }
            if (OP == 12) { a = __vhaddu4(a, b); c = __vhaddu4(c, d); e = __vhaddu4(e, f); g = __vhaddu4(g, h); }
            if (OP == 13) { a = __match_any_sync(0xffffffffu, a & 63u) + a; c = __match_any_sync(0xffffffffu, c & 63u) + c; e = __match_any_sync(0xffffffffu, e & 63u) + e; g = __match_any_sync(0xffffffffu, g & 63u) + g; }
            if (OP == 14) { if (threadIdx.x & 1) { atomicAdd(&sm[(a >> 7) & 8191], 1u); atomicAdd(&sm[(c >> 9) & 8191], 1u); atomicAdd(&sm[(e >> 5) & 8191], 1u); atomicAdd(&sm[(g >> 11) & 8191], 1u); }
                           a = a * 1664525u + 1013904223u; c = c * 22695477u + 1u; e = e * 1103515245u + 12345u; g = g * 134775813u + 1u; }
            if (OP == 15) { atomicAdd(&sm[it & 255], 1u); atomicAdd(&sm[256 + (it & 255)], 1u); atomicAdd(&sm[512 + (it & 255)], 1u); atomicAdd(&sm[768 + (it & 255)], 1u); }
            if (OP == 16) { atomicAdd(&sm[(threadIdx.x & 3) * 64 + (it & 63)], 1u); atomicAdd(&sm[256 + (threadIdx.x & 3) * 64 + (it & 63)], 1u);
                            atomicAdd(&sm[512 + (threadIdx.x & 3) * 64 + (it & 63)], 1u); atomicAdd(&sm[768 + (threadIdx.x & 3) * 64 + (it & 63)], 1u); }
            if (OP == 17) { atomicAdd(&sm[(a >> 7) & 255], 0x10000u); atomicAdd(&sm[(c >> 9) & 255], 0x10000u); atomicAdd(&sm[(e >> 5) & 255], 0x10000u); atomicAdd(&sm[(g >> 11) & 255], 0x10000u);
                           a = a * 1664525u + 1013904223u; c = c * 22695477u + 1u; e = e * 1103515245u + 12345u; g = g * 134775813u + 1u; }
            if (OP == 18) { atomicAdd(&sm[it & 255], 0x10000u); atomicAdd(&sm[256 + (it & 255)], 0x10000u); atomicAdd(&sm[512 + (it & 255)], 0x10000u); atomicAdd(&sm[768 + (it & 255)], 0x10000u); }
            if (OP == 19) { float x = __uint_as_float(a), y = __uint_as_float(c), z = __uint_as_float(e), w = __uint_as_float(g); uint32_t r0, r1, r2, r3;
                            asm volatile("cvt.rni.u8.f32 %0, %1;" : "=r"(r0) : "f"(x)); asm volatile("cvt.rni.u8.f32 %0, %1;" : "=r"(r1) : "f"(y));
                            asm volatile("cvt.rni.u8.f32 %0, %1;" : "=r"(r2) : "f"(z)); asm volatile("cvt.rni.u8.f32 %0, %1;" : "=r"(r3) : "f"(w));
                            a += r0; c += r1; e += r2; g += r3; }
            if (OP == 20) { a += __reduce_add_sync(0xffffffffu, a); c += __reduce_add_sync(0xffffffffu, c); e += __reduce_add_sync(0xffffffffu, e); g += __reduce_add_sync(0xffffffffu, g); }
            if (OP == 21) { a = __vabsdiffu4(a, b); c = __vabsdiffu4(c, d); __half2 z = __hadd2(*(__half2*)&e, *(__half2*)&f); e = *(uint32_t*)&z; __half2 w = __hadd2(*(__half2*)&g, *(__half2*)&h); g = *(uint32_t*)&w; }
            if (OP == 22) { a = __vabsdiffu4(a, b); c = __vabsdiffu4(c, d); e = e ^ (f & g); g = g ^ (h & e); }
            if (OP == 23) { a = __vabsdiffu4(a, b); c = __vabsdiffu4(c, d); e = e * 1103515245u + 12345u; g = g * 134775813u + 1u; }
            if (OP == 24) { __half2 x = __hadd2(*(__half2*)&a, *(__half2*)&b); a = *(uint32_t*)&x; __half2 y = __hadd2(*(__half2*)&c, *(__half2*)&d); c = *(uint32_t*)&y;
                            e = __hle2_mask(*(__half2*)&e, *(__half2*)&f); g = __hle2_mask(*(__half2*)&g, *(__half2*)&h); }
            if (OP == 25) { a = __byte_perm(a, b, 0x5140); c = __byte_perm(c, d, 0x5342); e = e ^ (f & g); g = g ^ (h & e); }
            if (OP == 26) { __half2 x = __hadd2(*(__half2*)&a, *(__half2*)&b); a = *(uint32_t*)&x; __half2 y = __hadd2(*(__half2*)&c, *(__half2*)&d); c = *(uint32_t*)&y;
                            e = e * 1103515245u + 12345u; g = g * 134775813u + 1u; }
            if (OP == 27) { e = __hle2_mask(*(__half2*)&e, *(__half2*)&f); g = __hle2_mask(*(__half2*)&g, *(__half2*)&h); a = a ^ (b & c); c = c ^ (d & a); }
            if (OP == 28) { asm volatile("{.reg .pred p; setp.ne.u32 p, %1, 0; @p red.shared.add.u32 [%0], 1;}" :: "r"((uint32_t)__cvta_generic_to_shared(&sm[(a >> 7) & 255])), "r"(a & 1u));
                            asm volatile("{.reg .pred p; setp.ne.u32 p, %1, 0; @p red.shared.add.u32 [%0], 1;}" :: "r"((uint32_t)__cvta_generic_to_shared(&sm[(c >> 7) & 255])), "r"(c & 1u));
                            asm volatile("{.reg .pred p; setp.ne.u32 p, %1, 0; @p red.shared.add.u32 [%0], 1;}" :: "r"((uint32_t)__cvta_generic_to_shared(&sm[(e >> 7) & 255])), "r"(e & 1u));
                            asm volatile("{.reg .pred p; setp.ne.u32 p, %1, 0; @p red.shared.add.u32 [%0], 1;}" :: "r"((uint32_t)__cvta_generic_to_shared(&sm[(g >> 7) & 255])), "r"(g & 1u));
                           a = a * 1664525u + 1013904223u; c = c * 22695477u + 1u; e = e * 1103515245u + 12345u; g = g * 134775813u + 1u; }
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a + c + e + g + sm[threadIdx.x];
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP> void run(const char* name, int ops_per_iter_unit) {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    k<OP><<<148, 1024>>>(out, cyc, 1); k<OP><<<148, 1024>>>(out, cyc, 2);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    double winst = 32.0 * ITER * 4 * ops_per_iter_unit;   // warp-instructions of the measured op per SM (32 warps)
    printf("%-28s %8.0f cycles  %6.3f warp-inst/cyc/SM  (%5.2f cyc per warp-inst)  err=%s\n", name, avg, winst / avg, avg / winst, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<0>("VABSDIFF4", 4); run<1>("LOP3", 4); run<2>("PRMT", 4); run<3>("HADD2", 4); run<4>("HSET2 mask", 4);
    run<5>("SHFL up", 4); run<7>("IMAD lcg (baseline for atom)", 4); run<6>("ATOMS spread 8192 + lcg", 4); run<9>("ATOMS 256 bins + lcg", 4);
    run<14>("ATOMS half lanes + lcg", 4); run<8>("LDS random", 4); run<10>("STS stride33", 1); run<11>("vsub4 (emulated)", 4); run<12>("vhaddu4 (emulated)", 4);
    run<13>("MATCH.ANY 64 values", 4);
    run<15>("ATOMS.INC all lanes 1 addr", 4); run<16>("ATOMS.INC 4 addrs per warp", 4); run<17>("ATOMS.ADD 0x10000 256 bins", 4); run<18>("ATOMS.ADD 0x10000 1 addr", 4);
    run<28>("ATOMS.INC 256 bins half pred", 4);
    run<19>("F2I u8 sat (cvt.rni.u8.f32)", 4); run<20>("REDUX.SUM", 4);
    run<21>("2 VABSDIFF4 + 2 HADD2", 4); run<22>("2 VABSDIFF4 + 2 LOP3", 4); run<23>("2 VABSDIFF4 + 2 IMAD", 4); run<24>("2 HADD2 + 2 HSET2", 4);
    run<25>("2 PRMT + 2 LOP3", 4); run<26>("2 HADD2 + 2 IMAD", 4); run<27>("2 HSET2 + 2 LOP3", 4);
    return 0;
}
