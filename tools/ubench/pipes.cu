// Throughput of the instruction classes the modmul could be built from, per SM per clock (B200 microbenchmark).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITER 4096
template <int OP> __global__ void k(uint64_t *out, uint64_t seed) {
    uint64_t a = seed + threadIdx.x, b = seed * 3 + blockIdx.x, c = seed ^ 0x9E3779B97F4A7C15ull, d = a ^ b;
    double x = (double)a, y = (double)b * 1e-9, z = 1.000001, w = 0.999999;
    uint32_t p = (uint32_t)a, q = (uint32_t)b, r = (uint32_t)c, s = (uint32_t)d;
#pragma unroll 1
    for (int i = 0; i < ITER; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (OP == 0) {  // IMAD.WIDE.U32 x4 independent
                a = (uint64_t)p * q + a; b = (uint64_t)q * r + b; c = (uint64_t)r * s + c; d = (uint64_t)s * p + d;
                p += (uint32_t)a; q ^= (uint32_t)b;
            } else if (OP == 1) {  // IMAD (32-bit) x4
                p = p * q + r; q = q * r + s; r = r * s + p; s = s * p + q;
            } else if (OP == 2) {  // DFMA x4
                x = fma(x, z, y); y = fma(y, w, x); z = fma(z, w, 1e-30); w = fma(w, z, 1e-30);
            } else if (OP == 3) {  // u64 -> f64 -> u64 conversions
                x = (double)(long long)a; y = (double)(long long)b; a = (uint64_t)(long long)(y * 0.5) + c; b = (uint64_t)(long long)(x * 0.5) + d;
            } else if (OP == 4) {  // __umul64hi
                a = __umul64hi(a, c) + b; b = __umul64hi(b, d) + a;
            } else if (OP == 5) {  // IADD3 / LOP3 (alu)
                p = (p + q) ^ r; q = (q + r) ^ s; r = (r + s) ^ p; s = (s + p) ^ q;
            } else if (OP == 6) {  // u32 -> f64 and f64 -> s32 conversions (I2F.F64.U32 / F2I.S32.F64)
                x = (double)p; y = (double)q; p = (uint32_t)(int)(y * 0.5) + r; q = (uint32_t)(int)(x * 0.5) + s;
            } else if (OP == 7) {  // IMAD.WIDE x4 and DFMA x4 interleaved in every warp: do the two pipes overlap?
                a = (uint64_t)p * q + a; x = fma(x, z, y); b = (uint64_t)q * r + b; y = fma(y, w, x);
                c = (uint64_t)r * s + c; z = fma(z, w, 1e-30); d = (uint64_t)s * p + d; w = fma(w, z, 1e-30);
                p += (uint32_t)a; q ^= (uint32_t)b;
            } else if (OP == 8) {  // even warps IMAD.WIDE x4, odd warps DFMA x4 (two CTAs' phases side by side)
                if ((threadIdx.x >> 5) & 1) {
                    x = fma(x, z, y); y = fma(y, w, x); z = fma(z, w, 1e-30); w = fma(w, z, 1e-30);
                } else {
                    a = (uint64_t)p * q + a; b = (uint64_t)q * r + b; c = (uint64_t)r * s + c; d = (uint64_t)s * p + d;
                    p += (uint32_t)a; q ^= (uint32_t)b;
                }
            } else if (OP == 11) {  // IMAD.HI.U32 x4: the high half of a 32 x 32 product
                p = __umulhi(p, q) + r; q = __umulhi(q, r) + s; r = __umulhi(r, s) + p; s = __umulhi(s, p) + q;
            } else if (OP == 9) {  // IMAD.WIDE x4 + IADD/LOP3 x4 interleaved
                a = (uint64_t)p * q + a; r = (r + s) ^ p; b = (uint64_t)q * r + b; s = (s + p) ^ q;
                c = (uint64_t)r * s + c; p = (p + q) ^ r; d = (uint64_t)s * p + d; q = (q + r) ^ s;
            } else if (OP == 10) {  // DFMA x4 + IADD/LOP3 x4 interleaved
                x = fma(x, z, y); p = (p + q) ^ r; y = fma(y, w, x); q = (q + r) ^ s;
                z = fma(z, w, 1e-30); r = (r + s) ^ p; w = fma(w, z, 1e-30); s = (s + p) ^ q;
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a ^ b ^ c ^ d ^ (uint64_t)(x + y + z + w) ^ p ^ q ^ r ^ s;
}
// Warp-specialised mixes: every warp runs ONE instruction class for the whole kernel (the branch is outside the loop, so
// nothing is predicated): half of the warps of each CTA class A, the other half class B, the same iteration count per
// warp as in the single-class runs.  If the pipes run side by side the time is max(A, B) / 2-ish, if an SMSP is one
// issue port it is (A + B) / 2.
template <int A, int B> __global__ void kmix(uint64_t *out, uint64_t seed) {
    uint64_t a = seed + threadIdx.x, b = seed * 3 + blockIdx.x, c = seed ^ 0x9E3779B97F4A7C15ull, d = a ^ b;
    double x = (double)a, y = (double)b * 1e-9, z = 1.000001, w = 0.999999;
    uint32_t p = (uint32_t)a, q = (uint32_t)b, r = (uint32_t)c, s = (uint32_t)d;
    const int cls = ((threadIdx.x >> 5) & 1) ? B : A;
    if (cls == 0) {
#pragma unroll 1
        for (int i = 0; i < ITER; i++) {
#pragma unroll
            for (int j = 0; j < 8; j++) {
                a = (uint64_t)p * q + a; b = (uint64_t)q * r + b; c = (uint64_t)r * s + c; d = (uint64_t)s * p + d;
                p += (uint32_t)a; q ^= (uint32_t)b;
            }
        }
    } else if (cls == 2) {
#pragma unroll 1
        for (int i = 0; i < ITER; i++) {
#pragma unroll
            for (int j = 0; j < 8; j++) { x = fma(x, z, y); y = fma(y, w, x); z = fma(z, w, 1e-30); w = fma(w, z, 1e-30); }
        }
    } else {
#pragma unroll 1
        for (int i = 0; i < ITER; i++) {
#pragma unroll
            for (int j = 0; j < 8; j++) { p = (p + q) ^ r; q = (q + r) ^ s; r = (r + s) ^ p; s = (s + p) ^ q; }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a ^ b ^ c ^ d ^ (uint64_t)(x + y + z + w) ^ p ^ q ^ r ^ s;
}
template <int A, int B> void runmix(const char *name) {
    uint64_t *out; cudaMalloc(&out, 148 * 8 * 1024 * 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    kmix<A, B><<<148 * 8, 256>>>(out, 12345);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    kmix<A, B><<<148 * 8, 256>>>(out, 12345);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-44s %8.3f ms\n", name, ms);
    cudaFree(out);
}
template <int OP> void run(const char *name, double ops_per_iter) {
    uint64_t *out; cudaMalloc(&out, 148 * 8 * 1024 * 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP><<<148 * 8, 256>>>(out, 12345);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<OP><<<148 * 8, 256>>>(out, 12345);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double total = 148.0 * 8 * 256 * ITER * 8 * ops_per_iter;
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("%-28s %8.3f ms  %7.1f Gop/s  %6.1f op/clk/SM (at %d MHz nominal)\n", name, ms, total / ms / 1e6,
           total / (ms * 1e-3) / 148 / (clk * 1e3), clk / 1000);
    cudaFree(out);
}
int main() {
    run<0>("IMAD.WIDE.U32 (32x32+64)", 4);
    run<1>("IMAD 32-bit", 4);
    run<2>("DFMA", 4);
    run<3>("I2F.F64.S64 + F2I.S64.F64", 4);   // 2 each
    run<4>("__umul64hi (+add)", 2);
    run<5>("IADD+LOP3 pair", 4);
    run<11>("IMAD.HI.U32 (+add)", 4);
    run<6>("I2F.F64.U32 + F2I.S32.F64", 4);
    // mixes: ops counted per class (4 + 4 per inner step); compare the ms with the single-class runs above
    run<7>("IMAD.WIDE x4 + DFMA x4, interleaved", 8);
    run<8>("IMAD.WIDE | DFMA by warp parity", 4);
    run<9>("IMAD.WIDE x4 + ALU x4, interleaved", 8);
    run<10>("DFMA x4 + ALU x4, interleaved", 8);
    // warp-specialised: half of the warps one class, half the other (compare with HALF the single-class times)
    runmix<0, 0>("all warps IMAD.WIDE (kmix)");
    runmix<2, 2>("all warps DFMA (kmix)");
    runmix<5, 5>("all warps ALU (kmix)");
    runmix<0, 2>("even warps IMAD.WIDE | odd warps DFMA");
    runmix<0, 5>("even warps IMAD.WIDE | odd warps ALU");
    runmix<2, 5>("even warps DFMA | odd warps ALU");
    return 0;
}
