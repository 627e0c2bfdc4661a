// fma_lab.cu -- developer probe: FMA-pipe cost of the operand forms the scoring kernel uses.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scratch/fma_lab scratch/fma_lab.cu && scratch/fma_lab
// Each variant runs 148*4 CTAs x 256 threads (8 warps per SMSP), ITER iterations of 16 independent instructions,
// and reports cycles per warp-instruction per SMSP (1.0 = one instruction per cycle).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float lo, float hi) { u64 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi)); return d; }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ float ffma(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }

template <int V>
__global__ void __launch_bounds__(256) lab(int iters, float seed, float* out, long long* cyc) {
    const float s = seed + 1e-6f * threadIdx.x;
    float x[16], y[16], acc[16];
    u64 p2[8], a2[8], m2[8];
#pragma unroll
    for (int k = 0; k < 16; ++k) { x[k] = s + 0.01f * k; y[k] = s * 0.5f - 0.02f * k; acc[k] = s + k; }
#pragma unroll
    for (int k = 0; k < 8; ++k) { p2[k] = pack2(s + k, s - k); a2[k] = pack2(s * k, s + 2 * k); m2[k] = pack2(x[k], x[k]); }
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        if (V == 0) {            // scalar, 3 distinct registers per instruction, nothing reusable
#pragma unroll
            for (int k = 0; k < 16; ++k) acc[k] = ffma(x[k], y[(k + 5) & 15], acc[k]);
        } else if (V == 1) {     // scalar, multiplier shared by consecutive instructions (reuse cache)
#pragma unroll
            for (int k = 0; k < 16; ++k) acc[k] = ffma(x[0], y[0], acc[k]);
        } else if (V == 2) {     // scalar, d = a*a + c (two distinct registers)
#pragma unroll
            for (int k = 0; k < 16; ++k) acc[k] = ffma(x[k], x[k], acc[k]);
        } else if (V == 3) {     // packed, (broadcast 32, 64, 64): matrix entry x point pair + accumulator pair
#pragma unroll
            for (int k = 0; k < 16; ++k) a2[k & 7] = ffma2(m2[(k * 3) & 7], p2[(k + 1) & 7], a2[k & 7]);
        } else if (V == 4) {     // packed, three full 64-bit operands, all distinct
#pragma unroll
            for (int k = 0; k < 16; ++k) a2[k & 7] = ffma2(p2[(k * 3) & 7], p2[(k * 3 + 1) & 7], a2[k & 7]);
        } else if (V == 5) {     // packed, (broadcast 32, 64, broadcast 32): first FMA of a chain
#pragma unroll
            for (int k = 0; k < 16; ++k) p2[k & 7] = ffma2(m2[(k * 3) & 7], p2[k & 7], m2[(k * 5 + 1) & 7]);
        } else if (V == 8) {     // packed, point pair shared by consecutive instructions (the scoring loop's order)
#pragma unroll
            for (int k = 0; k < 16; ++k) a2[k & 7] = ffma2(m2[(k * 3) & 7], p2[k >> 3], a2[k & 7]);
        } else if (V == 9) {     // packed, matrix entry shared by consecutive instructions (points vary)
#pragma unroll
            for (int k = 0; k < 16; ++k) a2[k & 7] = ffma2(m2[k >> 3], p2[(k * 3) & 7], a2[k & 7]);
        } else if (V == 10) {    // scalar, (fresh, shared, fresh): outer-product order
#pragma unroll
            for (int k = 0; k < 16; ++k) acc[k] = ffma(x[k], y[k >> 3], acc[k]);
        } else if (V == 11) {    // scalar tail form: d = n*w + x with w shared by two consecutive instructions
#pragma unroll
            for (int k = 0; k < 16; ++k) acc[k] = ffma(x[k], y[k >> 1], acc[k]);
        } else if (V == 6) {     // packed, d = a*a + c (two distinct 64-bit registers)
#pragma unroll
            for (int k = 0; k < 16; ++k) a2[k & 7] = ffma2(p2[k & 7], p2[k & 7], a2[k & 7]);
        } else if (V == 7) {     // scalar with two operands from one aligned 64-bit pair: d = lo*hi + c
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                float lo, hi; asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p2[k & 7]));
                acc[k] = ffma(lo, hi, acc[k]);
            }
        }
    }
    const long long t1 = clock64();
    float r = 0.f; u64 z = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) r += acc[k];
#pragma unroll
    for (int k = 0; k < 8; ++k) z ^= a2[k] ^ p2[k];
    if (r == 123.456f || z == 0x1234567812345678ull) out[0] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int V>
void run(const char* name, float* out, long long* cyc, int cps = 4) {
    const int iters = 4096, ctas = 148 * cps;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    lab<V><<<ctas, 256>>>(iters, 0.5f, out, cyc);
    cudaEventRecord(e0);
    lab<V><<<ctas, 256>>>(iters, 0.5f, out, cyc);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    // 4 CTAs x 8 warps per SM = 8 warps per SMSP, each iters*16 instructions
    (void)c;
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double per = ms * 1e-3 * khz * 1e3 / (2.0 * cps * iters * 16.0);   // 2*cps warps per SMSP
    printf("%-58s %d warps/SMSP %7.3f ms  %6.3f cycles per warp-instruction per SMSP (at %d MHz)\n", name, 2 * cps, ms, per, khz / 1000);
}

int main() {
    float* out; long long* cyc; cudaMalloc(&out, 4); cudaMalloc(&cyc, 8);
    run<0>("V0 FFMA  3 distinct registers", out, cyc);
    run<1>("V1 FFMA  shared multiplier pair (reuse)", out, cyc);
    run<2>("V2 FFMA  a*a+c", out, cyc);
    run<3>("V3 FFMA2 (bcast32, 64, 64)", out, cyc);
    run<4>("V4 FFMA2 (64, 64, 64) distinct", out, cyc);
    run<5>("V5 FFMA2 (bcast32, 64, bcast32)", out, cyc);
    run<6>("V6 FFMA2 a*a+c", out, cyc);
    run<7>("V7 FFMA  lo*hi+c from one 64-bit pair", out, cyc);
    run<8>("V8 FFMA2 (bcast32, 64 shared by 8 in a row, 64)", out, cyc);
    run<9>("V9 FFMA2 (bcast32 shared by 8 in a row, 64, 64)", out, cyc);
    run<10>("V10 FFMA (fresh, shared by 8 in a row, fresh)", out, cyc);
    run<11>("V11 FFMA (fresh, shared by 2 in a row, fresh)", out, cyc);
    for (int cps = 1; cps <= 3; ++cps) {
        run<0>("V0 FFMA  3 distinct registers", out, cyc, cps);
        run<3>("V3 FFMA2 (bcast32, 64, 64)", out, cyc, cps);
        run<8>("V8 FFMA2 (bcast32, 64 shared, 64)", out, cyc, cps);
        run<10>("V10 FFMA (fresh, shared, fresh)", out, cyc, cps);
    }
    return cudaDeviceSynchronize() != cudaSuccess;
}
