// k3_lab.cu -- developer probe: the scoring kernel's inner loop in isolation (points resident in shared memory),
// in several instruction forms, to see what the FMA pipe / register file sustains.
//   nvcc -O3 -fmad=false -gencode arch=compute_100a,code=sm_100a -o scratch/k3_lab scratch/k3_lab.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 fmul2(f32x2 a, f32x2 b) { f32x2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi)); return d; }
__device__ __forceinline__ float lo32(f32x2 v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float hi32(f32x2 v) { return __uint_as_float((uint32_t)(v >> 32)); }

constexpr int NP = 1024;     // points in shared memory
constexpr int ST = 128;

// ---- L0: the shipped form (9 packed matrix FMAs per point pair, scalar tail)
__device__ __forceinline__ void inlier_s2(const f32x2 m[12], f32x2 X, f32x2 Y, f32x2 Z, f32x2 nus, f32x2 nvs, uint32_t& s0, uint32_t& s1) {
    const f32x2 x = ffma2(m[0], X, ffma2(m[1], Y, ffma2(m[2], Z, m[3])));
    const f32x2 y = ffma2(m[4], X, ffma2(m[5], Y, ffma2(m[6], Z, m[7])));
    const f32x2 w = ffma2(m[8], X, ffma2(m[9], Y, ffma2(m[10], Z, m[11])));
    const float w0 = lo32(w), dx0 = fmaf(lo32(nus), w0, lo32(x)), dy0 = fmaf(lo32(nvs), w0, lo32(y));
    const float nq0 = w0 * -fabsf(w0);
    s0 = __float_as_uint(fmaf(dx0, dx0, fmaf(dy0, dy0, nq0)));
    const float w1 = hi32(w), dx1 = fmaf(hi32(nus), w1, hi32(x)), dy1 = fmaf(hi32(nvs), w1, hi32(y));
    const float nq1 = w1 * -fabsf(w1);
    s1 = __float_as_uint(fmaf(dx1, dx1, fmaf(dy1, dy1, nq1)));
}
// ---- L3: tail packed as well
__device__ __forceinline__ void inlier_s2p(const f32x2 m[12], f32x2 X, f32x2 Y, f32x2 Z, f32x2 nus, f32x2 nvs, uint32_t& s0, uint32_t& s1) {
    const f32x2 x = ffma2(m[0], X, ffma2(m[1], Y, ffma2(m[2], Z, m[3])));
    const f32x2 y = ffma2(m[4], X, ffma2(m[5], Y, ffma2(m[6], Z, m[7])));
    const f32x2 w = ffma2(m[8], X, ffma2(m[9], Y, ffma2(m[10], Z, m[11])));
    const f32x2 dx = ffma2(nus, w, x), dy = ffma2(nvs, w, y);
    const f32x2 nq = pack2(lo32(w) * -fabsf(lo32(w)), hi32(w) * -fabsf(hi32(w)));
    const f32x2 s = ffma2(dx, dx, ffma2(dy, dy, nq));
    s0 = (uint32_t)s; s1 = (uint32_t)(s >> 32);
}
__device__ __forceinline__ float inlier_s(const float M[12], float X, float Y, float Z, float nus, float nvs) {
    const float x = fmaf(M[0], X, fmaf(M[1], Y, fmaf(M[2], Z, M[3])));
    const float y = fmaf(M[4], X, fmaf(M[5], Y, fmaf(M[6], Z, M[7])));
    const float w = fmaf(M[8], X, fmaf(M[9], Y, fmaf(M[10], Z, M[11])));
    const float dx = fmaf(nus, w, x), dy = fmaf(nvs, w, y);
    const float nq = w * -fabsf(w);
    return fmaf(dx, dx, fmaf(dy, dy, nq));
}

template <int V, int MINB>
__global__ void __launch_bounds__(ST, MINB) lab(const float* __restrict__ pts, const float* __restrict__ hyp, int reps, int* out) {
    __shared__ __align__(16) float sm[NP * 5];
    const int tid = threadIdx.x;
    // pair layout for the packed forms: sA[i] = (X0,X1,Y0,Y1), sB[i] = (Z0,Z1,nU0,nU1), sC[i] = (nV0,nV1)
    // SoA layout for the scalar forms: X[NP] Y[NP] Z[NP] nU[NP] nV[NP]
    for (int i = tid; i < NP; i += ST) {
        const float X = pts[i], Y = pts[NP + i], Z = pts[2 * NP + i], U = -pts[3 * NP + i], Vv = -pts[4 * NP + i];
        if (V == 0 || V == 3) {
            const int o4 = (i >> 1) * 4 + (i & 1);
            sm[o4] = X; sm[o4 + 2] = Y; sm[2 * NP + o4] = Z; sm[2 * NP + o4 + 2] = U; sm[4 * NP + i] = Vv;
        } else {
            sm[i] = X; sm[NP + i] = Y; sm[2 * NP + i] = Z; sm[3 * NP + i] = U; sm[4 * NP + i] = Vv;
        }
    }
    float mj[2][12];
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int k = 0; k < 12; ++k) mj[j][k] = hyp[((blockIdx.x * 2 + j) * ST + tid) % 4096 * 12 + k];
    __syncthreads();
    int c[2] = {0, 0};
    if (V == 0 || V == 3) {
        f32x2 m[2][12];
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int k = 0; k < 12; ++k) m[j][k] = pack2(mj[j][k], mj[j][k]);
        const ulonglong2* a2 = reinterpret_cast<const ulonglong2*>(sm);
        const ulonglong2* b2 = reinterpret_cast<const ulonglong2*>(sm + 2 * NP);
        const ulonglong2* c2 = reinterpret_cast<const ulonglong2*>(sm + 4 * NP);
        for (int r = 0; r < reps; ++r) {
#pragma unroll 1
            for (int i = 0; i < NP / 2; i += 4) {
                const ulonglong2 v01 = c2[i >> 1], v23 = c2[(i >> 1) + 1];
                const f32x2 nvs[4] = {v01.x, v01.y, v23.x, v23.y};
                uint32_t acc[2] = {0u, 0u};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const ulonglong2 pa = a2[i + k], pbv = b2[i + k];
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        uint32_t s0, s1;
                        if (V == 0) inlier_s2(m[j], pa.x, pa.y, pbv.x, pbv.y, nvs[k], s0, s1);
                        else inlier_s2p(m[j], pa.x, pa.y, pbv.x, pbv.y, nvs[k], s0, s1);
                        acc[j] = __funnelshift_l(s0, acc[j], 1);
                        acc[j] = __funnelshift_l(s1, acc[j], 1);
                    }
                }
                c[0] += __popc(acc[0]); c[1] += __popc(acc[1]);
            }
        }
    } else if (V == 2) {     // scalar, per-point helper, compiler's order; 8 points per iteration
        const float4* X4 = reinterpret_cast<const float4*>(sm);
        for (int r = 0; r < reps; ++r) {
#pragma unroll 1
            for (int i = 0; i < NP / 4; i += 2) {
                uint32_t acc[2] = {0u, 0u};
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const float4 x = X4[i + h], y = X4[NP / 4 + i + h], z = X4[2 * NP / 4 + i + h], u = X4[3 * NP / 4 + i + h], v = X4[NP + i + h];
                    const float xs[4] = {x.x, x.y, x.z, x.w}, ys[4] = {y.x, y.y, y.z, y.w}, zs[4] = {z.x, z.y, z.z, z.w};
                    const float us[4] = {u.x, u.y, u.z, u.w}, vs[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int j = 0; j < 2; ++j)
                            acc[j] = __funnelshift_l(__float_as_uint(inlier_s(mj[j], xs[p], ys[p], zs[p], us[p], vs[p])), acc[j], 1);
                }
                c[0] += __popc(acc[0]); c[1] += __popc(acc[1]);
            }
        }
    } else if (V == 1) {     // scalar, outer-product order: 4 points x 6 rows, coordinate-major so one point value is shared
        const float4* X4 = reinterpret_cast<const float4*>(sm);
        for (int r = 0; r < reps; ++r) {
#pragma unroll 1
            for (int i = 0; i < NP / 4; i += 2) {
                uint32_t acc[2] = {0u, 0u};
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const float4 x = X4[i + h], y = X4[NP / 4 + i + h], z = X4[2 * NP / 4 + i + h], u = X4[3 * NP / 4 + i + h], v = X4[NP + i + h];
                    const float xs[4] = {x.x, x.y, x.z, x.w}, ys[4] = {y.x, y.y, y.z, y.w}, zs[4] = {z.x, z.y, z.z, z.w};
                    const float us[4] = {u.x, u.y, u.z, u.w}, vs[4] = {v.x, v.y, v.z, v.w};
                    float a[4][6];
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int q = 0; q < 6; ++q) a[p][q] = fmaf(mj[q / 3][(q % 3) * 4 + 2], zs[p], mj[q / 3][(q % 3) * 4 + 3]);
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int q = 0; q < 6; ++q) a[p][q] = fmaf(mj[q / 3][(q % 3) * 4 + 1], ys[p], a[p][q]);
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int q = 0; q < 6; ++q) a[p][q] = fmaf(mj[q / 3][(q % 3) * 4 + 0], xs[p], a[p][q]);
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            const float w = a[p][j * 3 + 2];
                            const float dx = fmaf(us[p], w, a[p][j * 3]), dy = fmaf(vs[p], w, a[p][j * 3 + 1]);
                            const float nq = w * -fabsf(w);
                            acc[j] = __funnelshift_l(__float_as_uint(fmaf(dx, dx, fmaf(dy, dy, nq))), acc[j], 1);
                        }
                }
                c[0] += __popc(acc[0]); c[1] += __popc(acc[1]);
            }
        }
    }
    if (c[0] + c[1] == -7) out[0] = c[0];
    atomicAdd(out + 1 + (V & 3), c[0] + c[1]);
}

template <int V, int MINB>
void run(const char* name, const float* pts, const float* hyp, int* out) {
    const int reps = 8, ctas = 148 * MINB;
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, lab<V, MINB>);
    int occ = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, lab<V, MINB>, ST, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaMemset(out, 0, 64);
    lab<V, MINB><<<ctas, ST>>>(pts, hyp, reps, out);
    cudaEventRecord(e0);
    lab<V, MINB><<<ctas, ST>>>(pts, hyp, reps, out);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    int h[8]; cudaMemcpy(h, out, 32, cudaMemcpyDeviceToHost);
    // per SMSP: MINB warps, each reps * NP points * 2 hypotheses
    const double hp = (double)MINB * reps * NP * 2;
    printf("%-44s regs %3d  CTAs/SM %d (occ %d)  %7.3f ms  %6.2f cycles per warp-level hyp*pt (ideal 14)  count %d\n",
           name, fa.numRegs, MINB, occ, ms, ms * 1e-3 * khz * 1e3 / hp, h[1 + (V & 3)]);
}

int main() {
    float *pts, *hyp; int* out;
    cudaMalloc(&pts, NP * 5 * 4); cudaMalloc(&hyp, 4096 * 12 * 4); cudaMalloc(&out, 64);
    float* hp = new float[NP * 5]; float* hh = new float[4096 * 12];
    unsigned s = 12345u;
    auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (float)(s >> 8) / 16777216.0f; };
    for (int i = 0; i < NP; ++i) { hp[i] = rnd() * 0.1f - 0.05f; hp[NP + i] = rnd() * 0.1f - 0.05f; hp[2 * NP + i] = rnd() * 0.1f - 0.05f; hp[3 * NP + i] = 300 + rnd() * 40; hp[4 * NP + i] = 220 + rnd() * 40; }
    for (int i = 0; i < 4096; ++i) {
        float* m = hh + i * 12;
        for (int k = 0; k < 12; ++k) m[k] = rnd() * 2 - 1;
        m[0] = 572 + rnd(); m[5] = 573 + rnd(); m[3] = 320 * 0.9f; m[7] = 240 * 0.9f; m[8] = m[9] = 0.01f; m[10] = 1; m[11] = 0.9f;
        m[2] = 325; m[6] = 242;
    }
    cudaMemcpy(pts, hp, NP * 5 * 4, cudaMemcpyHostToDevice); cudaMemcpy(hyp, hh, 4096 * 12 * 4, cudaMemcpyHostToDevice);
    run<0, 6>("L0 shipped: packed matrix, scalar tail", pts, hyp, out);
    run<0, 4>("L0 shipped: packed matrix, scalar tail", pts, hyp, out);
    run<0, 8>("L0 shipped: packed matrix, scalar tail", pts, hyp, out);
    run<3, 6>("L3 all packed", pts, hyp, out);
    run<3, 8>("L3 all packed", pts, hyp, out);
    run<2, 6>("L2 scalar, compiler's order", pts, hyp, out);
    run<2, 8>("L2 scalar, compiler's order", pts, hyp, out);
    run<1, 6>("L1 scalar, outer-product order", pts, hyp, out);
    run<1, 4>("L1 scalar, outer-product order", pts, hyp, out);
    run<1, 8>("L1 scalar, outer-product order", pts, hyp, out);
    return cudaDeviceSynchronize() != cudaSuccess;
}
