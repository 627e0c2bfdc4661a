// hmma_acc_lab.cu -- how accurately does the tensor core add the 16 products of an FP16 x FP16 -> FP32 dot product?
// For random operand rows with a wide spread of magnitudes (and an adversarial family: one large term, many small ones)
// the m16n8k16 result is compared with the exactly rounded sum (products of two FP16 are exact in FP64; the sum of 16 of
// them is accumulated in FP64, whose error is far below FP32's).  Reports the largest |d - exact| / sum|terms| in units
// of u = 2^-24 -- the quantity the K3 band allows 16 u for (csrc/score_tc_common.cuh).
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o scratch/hmma_acc_lab scratch/hmma_acc_lab.cu
#include <cstdio>
#include <cstdint>
#include <cmath>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t rng(uint32_t& s) { s ^= s << 13; s ^= s >> 17; s ^= s << 5; return s; }
__device__ __forceinline__ float rnd_val(uint32_t& s, int spread) {       // +-2^e * (1 + m), e uniform in [-spread, 0]
    const uint32_t r = rng(s);
    const int e = -(int)(r % (uint32_t)(spread + 1));
    const float m = 1.0f + (float)((r >> 8) & 0x3ff) / 1024.0f;
    return ((r >> 31) ? -1.0f : 1.0f) * ldexpf(m, e);
}

// every warp: its own 16 x 16 A and 16 x 8 B tiles (in shared memory), one HMMA, exact reference per output element
__global__ void __launch_bounds__(128) acc_kernel(int iters, int spread, int family, double* out_max, unsigned long long* out_n) {
    __shared__ __half sA[4][16][16];
    __shared__ __half sB[4][8][16];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    uint32_t s = 0x9E3779B9u * (blockIdx.x * 128 + threadIdx.x + 1) + 12345u;
    double worst = 0.0;
    unsigned long long n = 0;
    for (int it = 0; it < iters; ++it) {
        for (int e = lane; e < 16 * 16; e += 32) {
            float v = rnd_val(s, spread);
            if (family == 1 && (e & 15) == 0) v *= 1024.0f;            // one dominant term per row
            sA[warp][e >> 4][e & 15] = __float2half_rn(v);
        }
        for (int e = lane; e < 8 * 16; e += 32) sB[warp][e >> 4][e & 15] = __float2half_rn(rnd_val(s, family == 2 ? 0 : spread));
        __syncwarp();
        uint32_t a[4], b[2];
        a[0] = *reinterpret_cast<const uint32_t*>(&sA[warp][g][2 * t]);
        a[1] = *reinterpret_cast<const uint32_t*>(&sA[warp][g + 8][2 * t]);
        a[2] = *reinterpret_cast<const uint32_t*>(&sA[warp][g][2 * t + 8]);
        a[3] = *reinterpret_cast<const uint32_t*>(&sA[warp][g + 8][2 * t + 8]);
        b[0] = *reinterpret_cast<const uint32_t*>(&sB[warp][g][2 * t]);
        b[1] = *reinterpret_cast<const uint32_t*>(&sB[warp][g][2 * t + 8]);
        float d[4];
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                     : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]), "f"(0.0f));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int row = g + 8 * (k >> 1), col = 2 * t + (k & 1);
            double ex = 0.0, sa = 0.0;
            for (int kk = 0; kk < 16; ++kk) {
                const double p = (double)__half2float(sA[warp][row][kk]) * (double)__half2float(sB[warp][col][kk]);
                ex += p; sa += fabs(p);
            }
            const double r = fabs((double)d[k] - ex) / sa * 16777216.0;   // in units of 2^-24 of sum|terms|
            worst = fmax(worst, r);
            ++n;
        }
        __syncwarp();
    }
    for (int off = 16; off >= 1; off >>= 1) worst = fmax(worst, __shfl_xor_sync(0xffffffffu, worst, off));
    if (lane == 0) {
        atomicMax(reinterpret_cast<unsigned long long*>(out_max), (unsigned long long)__double_as_longlong(worst));
        atomicAdd(out_n, n * 32);
    }
}

int main() {
    double* dmax; unsigned long long* dn;
    cudaMalloc(&dmax, 8); cudaMalloc(&dn, 8);
    const char* names[3] = {"random magnitudes", "one dominant term per row", "A spread, B near 1"};
    for (int family = 0; family < 3; ++family)
        for (int spread : {0, 4, 12, 24}) {
            cudaMemset(dmax, 0, 8); cudaMemset(dn, 0, 8);
            acc_kernel<<<592, 128>>>(400, spread, family, dmax, dn);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            double h; unsigned long long n;
            cudaMemcpy(&h, dmax, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&n, dn, 8, cudaMemcpyDeviceToHost);
            printf("{\"family\": \"%s\", \"exponent_spread\": %d, \"dot_products\": %llu, \"max_error_in_u_of_sum_abs_terms\": %.3f}\n",
                   names[family], spread, n, h);
        }
    return 0;
}
