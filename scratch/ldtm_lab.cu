// ldtm_lab.cu -- TMEM read bandwidth of tcgen05.ld on B200 (sm_100a): bytes per clock per SM for the 32x32b shape
// with 16 / 32 / 64 / 128 columns per instruction and 4 / 8 / 16 warps per CTA (one CTA per SM, 512 columns).
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o scratch/ldtm_lab scratch/ldtm_lab.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int NCOL>
__device__ __forceinline__ uint32_t ld_cols(uint32_t taddr);

template <>
__device__ __forceinline__ uint32_t ld_cols<16>(uint32_t taddr) {
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) x ^= r[i];
    return x;
}
template <>
__device__ __forceinline__ uint32_t ld_cols<32>(uint32_t taddr) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) x ^= r[i];
    return x;
}
// three x16 loads in flight before one wait (the scoring kernel's pattern)
__device__ __forceinline__ uint32_t ld_3x16(uint32_t taddr) {
    uint32_t r[48];
#pragma unroll
    for (int k = 0; k < 3; ++k)
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(r[16 * k + 0]), "=r"(r[16 * k + 1]), "=r"(r[16 * k + 2]), "=r"(r[16 * k + 3]), "=r"(r[16 * k + 4]),
                       "=r"(r[16 * k + 5]), "=r"(r[16 * k + 6]), "=r"(r[16 * k + 7]), "=r"(r[16 * k + 8]), "=r"(r[16 * k + 9]),
                       "=r"(r[16 * k + 10]), "=r"(r[16 * k + 11]), "=r"(r[16 * k + 12]), "=r"(r[16 * k + 13]),
                       "=r"(r[16 * k + 14]), "=r"(r[16 * k + 15])
                     : "r"(taddr + 64 * k) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 48; ++i) x ^= r[i];
    return x;
}

template <int MODE>   // 0: x16, 1: x32, 2: 3 x x16 per wait
__global__ void __launch_bounds__(512, 1) ldtm_kernel(int iters, unsigned long long* cyc, uint32_t* sink) {
    __shared__ uint32_t tmem_base;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t t = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 16);
    uint32_t x = 0;
    __syncthreads();
    const long long c0 = clock64();
    for (int i = 0; i < iters; ++i) {
        if (MODE == 0) x ^= ld_cols<16>(t + ((i & 3) * 64));
        else if (MODE == 1) x ^= ld_cols<32>(t + ((i & 3) * 64));
        else x ^= ld_3x16(t + ((i & 1) * 192));
    }
    __syncthreads();
    const long long c1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = (unsigned long long)(c1 - c0);
    if (x == 0x12345u) sink[0] = x;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
}

int main() {
    unsigned long long* cyc;
    uint32_t* sink;
    cudaMalloc(&cyc, 148 * 8);
    cudaMalloc(&sink, 4);
    const int iters = 20000;
    for (int mode = 0; mode < 3; ++mode)
        for (int warps = 4; warps <= 16; warps *= 2) {
            for (int rep = 0; rep < 2; ++rep) {
                if (mode == 0) ldtm_kernel<0><<<148, warps * 32>>>(iters, cyc, sink);
                else if (mode == 1) ldtm_kernel<1><<<148, warps * 32>>>(iters, cyc, sink);
                else ldtm_kernel<2><<<148, warps * 32>>>(iters, cyc, sink);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
            }
            unsigned long long h[148];
            cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
            double mean = 0;
            for (int i = 0; i < 148; ++i) mean += (double)h[i];
            mean /= 148;
            const double cols = mode == 0 ? 16 : mode == 1 ? 32 : 48;
            const double bytes = (double)iters * warps * 32 * cols * 4;
            printf("{\"mode\": \"%s\", \"warps\": %d, \"cycles\": %.0f, \"bytes_per_clk_per_sm\": %.1f, \"cycles_per_warp_instr_group\": %.1f}\n",
                   mode == 0 ? "x16" : mode == 1 ? "x32" : "3x16", warps, mean, bytes / mean, mean / iters);
        }
    return 0;
}
