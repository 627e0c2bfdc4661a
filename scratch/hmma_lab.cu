// hmma_lab.cu -- throughput of the legacy warp-level tensor path (mma.sync.m16n8k16, f16 in, f32 accumulate) on B200:
// cycles per instruction per SM sub-partition with 1 / 2 / 4 / 6 warps per scheduler and 3 / 6 / 12 independent
// accumulators.  Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o scratch/hmma_lab scratch/hmma_lab.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int NACC>
__global__ void __launch_bounds__(1024) hmma_kernel(int iters, unsigned long long* cyc, float* sink) {
    uint32_t a[4] = {0x3c003c00u + threadIdx.x, 0x3c003c00u, 0x38003800u, 0x34003400u};
    uint32_t b[2] = {0x3c003c00u, 0x38003800u + threadIdx.x};
    float c[NACC][4];
#pragma unroll
    for (int k = 0; k < NACC; ++k) c[k][0] = c[k][1] = c[k][2] = c[k][3] = 0.0f;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < NACC; ++k)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[k][0]), "+f"(c[k][1]), "+f"(c[k][2]), "+f"(c[k][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = (unsigned long long)(t1 - t0);
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < NACC; ++k) s += c[k][0] + c[k][1] + c[k][2] + c[k][3];
    if (s == 12345.0f) sink[0] = s;
}

template <int NACC>
void run(int warps, unsigned long long* cyc, float* sink) {
    const int iters = 4000;
    for (int rep = 0; rep < 2; ++rep) {
        hmma_kernel<NACC><<<148, warps * 32>>>(iters, cyc, sink);
        cudaDeviceSynchronize();
    }
    unsigned long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < 148; ++i) mean += (double)h[i];
    mean /= 148;
    const double per_smsp = (double)iters * NACC * warps / 4.0;       // instructions per scheduler
    printf("{\"warps_per_sm\": %d, \"nacc\": %d, \"cycles_per_hmma_per_scheduler\": %.2f, \"tflops_chip_at_1965MHz\": %.0f}\n", warps, NACC,
           mean / per_smsp, 4096.0 * 4.0 / (mean / per_smsp) * 148 * 1.965e9 / 1e12);
}

int main() {
    unsigned long long* cyc; float* sink;
    cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 4);
    for (int warps : {4, 8, 16, 24}) { run<3>(warps, cyc, sink); run<6>(warps, cyc, sink); run<12>(warps, cyc, sink); }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
