// tc_debug.cu -- developer build of the tensor-core nearest-neighbour TU with its probes (NOT part of libposeb200.so):
// the per-CTA timeline / cycle counters of knn_tc_kernel (TCDBG / TCCLK in csrc/knn_tc.cu, compiled in by PB_TC_DEBUG)
// and the tcgen05.mma issue-rate probe.  Build a debug library by compiling this file INSTEAD of csrc/knn_tc.cu:
//   nvcc <flags of pose_estimation_b200/build.py> -c scratch/tc_debug.cu -o pose_estimation_b200/build/knn_tc.o  (then relink)
// scratch/tc_mma_probe.py and scratch/tc_timeline.py use the exports below.
#define PB_TC_DEBUG 1
#include "../pose_estimation_b200/csrc/knn_tc.cu"

namespace {
// MMA issue-rate probe: reps x (two K=16 MMAs, M=128, N=n) on garbage operands, one commit at the end.
__global__ void __launch_bounds__(128) tc_mma_probe_kernel(int n, uint32_t layout, uint32_t lbo, uint32_t sbo, uint32_t kstep,
                                                            int reps, int nacc, long long* out) {
    extern __shared__ __align__(1024) unsigned char sm[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 64 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(sm)[i] = 0x3c003c00u;
    if (tid == 0) { tc::mbar_init(&bar, 1); tc::fence_mbar_init(); }
    if (warp == 0) tc::tmem_alloc<512>(&tmem_base);
    tc::fence_proxy_async();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = tmem_base;
    if (tid == 0) {
        const uint32_t idesc = tc::idesc_bf16_f32(128, n);
        const uint64_t ad = tc::smem_desc(tc::smem_u32(sm), lbo, sbo) | ((uint64_t)layout << 61);
        const uint64_t bd = tc::smem_desc(tc::smem_u32(sm) + 32768, lbo, sbo) | ((uint64_t)layout << 61);
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            const uint32_t d = tmem + (uint32_t)(r % nacc) * (uint32_t)n;
            tc::mma_bf16(d, ad, bd, idesc, 0);
            tc::mma_bf16(d, ad + (kstep >> 4), bd + (kstep >> 4), idesc, 1);
        }
        const long long t1 = clock64();
        tc::mma_commit(&bar);
        tc::mbar_wait(&bar, 0);
        const long long t2 = clock64();
        if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    } 
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<512>(tmem);
}
}  // namespace

extern "C" int pb_tc_probe_mma(int ctas, int n, unsigned layout, unsigned lbo, unsigned sbo, unsigned kstep, int reps, int nacc,
                               long long* host_out) {
    long long* d;
    cudaMalloc(&d, 16);
    cudaFuncSetAttribute(tc_mma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    tc_mma_probe_kernel<<<ctas, 128, 64 * 1024>>>(n, layout, lbo, sbo, kstep, reps, nacc, d);
    int rc = (int)cudaDeviceSynchronize();
    cudaMemcpy(host_out, d, 16, cudaMemcpyDeviceToHost);
    cudaFree(d);
    return rc;
}
extern "C" int pb_tc_debug_read(unsigned long long* host) {
    return (int)cudaMemcpyFromSymbol(host, g_tcdbg, sizeof(unsigned long long) * 256 * 8);
}
extern "C" int pb_tc_debug_read_cycles(unsigned long long* host) {
    return (int)cudaMemcpyFromSymbol(host, g_tccyc, sizeof(unsigned long long) * 256 * 8);
}
