// knn_tc.cu -- K5 on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
#include "common.cuh"
#include "tc.cuh"

namespace {

// ---- self test: D[128x128] = A[128x32] * B[128x32]^T, operands given in the canonical tile layout
// [kc 0..3][row 0..127][8 bf16]; exercises descriptors, two accumulating K=16 MMAs, commit, tcgen05.ld.
__global__ void __launch_bounds__(128) tc_selftest_kernel(const uint4* __restrict__ A, const uint4* __restrict__ B,
                                                          float* __restrict__ D, int* __restrict__ err) {
    __shared__ __align__(128) uint4 sA[4 * 128];
    __shared__ __align__(128) uint4 sB[4 * 128];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 4 * 128; i += 128) { sA[i] = A[i]; sB[i] = B[i]; }
    if (tid == 0) { tc::mbar_init(&bar, 1); tc::fence_mbar_init(); }
    if (warp == 0) tc::tmem_alloc<128>(&tmem_base);
    tc::fence_proxy_async();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = tmem_base;
    if (tid == 0) {
        const uint32_t idesc = tc::idesc_bf16_f32(128, 128);
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const uint64_t ad = tc::smem_desc(tc::smem_u32(sA) + k * 2 * 2048, 2048, 128);
            const uint64_t bd = tc::smem_desc(tc::smem_u32(sB) + k * 2 * 2048, 2048, 128);
            tc::mma_bf16(tmem, ad, bd, idesc, k > 0);
        }
        tc::mma_commit(&bar);
    }
    if (!tc::mbar_wait(&bar, 0)) { if (tid == 0) *err = 1; }
    tc::fence_after_sync();
#pragma unroll 1
    for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t r[32];
        tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, r);
        tc::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) D[tid * 128 + c0 + j] = __uint_as_float(r[j]);
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<128>(tmem);
}

}  // namespace

extern "C" int pb_tc_selftest(const void* A, const void* B, float* D, int* err, void* stream) {
    PB_REQUIRE(A && B && D && err, "pb_tc_selftest: null pointer argument");
    tc_selftest_kernel<<<1, 128, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const uint4*>(A),
                                                                          static_cast<const uint4*>(B), D, err);
    return pb_check_launch("tc_selftest_kernel") ? -2 : 0;
}
