// nn.cuh -- exact FP32 nearest-neighbour scan shared by knn.cu and metrics.cu.
#pragma once
#include "common.cuh"

namespace {

constexpr int KT = 128;    // threads per CTA
constexpr int QPT = 4;     // queries per thread
constexpr int TY = 1024;   // database points per shared-memory tile

struct LoadXYZ {  // database points stored [M,3]
    const float* y;
    __device__ __forceinline__ float4 operator()(int j) const {
        const float* p = y + 3 * (size_t)j;
        return make_float4(p[0], p[1], p[2], 0.0f);
    }
};

struct LoadTransformed {  // database points = R * model_j + t, computed while staging
    const float* m; float R[9]; float t[3];
    __device__ __forceinline__ float4 operator()(int j) const {
        const float* p = m + 3 * (size_t)j;
        const float x = p[0], y = p[1], z = p[2];
        return make_float4(fmaf(R[0], x, fmaf(R[1], y, fmaf(R[2], z, t[0]))),
                           fmaf(R[3], x, fmaf(R[4], y, fmaf(R[5], z, t[1]))),
                           fmaf(R[6], x, fmaf(R[7], y, fmaf(R[8], z, t[2]))), 0.0f);
    }
};

// For QPT register-resident queries per thread, scan all M database points (staged as float4 in
// shared memory, one broadcast LDS.128 per point).  Strict '<' keeps the lowest index on ties.
template <bool WANT_IDX, class Load>
__device__ __forceinline__ void nn_scan(const Load& load, int M, float4* tile, const float qx[QPT],
                                        const float qy[QPT], const float qz[QPT], float best[QPT], int bidx[QPT]) {
    for (int j0 = 0; j0 < M; j0 += TY) {
        const int n = min(TY, M - j0);
        __syncthreads();
        for (int j = threadIdx.x; j < n; j += blockDim.x) tile[j] = load(j0 + j);
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < n; ++j) {
            const float4 p = tile[j];
#pragma unroll
            for (int q = 0; q < QPT; ++q) {
                const float d = sqdist3(qx[q], qy[q], qz[q], p.x, p.y, p.z);
                if (d < best[q]) {
                    best[q] = d;
                    if (WANT_IDX) bidx[q] = j0 + j;
                }
            }
        }
    }
}

// deterministic block sum (KT threads): butterfly per warp, then warps in order
__device__ __forceinline__ float block_sum_kt(float v, float* red) {
    v = warp_sum_bfly(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float s = red[0];
    for (int w = 1; w < KT / 32; ++w) s = s + red[w];
    return s;
}

}  // namespace
