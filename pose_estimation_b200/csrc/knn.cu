// knn.cu -- K5: nearest neighbour (K=1) and ADD-S, exact FP32 (sm_100a).
//
// pb_knn1 replaces the KeOps handle generic_argkmin('SqDist(x, y)', 'a = Vi(1)', 'x = Vi(3)',
// 'y = Vj(3)') (train.py:126, lib/network/loss.py:34, version/transparent/lib/utils.py:24).
// pb_add_adds replaces Metric.cal_adds_cuda (lib/utils/metric.py:17-36), which materialises two
// [N,N,3] tensors; here the N x N distances never leave registers.
// FP32-pipe bound: database points are staged as float4 in shared memory and read as one
// broadcast LDS.128 per QPT pairs; each thread keeps QPT queries in registers.
#include "common.cuh"
#include "nn.cuh"

extern "C" size_t pb_knn1_tc_workspace_bytes(int B, int N, int M);

namespace {

__global__ void __launch_bounds__(KT) knn1_kernel(const float* __restrict__ x, const float* __restrict__ y, int N,
                                                  int M, int y_shared, int64_t* __restrict__ idx,
                                                  float* __restrict__ d2) {
    __shared__ float4 tile[TY];
    const int b = blockIdx.y;
    const float* xb = x + (size_t)b * N * 3;
    const float* yb = y + (y_shared ? 0 : (size_t)b * M * 3);
    float qx[QPT], qy[QPT], qz[QPT], best[QPT];
    int bidx[QPT];
    const int i0 = blockIdx.x * (KT * QPT) + threadIdx.x;
#pragma unroll
    for (int q = 0; q < QPT; ++q) {
        const int i = i0 + q * KT;
        const bool ok = i < N;
        qx[q] = ok ? xb[3 * (size_t)i] : 0.0f;
        qy[q] = ok ? xb[3 * (size_t)i + 1] : 0.0f;
        qz[q] = ok ? xb[3 * (size_t)i + 2] : 0.0f;
        best[q] = INFINITY; bidx[q] = 0;
    }
    nn_scan<true>(LoadXYZ{yb}, M, tile, qx, qy, qz, best, bidx);
#pragma unroll
    for (int q = 0; q < QPT; ++q) {
        const int i = i0 + q * KT;
        if (i < N) {
            idx[(size_t)b * N + i] = bidx[q];
            if (d2) d2[(size_t)b * N + i] = best[q];
        }
    }
}

// one CTA per batch element: ADD and (optionally) ADD-S
__global__ void __launch_bounds__(KT) add_adds_kernel(const float* __restrict__ pred, const float* __restrict__ target,
                                                      int N, const uint8_t* __restrict__ want_adds,
                                                      float* __restrict__ add, float* __restrict__ adds) {
    __shared__ float4 tile[TY];
    __shared__ float red[KT / 32];
    const int b = blockIdx.x;
    const float* pb = pred + (size_t)b * N * 3;
    const float* tb = target + (size_t)b * N * 3;
    float s = 0.0f;
    for (int i = threadIdx.x; i < N; i += KT)
        s += sqrtf(sqdist3(pb[3 * (size_t)i], pb[3 * (size_t)i + 1], pb[3 * (size_t)i + 2], tb[3 * (size_t)i],
                           tb[3 * (size_t)i + 1], tb[3 * (size_t)i + 2]));
    const float add_v = block_sum_kt(s, red) / (float)N;
    if (threadIdx.x == 0 && add) add[b] = add_v;
    if (!adds) return;
    const bool want = want_adds ? (want_adds[b] != 0) : true;
    if (!want) {
        if (threadIdx.x == 0) adds[b] = add_v;
        return;
    }
    float s2 = 0.0f;
    for (int base = 0; base < N; base += KT * QPT) {
        float qx[QPT], qy[QPT], qz[QPT], best[QPT];
        int bidx[QPT];
#pragma unroll
        for (int q = 0; q < QPT; ++q) {
            const int i = base + q * KT + threadIdx.x;
            const bool ok = i < N;
            qx[q] = ok ? tb[3 * (size_t)i] : 0.0f;
            qy[q] = ok ? tb[3 * (size_t)i + 1] : 0.0f;
            qz[q] = ok ? tb[3 * (size_t)i + 2] : 0.0f;
            best[q] = INFINITY; bidx[q] = 0;
        }
        nn_scan<false>(LoadXYZ{pb}, N, tile, qx, qy, qz, best, bidx);
#pragma unroll
        for (int q = 0; q < QPT; ++q)
            if (base + q * KT + threadIdx.x < N) s2 += sqrtf(best[q]);
    }
    const float adds_v = block_sum_kt(s2, red) / (float)N;
    if (threadIdx.x == 0) adds[b] = adds_v;
}

__global__ void transform_kernel(const float* __restrict__ model, int model_shared, const float* __restrict__ R,
                                 const float* __restrict__ t, int N, float* __restrict__ out) {
    const int b = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float* m = model + (model_shared ? 0 : (size_t)b * N * 3) + 3 * (size_t)i;
    const float* Rb = R + 9 * (size_t)b;
    const float* tb = t + 3 * (size_t)b;
    const float x = m[0], y = m[1], z = m[2];
    float* o = out + ((size_t)b * N + i) * 3;
    o[0] = fmaf(Rb[0], x, fmaf(Rb[1], y, fmaf(Rb[2], z, tb[0])));
    o[1] = fmaf(Rb[3], x, fmaf(Rb[4], y, fmaf(Rb[5], z, tb[1])));
    o[2] = fmaf(Rb[6], x, fmaf(Rb[7], y, fmaf(Rb[8], z, tb[2])));
}

}  // namespace

extern "C" int pb_knn1(const float* x, const float* y, int B, int N, int M, int y_shared, int64_t* idx, float* d2,
                       void* stream) {
    PB_REQUIRE(B >= 0 && N >= 0 && M > 0, "pb_knn1: bad sizes B=%d N=%d M=%d", B, N, M);
    PB_REQUIRE(B <= 65535, "pb_knn1: B too large (<= 65535)");
    if (B == 0 || N == 0) return 0;
    PB_REQUIRE(x && y && idx, "pb_knn1: null pointer argument");
    dim3 grid((N + KT * QPT - 1) / (KT * QPT), B);
    knn1_kernel<<<grid, KT, 0, static_cast<cudaStream_t>(stream)>>>(x, y, N, M, y_shared, idx, d2);
    return pb_check_launch("knn1_kernel") ? -2 : 0;
}

extern "C" int pb_add_adds(const float* pred, const float* target, int B, int N, const uint8_t* want_adds, float* add,
                           float* adds, void* stream) {
    PB_REQUIRE(B >= 0 && N > 0, "pb_add_adds: bad sizes B=%d N=%d", B, N);
    if (B == 0) return 0;
    PB_REQUIRE(pred && target && (add || adds), "pb_add_adds: null pointer argument");
    add_adds_kernel<<<B, KT, 0, static_cast<cudaStream_t>(stream)>>>(pred, target, N, want_adds, add, adds);
    return pb_check_launch("add_adds_kernel") ? -2 : 0;
}

// ADD as in add_adds_kernel; ADD-S from the nearest-neighbour squared distances of the tensor path, summed in
// add_adds_kernel's order (thread-strided, then the deterministic block sum) so both paths return the same bits.
__global__ void __launch_bounds__(KT) add_adds_finish_kernel(const float* __restrict__ pred, const float* __restrict__ target,
                                                             int N, const uint8_t* __restrict__ want_adds,
                                                             const float* __restrict__ d2, float* __restrict__ add,
                                                             float* __restrict__ adds) {
    __shared__ float red[KT / 32];
    const int b = blockIdx.x;
    const float* pb = pred + (size_t)b * N * 3;
    const float* tb = target + (size_t)b * N * 3;
    float s = 0.0f;
    for (int i = threadIdx.x; i < N; i += KT)
        s += sqrtf(sqdist3(pb[3 * (size_t)i], pb[3 * (size_t)i + 1], pb[3 * (size_t)i + 2], tb[3 * (size_t)i],
                           tb[3 * (size_t)i + 1], tb[3 * (size_t)i + 2]));
    const float add_v = block_sum_kt(s, red) / (float)N;
    if (threadIdx.x == 0 && add) add[b] = add_v;
    if (!adds) return;
    const bool want = want_adds ? (want_adds[b] != 0) : true;
    if (!want) {
        if (threadIdx.x == 0) adds[b] = add_v;
        return;
    }
    float s2 = 0.0f;
    for (int base = 0; base < N; base += KT * QPT)
#pragma unroll
        for (int q = 0; q < QPT; ++q) {
            const int i = base + q * KT + threadIdx.x;
            if (i < N) s2 += sqrtf(d2[(size_t)b * N + i]);
        }
    const float adds_v = block_sum_kt(s2, red) / (float)N;
    if (threadIdx.x == 0) adds[b] = adds_v;
}

static size_t d2_bytes(int B, int N) { return (((size_t)B * N * sizeof(float)) + 127) / 128 * 128; }

extern "C" size_t pb_add_adds_tc_workspace_bytes(int B, int N) {
    if (B <= 0 || N <= 0) return 0;
    return d2_bytes(B, N) + pb_knn1_tc_workspace_bytes(B, N, N);
}

extern "C" int pb_add_adds_tc(const float* pred, const float* target, int B, int N, const uint8_t* want_adds,
                              const int* sym_list, int n_sym, float* add, float* adds, void* ws, size_t ws_bytes,
                              void* stream) {
    PB_REQUIRE(B >= 0 && N > 0, "pb_add_adds_tc: bad sizes B=%d N=%d", B, N);
    if (B == 0) return 0;
    PB_REQUIRE(pred && target && adds, "pb_add_adds_tc: null pointer argument");
    PB_REQUIRE(!want_adds || (n_sym >= 0 && n_sym <= B && (sym_list || n_sym == 0)), "pb_add_adds_tc: want_adds needs its list of set elements");
    PB_REQUIRE(ws && ws_bytes >= pb_add_adds_tc_workspace_bytes(B, N), "pb_add_adds_tc: workspace too small");
    PB_REQUIRE(reinterpret_cast<uintptr_t>(ws) % 128 == 0, "pb_add_adds_tc: workspace must be 128-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* d2 = static_cast<float*>(ws);
    char* kws = static_cast<char*>(ws) + d2_bytes(B, N);
    // ADD-S: for every target point the nearest predicted point (lib/utils/metric.py:26-33)
    if (!want_adds || n_sym > 0) {
        const int rc = pb_internal_knn1_tc("pb_add_adds_tc", target, pred, B, N, N, 0, nullptr, d2, kws,
                                           ws_bytes - d2_bytes(B, N), want_adds ? sym_list : nullptr, n_sym, st);
        if (rc) return rc;
    }
    add_adds_finish_kernel<<<B, KT, 0, st>>>(pred, target, N, want_adds, d2, add, adds);
    return pb_check_launch("add_adds_finish_kernel") ? -2 : 0;
}

extern "C" int pb_transform_points(const float* model, int model_shared, const float* R, const float* t, int B, int N,
                                   float* out, void* stream) {
    PB_REQUIRE(B >= 0 && N > 0, "pb_transform_points: bad sizes");
    PB_REQUIRE(B <= 65535, "pb_transform_points: B too large (<= 65535)");
    if (B == 0) return 0;
    PB_REQUIRE(model && R && t && out, "pb_transform_points: null pointer argument");
    dim3 grid((N + 127) / 128, B);
    transform_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(model, model_shared, R, t, N, out);
    return pb_check_launch("transform_kernel") ? -2 : 0;
}
