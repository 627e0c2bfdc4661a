// fps.cu -- farthest point sampling of model clouds, sm_100a.
//
// Replaces tools/script/sample_model.py:33-46 (farthest_point_sampling) + :49-63 (pairwise_distance), which
// materialise the N x N float32 distance matrix in numpy.  Same arithmetic per pair -- float32
// sqrt((dx*dx + dy*dy) + dz*dz), no fused multiply-add -- same start (index 0) and np.argmax's tie rule
// (first maximum), so the selected indices are identical.  One CTA per cloud: the running distance-to-set
// lives in shared memory (or in the caller's workspace for very large clouds), each of the n_samples steps
// is one pass over the cloud plus a block-wide arg-max on a packed (distance bits, ~index) key.
#include "common.cuh"

namespace {

constexpr int FPS_T = 1024;
constexpr int FPS_SMEM_MAX_PTS = 56 * 1024;      // 224 KB of float distances

__global__ void __launch_bounds__(FPS_T) fps_kernel(const float* __restrict__ pts, const int32_t* __restrict__ count, int cap,
                                                    int n_samples, int32_t* __restrict__ out, float* __restrict__ gdist) {
    extern __shared__ float sdist[];
    __shared__ unsigned long long wkey[FPS_T / 32];
    __shared__ int s_next;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(count ? count[b] : cap, cap);
    const float* p = pts + (size_t)b * cap * 3;
    float* dist = gdist ? gdist + (size_t)b * cap : sdist;
    int32_t* o = out + (size_t)b * n_samples;
    if (n <= 0) {
        for (int i = tid; i < n_samples; i += FPS_T) o[i] = -1;
        return;
    }
    for (int i = tid; i < n; i += FPS_T) dist[i] = INFINITY;      // min(inf, d(., p0)) = dist_mat[:, 0]
    int cur = 0;
    for (int it = 0; it < n_samples; ++it) {
        if (tid == 0) o[it] = cur;
        const float cx = p[3 * (size_t)cur], cy = p[3 * (size_t)cur + 1], cz = p[3 * (size_t)cur + 2];
        unsigned long long key = 0ull;
        for (int i = tid; i < n; i += FPS_T) {
            const float dx = p[3 * (size_t)i] - cx, dy = p[3 * (size_t)i + 1] - cy, dz = p[3 * (size_t)i + 2] - cz;
            const float d = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
            const float m = fminf(dist[i], d);
            dist[i] = m;
            // distances are >= 0: their bit patterns order like the values; ~i makes the lowest index win ties
            const unsigned long long k = ((unsigned long long)__float_as_uint(m) << 32) | (uint32_t)(~(uint32_t)i);
            key = k > key ? k : key;
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, off);
            key = other > key ? other : key;
        }
        if (lane == 0) wkey[warp] = key;
        __syncthreads();
        if (warp == 0) {
            unsigned long long k = wkey[lane];
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, k, off);
                k = other > k ? other : k;
            }
            if (lane == 0) s_next = (int)(~(uint32_t)k);
        }
        __syncthreads();
        cur = s_next;
    }
}

}  // namespace

extern "C" size_t pb_fps_workspace_bytes(int B, int cap) {
    if (B <= 0 || cap <= FPS_SMEM_MAX_PTS) return 0;
    return (size_t)B * cap * sizeof(float);
}

extern "C" int pb_fps(const float* pts, const int32_t* count, int B, int cap, int n_samples, int32_t* idx, void* ws,
                      size_t ws_bytes, void* stream) {
    PB_REQUIRE(B >= 0 && cap > 0 && n_samples >= 0, "pb_fps: bad sizes B=%d cap=%d n_samples=%d", B, cap, n_samples);
    if (B == 0 || n_samples == 0) return 0;
    PB_REQUIRE(pts && idx, "pb_fps: null pointer argument");
    const size_t need = pb_fps_workspace_bytes(B, cap);
    PB_REQUIRE(need == 0 || (ws && ws_bytes >= need), "pb_fps: clouds of more than %d points need a workspace of %zu bytes",
               FPS_SMEM_MAX_PTS, need);
    const size_t smem = need ? 0 : (size_t)cap * sizeof(float);
    cudaFuncSetAttribute(fps_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(FPS_SMEM_MAX_PTS * sizeof(float)));
    fps_kernel<<<B, FPS_T, smem, static_cast<cudaStream_t>(stream)>>>(pts, count, cap, n_samples, idx,
                                                                     need ? static_cast<float*>(ws) : nullptr);
    return pb_check_launch("fps_kernel") ? -2 : 0;
}
