// cloud.cu -- depth back-projection: the camera-frame cloud the T-offset vote consumes (sm_100a).
//
// Replaces dataset/linemod/batchdataset.py:712-721 (depth_choosed / x_map_choosed / y_map_choosed ->
// pt0 = (x - cx) * d / fx, pt1 = (y - cy) * d / fy, pt2 = d -> cloud [NP,3]) and
// lib/transform/coordinate.py:91-101 (uvd2cloud, the same formula on a [bs,n,3] list).  HBM-trivial gather:
// one thread per chosen pixel; float32 with the reference's operation order (sub, mul, IEEE div).
#include "common.cuh"

namespace {

template <typename IDX>
__global__ void __launch_bounds__(128) backproject_kernel(const float* __restrict__ depth, const IDX* __restrict__ choose,
                                                          const float* __restrict__ xmap, const float* __restrict__ ymap,
                                                          const int32_t* __restrict__ roi, const float* __restrict__ K4,
                                                          float depth_scale, int HW, int Ww, int NP, float* __restrict__ cloud) {
    const int f = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= NP) return;
    long long p = (long long)choose[(size_t)f * NP + i];
    p = p < 0 ? 0 : (p >= HW ? HW - 1 : p);
    const float d = depth[(size_t)f * HW + p] * depth_scale;
    float x, y;
    if (xmap) { x = xmap[(size_t)f * NP + i]; y = ymap[(size_t)f * NP + i]; }
    else {
        const int row = (int)(p / Ww), col = (int)(p - (long long)row * Ww);
        x = (float)(col + (roi ? roi[2 * f + 1] : 0));
        y = (float)(row + (roi ? roi[2 * f] : 0));
    }
    const float fx = K4[4 * f], fy = K4[4 * f + 1], cx = K4[4 * f + 2], cy = K4[4 * f + 3];
    float* o = cloud + ((size_t)f * NP + i) * 3;
    o[0] = __fdiv_rn((x - cx) * d, fx);
    o[1] = __fdiv_rn((y - cy) * d, fy);
    o[2] = d;
}

__global__ void __launch_bounds__(256) uvd2cloud_kernel(const float* __restrict__ uvd, const float* __restrict__ K9,
                                                        long long n, float* __restrict__ xyz) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const float fx = K9[0], fy = K9[4], cx = K9[2], cy = K9[5];
    const float u = uvd[3 * i], v = uvd[3 * i + 1], d = uvd[3 * i + 2];
    xyz[3 * i] = __fdiv_rn((u - cx) * d, fx);
    xyz[3 * i + 1] = __fdiv_rn((v - cy) * d, fy);
    xyz[3 * i + 2] = d;
}

}  // namespace

extern "C" int pb_backproject_cloud(const float* depth, const void* choose, int choose_is_i32, const float* xmap,
                                    const float* ymap, const int32_t* roi, const float* K4, float depth_scale, int F,
                                    int Hh, int Ww, int NP, float* cloud, void* stream) {
    PB_REQUIRE(F >= 0 && Hh > 0 && Ww > 0 && NP > 0, "pb_backproject_cloud: bad sizes");
    PB_REQUIRE(F <= 65535, "pb_backproject_cloud: F too large (<= 65535)");
    PB_REQUIRE((long long)Hh * Ww < (1ll << 30), "pb_backproject_cloud: map too large");
    if (F == 0) return 0;
    PB_REQUIRE(depth && choose && K4 && cloud, "pb_backproject_cloud: null pointer argument");
    PB_REQUIRE((xmap == nullptr) == (ymap == nullptr), "pb_backproject_cloud: xmap and ymap go together");
    dim3 grid((NP + 127) / 128, F);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (choose_is_i32)
        backproject_kernel<int32_t><<<grid, 128, 0, st>>>(depth, static_cast<const int32_t*>(choose), xmap, ymap, roi, K4,
                                                          depth_scale, Hh * Ww, Ww, NP, cloud);
    else
        backproject_kernel<int64_t><<<grid, 128, 0, st>>>(depth, static_cast<const int64_t*>(choose), xmap, ymap, roi, K4,
                                                          depth_scale, Hh * Ww, Ww, NP, cloud);
    return pb_check_launch("backproject_kernel") ? -2 : 0;
}

extern "C" int pb_uvd2cloud(const float* uvd, const float* K9, int B, int n, float* xyz, void* stream) {
    PB_REQUIRE(B >= 0 && n >= 0, "pb_uvd2cloud: bad sizes");
    const long long tot = (long long)B * n;
    if (tot == 0) return 0;
    PB_REQUIRE(uvd && K9 && xyz, "pb_uvd2cloud: null pointer argument");
    uvd2cloud_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(uvd, K9, tot, xyz);
    return pb_check_launch("uvd2cloud_kernel") ? -2 : 0;
}
