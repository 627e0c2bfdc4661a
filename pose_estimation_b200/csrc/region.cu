// region.cu -- region-coded coordinate maps: KRRN predicts a per-pixel OFFSET map plus region logits, and the
// model-coordinate map the pose solver consumes is  xyz = xyz_off + base(region)  (sm_100a).
//
//   hard  tools/script/eval.py:93-103 (and tools/trainer.py:391-398, commented out there):
//         region class = argmax_k softmax(region)[k];  base = region_point[class]
//   soft  lib/network/loss_utils.py:18-28 (get_xyz):
//         base = mean_k softmax(region)[k] * region_point[k]            (mean over the n regions, as written)
//
// Two entry points.  pb_region_decode_points re-derives X,Y,Z of ALREADY COMPACTED correspondences (pb_compact_map /
// pb_gather_correspondences output): only foreground pixels touch the n logit planes (n * 4 B per pixel, read as
// 32-byte sectors shared by neighbouring pixels of a row), so the dense [F,n,H,W] logit tensor is never streamed.
// pb_region_decode_map is the dense form (every pixel), HBM-bound: (n + 3) * 4 B read, 12 B written per pixel.
// softmax is monotone, so the hard class is the arg-max of the logits (lowest index on ties).
#include "common.cuh"

namespace {

struct Base3 { float x, y, z; };

// rp: this frame's [n,3] region points; logit(k) = lp[k * stride]
template <bool SOFT>
__device__ __forceinline__ Base3 region_base(const float* __restrict__ lp, size_t stride, const float* __restrict__ rp, int n) {
    Base3 b;
    if (!SOFT) {
        float best = lp[0];
        int idx = 0;
        for (int k = 1; k < n; ++k) {
            const float v = lp[(size_t)k * stride];
            if (v > best) { best = v; idx = k; }
        }
        b.x = rp[3 * idx]; b.y = rp[3 * idx + 1]; b.z = rp[3 * idx + 2];
    } else {
        // one pass over the logits (they are the traffic), eight at a time: independent loads and exponentials inside a
        // chunk, one rescale of the partial sums when a chunk raises the running maximum
        const float ninf = __int_as_float(0xff800000);
        float m = ninf, s = 0.0f, ax = 0.0f, ay = 0.0f, az = 0.0f;
        for (int k0 = 0; k0 < n; k0 += 8) {
            float v[8];
            float cm = ninf;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                v[j] = (k0 + j < n) ? lp[(size_t)(k0 + j) * stride] : ninf;
                cm = fmaxf(cm, v[j]);
            }
            if (cm > m) {
                const float sc = expf(m - cm);                     // 0 for the first chunk (m = -inf)
                s *= sc; ax *= sc; ay *= sc; az *= sc;
                m = cm;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (k0 + j < n) {
                    const float e = expf(v[j] - m);
                    s += e;
                    ax = fmaf(e, rp[3 * (k0 + j)], ax); ay = fmaf(e, rp[3 * (k0 + j) + 1], ay); az = fmaf(e, rp[3 * (k0 + j) + 2], az);
                }
            }
        }
        const float inv = 1.0f / (s * (float)n);
        b.x = ax * inv; b.y = ay * inv; b.z = az * inv;
    }
    return b;
}

template <bool SOFT>
__global__ void __launch_bounds__(128) region_points_kernel(const float* __restrict__ xyz_off, const float* __restrict__ region,
                                                            const float* __restrict__ region_point,
                                                            const int64_t* __restrict__ choose, int NP,
                                                            const int32_t* __restrict__ roi, const float* __restrict__ extent,
                                                            const float* __restrict__ lfborder, int HW, int Ww, int n, int cap,
                                                            float* __restrict__ pts, const int32_t* __restrict__ count) {
    extern __shared__ float s_rp[];                               // [n,3]
    const int f = blockIdx.y;
    for (int k = threadIdx.x; k < 3 * n; k += blockDim.x) s_rp[k] = region_point[(size_t)f * 3 * n + k];
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int cnt = min(count[f], cap);
    if (i >= cnt) return;
    float* X = pts + (size_t)f * 5 * cap;
    int p;
    if (choose) {
        p = (int)choose[(size_t)f * NP + i];
    } else {
        const int rmin = roi ? roi[2 * f] : 0, cmin = roi ? roi[2 * f + 1] : 0;
        const int col = (int)X[3 * (size_t)cap + i] - cmin, row = (int)X[4 * (size_t)cap + i] - rmin;
        p = row * Ww + col;
    }
    if (p < 0 || p >= HW) return;                                 // an index outside the map keeps its offset-only value
    const Base3 b = region_base<SOFT>(region + (size_t)f * n * HW + p, (size_t)HW, s_rp, n);
    const float* po = xyz_off + (size_t)f * 3 * HW + p;
    X[i] = fmaf(po[0] + b.x, extent[3 * f], lfborder[3 * f]);
    X[(size_t)cap + i] = fmaf(po[HW] + b.y, extent[3 * f + 1], lfborder[3 * f + 1]);
    X[2 * (size_t)cap + i] = fmaf(po[2 * (size_t)HW] + b.z, extent[3 * f + 2], lfborder[3 * f + 2]);
}

template <bool SOFT>
__global__ void __launch_bounds__(256) region_map_kernel(const float* __restrict__ xyz_off, const float* __restrict__ region,
                                                         const float* __restrict__ region_point, int HW, int n,
                                                         float* __restrict__ out) {
    extern __shared__ float s_rp[];
    const int f = blockIdx.y;
    for (int k = threadIdx.x; k < 3 * n; k += blockDim.x) s_rp[k] = region_point[(size_t)f * 3 * n + k];
    __syncthreads();
    const float* rg = region + (size_t)f * n * HW;
    const float* po = xyz_off + (size_t)f * 3 * HW;
    float* o = out + (size_t)f * 3 * HW;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
        const Base3 b = region_base<SOFT>(rg + p, (size_t)HW, s_rp, n);
        o[p] = po[p] + b.x;
        o[(size_t)HW + p] = po[(size_t)HW + p] + b.y;
        o[2 * (size_t)HW + p] = po[2 * (size_t)HW + p] + b.z;
    }
}

}  // namespace

extern "C" int pb_region_decode_points(const float* xyz_off, const float* region, const float* region_point,
                                       const int64_t* choose, int NP, const int32_t* roi, const float* extent,
                                       const float* lfborder, int F, int Hh, int Ww, int n_region, int cap, int soft,
                                       float* pts, const int32_t* count, void* stream) {
    PB_REQUIRE(F >= 0 && Hh > 0 && Ww > 0 && cap > 0, "pb_region_decode_points: bad sizes");
    PB_REQUIRE(n_region >= 1 && n_region <= 4096, "pb_region_decode_points: n_region=%d out of range (1..4096)", n_region);
    PB_REQUIRE(F <= 65535, "pb_region_decode_points: F too large (<= 65535)");
    PB_REQUIRE(!choose || NP > 0, "pb_region_decode_points: NP must be positive with a choose list");
    if (F == 0) return 0;
    PB_REQUIRE(xyz_off && region && region_point && extent && lfborder && pts && count,
               "pb_region_decode_points: null pointer argument");
    const int n = (choose && NP < cap) ? NP : cap;
    dim3 grid((n + 127) / 128, F);
    const size_t sm = (size_t)3 * n_region * sizeof(float);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (soft) region_points_kernel<true><<<grid, 128, sm, st>>>(xyz_off, region, region_point, choose, NP, roi, extent, lfborder,
                                                                 Hh * Ww, Ww, n_region, cap, pts, count);
    else region_points_kernel<false><<<grid, 128, sm, st>>>(xyz_off, region, region_point, choose, NP, roi, extent, lfborder,
                                                             Hh * Ww, Ww, n_region, cap, pts, count);
    return pb_check_launch("region_points_kernel") ? -2 : 0;
}

extern "C" int pb_region_decode_map(const float* xyz_off, const float* region, const float* region_point, int F, int Hh,
                                    int Ww, int n_region, int soft, float* xyz, void* stream) {
    PB_REQUIRE(F >= 0 && Hh > 0 && Ww > 0, "pb_region_decode_map: bad sizes");
    PB_REQUIRE(n_region >= 1 && n_region <= 4096, "pb_region_decode_map: n_region=%d out of range (1..4096)", n_region);
    PB_REQUIRE(F <= 65535 && (long long)Hh * Ww < (1ll << 30), "pb_region_decode_map: map or batch too large");
    if (F == 0) return 0;
    PB_REQUIRE(xyz_off && region && region_point && xyz, "pb_region_decode_map: null pointer argument");
    const int HW = Hh * Ww;
    int bx = (HW + 255) / 256;
    const int want = (148 * 8 + F - 1) / F;                        // about one resident wave over the whole batch
    if (bx > want) bx = want;
    dim3 grid(bx, F);
    const size_t sm = (size_t)3 * n_region * sizeof(float);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (soft) region_map_kernel<true><<<grid, 256, sm, st>>>(xyz_off, region, region_point, HW, n_region, xyz);
    else region_map_kernel<false><<<grid, 256, sm, st>>>(xyz_off, region, region_point, HW, n_region, xyz);
    return pb_check_launch("region_map_kernel") ? -2 : 0;
}
