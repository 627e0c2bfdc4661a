// synthgen.cu -- device-side synthetic frame generator for the sharded sweep, sm_100a.
//
// Measurement infrastructure, not a reference call site: config C5 evaluates 65,536 DISTINCT LineMOD-shaped frames,
// 262 GB of dense maps that cannot be kept (or built on the host) -- they are generated on the device, chunk by
// chunk, into a slab that stays zero outside the objects' silhouettes: a chunk's foreground pixels are written
// before it is evaluated and erased (the same kernel, erase = 1) afterwards, so generating a 2048-frame chunk
// costs ~2 M pixel writes instead of 8 GB.
//
// A frame shows the ellipsoid inscribed in its class box under a pose given by the host (pose_estimation_b200/
// synth.py::c5_params: 40 floats per frame); per pixel of the silhouette's bounding box the ray / ellipsoid hit
// gives the exact model coordinate; a pixel is kept with probability P / area (Philox4x32-10 keyed by the seed,
// counter = (pixel, frame)), a kept pixel is an outlier (uniform in the box) with probability `outlier` and
// otherwise gets ~N(0, sigma) noise (sum of four uniforms per coordinate).  Every operation is IEEE float32
// add / mul / div / sqrt in a written order (-fmad=false), so synth.py's NumPy twin reproduces every byte.
#include "common.cuh"

namespace {

constexpr int PARAM_STRIDE = 40;
// params[f]: 0-8 R (object -> camera, row major) | 9-11 t | 12-15 fx fy cx cy | 16-18 o = -R^T t | 19-21 oo = (o - centre) / semi
// | 22-24 semi | 25-27 lfb | 28-30 ext | 31 qc = oo.oo - 1 | 32-35 bbox r0 r1 c0 c1 | 36 P | 37 outlier | 38 sigma * sqrt(3) | 39 -
constexpr uint32_t GEN_TAG = 0x47454e35u;   // "GEN5"

struct GenArgs {
    const float* params;
    int F, Hh, Ww, bbox_max;
    uint32_t seed_lo, seed_hi, frame0;
    float* xyz; uint8_t* mask; int32_t* area; int erase;
};

__device__ __forceinline__ float u01(uint32_t r) { return (float)(r >> 8) * 5.9604644775390625e-08f; }   // 2^-24: exact

// ray / ellipsoid hit of pixel (row, col); Xo = exact model point
__device__ __forceinline__ bool hit_pixel(const float* __restrict__ p, int row, int col, float Xo[3]) {
    const float dcx = ((float)col - p[14]) / p[12];
    const float dcy = ((float)row - p[15]) / p[13];
    float d[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) d[k] = (p[k] * dcx + p[3 + k] * dcy) + p[6 + k];      // R^T (dcx, dcy, 1)
    const float ddx = d[0] / p[22], ddy = d[1] / p[23], ddz = d[2] / p[24];
    const float qa = (ddx * ddx + ddy * ddy) + ddz * ddz;
    const float qb = (ddx * p[19] + ddy * p[20]) + ddz * p[21];
    const float disc = qb * qb - qa * p[31];
    if (!(disc >= 0.0f)) return false;
    const float tau = (-qb - sqrtf(disc)) / qa;
    if (!(tau > 0.0f)) return false;
#pragma unroll
    for (int k = 0; k < 3; ++k) Xo[k] = p[16 + k] + tau * d[k];
    return true;
}

template <bool COUNT>
__global__ void __launch_bounds__(256) synth_kernel(GenArgs a) {
    const int f = blockIdx.y;
    const float* p = a.params + (size_t)f * PARAM_STRIDE;
    const int r0 = (int)p[32], r1 = (int)p[33], c0 = (int)p[34], c1 = (int)p[35];
    const int bw = c1 - c0, bh = r1 - r0;
    const int k = blockIdx.x * 256 + threadIdx.x;
    bool hit = false;
    float Xo[3] = {0.0f, 0.0f, 0.0f};
    int row = 0, col = 0;
    if (bw > 0 && k < bw * bh) {
        row = r0 + k / bw; col = c0 + k % bw;
        hit = hit_pixel(p, row, col, Xo);
    }
    if (COUNT) {
        const int n = __syncthreads_count(hit);
        if (threadIdx.x == 0 && n) atomicAdd(a.area + f, n);
        return;
    }
    if (!hit) return;
    const int pix = row * a.Ww + col;
    const uint32_t frame = a.frame0 + (uint32_t)f;
    uint32_t r[4];
    philox4x32_10((uint32_t)pix, frame, 0u, GEN_TAG, a.seed_lo, a.seed_hi, r);
    const float pk = p[36] / (float)a.area[f];
    if (!(u01(r[0]) < pk)) return;
    float* px = a.xyz + (size_t)f * 3 * a.Hh * a.Ww + pix;
    const size_t plane = (size_t)a.Hh * a.Ww;
    if (a.erase) {
        px[0] = 0.0f; px[plane] = 0.0f; px[2 * plane] = 0.0f;
        a.mask[(size_t)f * plane + pix] = 0;
        return;
    }
    float X[3];
    if (u01(r[1]) < p[37]) {                     // outlier: uniform in the box
        uint32_t q[4];
        philox4x32_10((uint32_t)pix, frame, 1u, GEN_TAG, a.seed_lo, a.seed_hi, q);
#pragma unroll
        for (int c = 0; c < 3; ++c) X[c] = p[25 + c] + u01(q[c]) * p[28 + c];
    } else {                                     // inlier: + sigma * sqrt(3) * (u1 + u2 + u3 + u4 - 2)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            uint32_t q[4];
            philox4x32_10((uint32_t)pix, frame, 2u + (uint32_t)c, GEN_TAG, a.seed_lo, a.seed_hi, q);
            const float s4 = ((u01(q[0]) + u01(q[1])) + (u01(q[2]) + u01(q[3]))) - 2.0f;
            X[c] = Xo[c] + p[38] * s4;
        }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float v = (X[c] - p[25 + c]) / p[28 + c];
        v = fminf(fmaxf(v, 1.0e-3f), 1.0f);
        px[c * plane] = v;
    }
    a.mask[(size_t)f * plane + pix] = 1;
}

}  // namespace

extern "C" int pb_synth_frames(const float* params, int F, int Hh, int Ww, int bbox_max, uint32_t seed_lo,
                               uint32_t seed_hi, uint32_t frame0, float* xyz, uint8_t* mask, int32_t* area, int erase,
                               void* stream) {
    PB_REQUIRE(F >= 0 && Hh > 0 && Ww > 0 && bbox_max > 0, "pb_synth_frames: bad sizes");
    PB_REQUIRE(F <= 65535, "pb_synth_frames: F too large (<= 65535)");
    if (F == 0) return 0;
    PB_REQUIRE(params && xyz && mask && area, "pb_synth_frames: null pointer argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    GenArgs a{params, F, Hh, Ww, bbox_max, seed_lo, seed_hi, frame0, xyz, mask, area, erase};
    dim3 grid((bbox_max + 255) / 256, F);
    if (!erase) {
        if (cudaMemsetAsync(area, 0, (size_t)F * sizeof(int32_t), st) != cudaSuccess) { pb_check_launch("pb_synth_frames memset"); return -2; }
        synth_kernel<true><<<grid, 256, 0, st>>>(a);
        if (pb_check_launch("synth_kernel<count>")) return -2;
    }
    synth_kernel<false><<<grid, 256, 0, st>>>(a);
    return pb_check_launch("synth_kernel") ? -2 : 0;
}
