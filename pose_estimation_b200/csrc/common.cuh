// common.cuh -- shared device helpers (sm_100a only; compiled with -fmad=false so that every
// fused multiply-add in the bit-defined paths is an explicit fmaf()).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "../../include/poseb200.h"

#define PB_SAMPLE_TAG 0x50335031u   /* Philox counter word 3 for minimal-set sampling */
#define PB_RESAMPLE_TAG 0x52534d50u /* ... for inlier resampling                       */
#define PB_SAMPLE_ATTEMPTS 16
#define PB_GN_EPS 1e-6f            /* Gauss-Newton stops once a step is below this (rad and m)        */
#define PB_THREADS 256              /* CTA size of the RANSAC kernels (fixes the refit reduction tree) */

void pb_set_error(const char* fmt, ...);
int pb_check_launch(const char* what);
int pb_sm_count();                                // multiprocessors of the current device (queried once per device)
void pb_stage_mark(int mark, cudaStream_t st);   // no-op unless pb_stage_timing(1) (abi.cu)
// tensor-core nearest neighbour over a list of batch elements (knn_tc.cu); ws sized by pb_knn1_tc_workspace_bytes
int pb_internal_knn1_tc(const char* who, const float* x, const float* y, int B, int N, int M, int y_shared, int64_t* idx,
                        float* d2, void* ws, size_t ws_bytes, const int* blist, int nlist, cudaStream_t st);

#define PB_REQUIRE(cond, ...)                 \
    do {                                      \
        if (!(cond)) {                        \
            pb_set_error(__VA_ARGS__);        \
            return -1;                        \
        }                                     \
    } while (0)

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// 4 distinct indices in [0,P): whole-set redraw on any collision (deterministic, counter based)
__device__ __forceinline__ bool sample4(uint32_t seed_lo, uint32_t seed_hi, uint32_t frame, uint32_t hyp,
                                        int P, int idx[4]) {
    if (P < 4) return false;
    for (uint32_t a = 0; a < PB_SAMPLE_ATTEMPTS; ++a) {
        uint32_t r[4];
        philox4x32_10(hyp, frame, a, PB_SAMPLE_TAG, seed_lo, seed_hi, r);
        int i0 = (int)__umulhi(r[0], (uint32_t)P), i1 = (int)__umulhi(r[1], (uint32_t)P);
        int i2 = (int)__umulhi(r[2], (uint32_t)P), i3 = (int)__umulhi(r[3], (uint32_t)P);
        if (i0 != i1 && i0 != i2 && i0 != i3 && i1 != i2 && i1 != i3 && i2 != i3) {
            idx[0] = i0; idx[1] = i1; idx[2] = i2; idx[3] = i3;
            return true;
        }
    }
    return false;
}

__device__ __forceinline__ float dot3(const float a[3], const float b[3]) {
    return fmaf(a[0], b[0], fmaf(a[1], b[1], a[2] * b[2]));
}
__device__ __forceinline__ void cross3(const float a[3], const float b[3], float c[3]) {
    c[0] = fmaf(a[1], b[2], -(a[2] * b[1]));
    c[1] = fmaf(a[2], b[0], -(a[0] * b[2]));
    c[2] = fmaf(a[0], b[1], -(a[1] * b[0]));
}

// projection matrix M = K [R|t]
__device__ __forceinline__ void make_proj(float fx, float fy, float cx, float cy, const float R[9],
                                          const float t[3], float M[12]) {
    M[0] = fmaf(fx, R[0], cx * R[6]); M[1] = fmaf(fx, R[1], cx * R[7]);
    M[2] = fmaf(fx, R[2], cx * R[8]); M[3] = fmaf(fx, t[0], cx * t[2]);
    M[4] = fmaf(fy, R[3], cy * R[6]); M[5] = fmaf(fy, R[4], cy * R[7]);
    M[6] = fmaf(fy, R[5], cy * R[8]); M[7] = fmaf(fy, t[1], cy * t[2]);
    M[8] = R[6]; M[9] = R[7]; M[10] = R[8]; M[11] = t[2];
}

// Projection matrix in threshold units: rows 0,1 of K[R|t] times inv_thr = 1/thr (thr = 1: unchanged).
__device__ __forceinline__ void make_proj_s(float fx, float fy, float cx, float cy, const float R[9],
                                            const float t[3], float inv_thr, float M[12]) {
    make_proj(fx, fy, cx, cy, R, t, M);
#pragma unroll
    for (int k = 0; k < 8; ++k) M[k] = M[k] * inv_thr;
}

// Division-free reprojection test in threshold units (cv2: |proj - uv|^2 <= thr^2):
//   s = (x - us w)^2 + (y - vs w)^2 - w|w|,  (us,vs) = (u,v)*inv_thr;   inlier <=> s < 0.
// w <= 0 gives -w|w| >= 0, so points behind the camera never pass; NaN never passes (the GPU's
// canonical NaN has a clear sign bit).  14 FMA-pipe instructions; the sign bit of s is the verdict.
// Operation order is a contract with the CPU oracle (its inlier_test): bit-exact inlier sets.
__device__ __forceinline__ float inlier_s(const float M[12], float X, float Y, float Z, float us, float vs) {
    const float x = fmaf(M[0], X, fmaf(M[1], Y, fmaf(M[2], Z, M[3])));
    const float y = fmaf(M[4], X, fmaf(M[5], Y, fmaf(M[6], Z, M[7])));
    const float w = fmaf(M[8], X, fmaf(M[9], Y, fmaf(M[10], Z, M[11])));
    const float dx = fmaf(-us, w, x);
    const float dy = fmaf(-vs, w, y);
    const float q = w * fabsf(w);
    const float t = fmaf(dy, dy, -q);
    return fmaf(dx, dx, t);
}
__device__ __forceinline__ bool inlier_test(const float M[12], float X, float Y, float Z, float us, float vs) {
    return __float_as_int(inlier_s(M, X, Y, Z, us, vs)) < 0;
}

__device__ __forceinline__ void quat_to_rot(const float q[4], float R[9]) {
    float w = q[0], x = q[1], y = q[2], z = q[3];
    float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z;
    float wx = w * x, wy = w * y, wz = w * z;
    R[0] = 1.0f - 2.0f * (yy + zz); R[1] = 2.0f * (xy - wz);        R[2] = 2.0f * (xz + wy);
    R[3] = 2.0f * (xy + wz);        R[4] = 1.0f - 2.0f * (xx + zz); R[5] = 2.0f * (yz - wx);
    R[6] = 2.0f * (xz - wy);        R[7] = 2.0f * (yz + wx);        R[8] = 1.0f - 2.0f * (xx + yy);
}

__device__ __forceinline__ void rot_to_quat(const float R[9], float q[4]) {
    float tr = R[0] + R[4] + R[8];
    float w, x, y, z;
    if (tr > 0.0f) {
        float s = sqrtf(tr + 1.0f) * 2.0f;
        w = 0.25f * s; x = (R[7] - R[5]) / s; y = (R[2] - R[6]) / s; z = (R[3] - R[1]) / s;
    } else if (R[0] > R[4] && R[0] > R[8]) {
        float s = sqrtf(1.0f + R[0] - R[4] - R[8]) * 2.0f;
        w = (R[7] - R[5]) / s; x = 0.25f * s; y = (R[1] + R[3]) / s; z = (R[2] + R[6]) / s;
    } else if (R[4] > R[8]) {
        float s = sqrtf(1.0f + R[4] - R[0] - R[8]) * 2.0f;
        w = (R[2] - R[6]) / s; x = (R[1] + R[3]) / s; y = 0.25f * s; z = (R[5] + R[7]) / s;
    } else {
        float s = sqrtf(1.0f + R[8] - R[0] - R[4]) * 2.0f;
        w = (R[3] - R[1]) / s; x = (R[2] + R[6]) / s; y = (R[5] + R[7]) / s; z = 0.25f * s;
    }
    float n = sqrtf(fmaf(w, w, fmaf(x, x, fmaf(y, y, z * z))));
    float in = 1.0f / n;
    q[0] = w * in; q[1] = x * in; q[2] = y * in; q[3] = z * in;
}

__device__ __forceinline__ float sqdist3(float ax, float ay, float az, float bx, float by, float bz) {
    float dx = ax - bx, dy = ay - by, dz = az - bz;
    return fmaf(dz, dz, fmaf(dy, dy, dx * dx));
}

// xor-butterfly sum: every lane ends with the same value (float add is commutative bit-for-bit)
__device__ __forceinline__ float warp_sum_bfly(float v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v = v + __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}
