// score_tc_common.cuh -- pieces shared by the two tensor-core formulations of K3 (ransac_score_tc.cu: tcgen05 + TMEM;
// ransac_score_mma.cu: warp-level mma.sync): FP16 operand splitting, packed FP32x2 arithmetic, the contract's matrix of a
// hypothesis with its centred rows, and the per-hypothesis band inside which the contract's own arithmetic decides.
// The band allows 16 u (u = 2^-24 of the sum of the absolute terms) for the tensor core's internal accumulation; measured
// on B200: <= 5.6 u for exponent spreads up to 2^12 per operand and a dominant term (scratch/hmma_acc_lab.cu,
// profiles/r02_hmma_acc_lab.txt).
#pragma once
#include <cuda_fp16.h>
#include "ransac_plan.h"

namespace k3tc {

typedef unsigned long long f32x2;   // two packed binary32 values
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
    return d;
}
__device__ __forceinline__ f32x2 pack2u(uint32_t lo, uint32_t hi) {      // adjacent registers of a tcgen05.ld: no instruction
    return (f32x2)lo | ((f32x2)hi << 32);
}
__device__ __forceinline__ float min3nan(float a, float b, float c) {   // NaN if any input is NaN
    float d;
    asm("min.NaN.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ uint32_t pkh(__half a, __half b) {
    return (uint32_t)__half_as_ushort(a) | ((uint32_t)__half_as_ushort(b) << 16);
}
__device__ __forceinline__ void split2h(float v, __half& h, __half& m) {
    h = __float2half_rn(v);
    m = __float2half_rn(v - __half2float(h));      // the difference is exact
}

// The contract's matrix of a hypothesis (or the never-inlier matrix when it has no model) and its centred rows.
struct HypM { float m[12]; float c[8]; bool valid; };
__device__ __forceinline__ HypM hyp_matrix(const ScoreArgs& a, int f, int h, float u0, float v0) {
    HypM o;
    o.valid = false;
    if (h < a.H) {
        const float4* p = reinterpret_cast<const float4*>(a.hypRt + ((size_t)f * a.H + h) * 12);
        const float4 r0 = p[0], r1 = p[1], r2 = p[2];
        o.valid = (r2.w == r2.w);
        if (o.valid) {
            const float R[9] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x};
            const float tt[3] = {r2.y, r2.z, r2.w};
            make_proj_s(a.K4[4 * f], a.K4[4 * f + 1], a.K4[4 * f + 2], a.K4[4 * f + 3], R, tt, a.inv_thr, o.m);
        }
    }
    if (!o.valid) {
#pragma unroll
        for (int k = 0; k < 12; ++k) o.m[k] = 0.0f;
        o.m[11] = -1.0f;                                            // w < 0 for every point: never an inlier
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        o.c[k] = fmaf(-u0, o.m[8 + k], o.m[k]);
        o.c[4 + k] = fmaf(-v0, o.m[8 + k], o.m[4 + k]);
    }
    return o;
}

// frame constants (sFr[buf][..], written by the hypothesis-operand producer)
enum { FR_U0 = 0, FR_V0, FR_SC, FR_MAXUS, FR_MAXVS, FR_MAXDU, FR_MAXDV, FR_BIG };

// Largest difference between the tensor-core s and the contract's s over every point of the frame, for one hypothesis:
// c2 Wb^2 + c1 Wb + c0 (see the header).  +inf: always re-evaluate.
__device__ __forceinline__ float hyp_band(const HypM& hm, const float* fr, float kappa) {
    const float u = 5.9604645e-8f, sub = 2.9802322e-8f;             // 2^-24, 2^-25
    const float sc = fr[FR_SC], mus = fr[FR_MAXUS], mvs = fr[FR_MAXVS], mdu = fr[FR_MAXDU], mdv = fr[FR_MAXDV];
    const float A1x = fabsf(hm.m[0]) + fabsf(hm.m[1]) + fabsf(hm.m[2]), A1y = fabsf(hm.m[4]) + fabsf(hm.m[5]) + fabsf(hm.m[6]);
    const float A1w = fabsf(hm.m[8]) + fabsf(hm.m[9]) + fabsf(hm.m[10]);
    const float A1xc = fabsf(hm.c[0]) + fabsf(hm.c[1]) + fabsf(hm.c[2]), A1yc = fabsf(hm.c[4]) + fabsf(hm.c[5]) + fabsf(hm.c[6]);
    float amax = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) amax = fmaxf(amax, fabsf(hm.c[k]));
#pragma unroll
    for (int k = 8; k < 12; ++k) amax = fmaxf(amax, fabsf(hm.m[k]));
    const float ax = u * (6.2f * mus + 51.0f * mdu + 30.1f), ay = u * (6.2f * mvs + 51.0f * mdv + 30.1f);
    const float bx = u * (6.2f * A1x + 50.0f * A1xc + (6.2f * mus + 50.0f * mdu) * A1w) + sub * ((A1xc + 3.0f + sc) + mdu * (A1w + 3.0f + sc));
    const float by = u * (6.2f * A1y + 50.0f * A1yc + (6.2f * mvs + 50.0f * mdv) * A1w) + sub * ((A1yc + 3.0f + sc) + mdv * (A1w + 3.0f + sc));
    const float ew = 28.1f * u;
    const float c2 = 1.02f * (2.0f * (ax + ay) + 2.0f * ew + 12.0f * u);
    const float c1 = 1.02f * (2.0f * (bx + by) + 4.0f * ew * A1w + 2.0f * sub * (A1w + 3.0f + sc));
    const float e0 = 2.0f * ew * A1w + sub * (A1w + 3.0f + sc);
    const float c0 = 4.0f * (bx * bx + by * by + e0 * e0);
    const float Wb = 1.001f * (A1w + fabsf(hm.m[11]) * sc);         // |w'| of every point of the frame (|P'| <= 1)
    float band = kappa * fmaf(fmaf(c2, Wb, c1), Wb, 0.0f) + kappa * kappa * c0;
    if (!(amax < 32768.0f) || fr[FR_BIG] != 0.0f || !(band == band)) band = __int_as_float(0x7f800000);
    return band;
}

}  // namespace k3tc
