// ransac_score_tc.cu -- K3 (score every correspondence against every hypothesis) on the 5th-generation tensor
// cores (tcgen05 + TMEM), sm_100a.  Same result as ransac_score_bal_kernel: integer inlier counts cnt_all[f][h] that
// are bit-identical to the FP32 contract common.cuh::inlier_s (cv2's PnPRansacCallback::computeError +
// findInliers inside cv2.solvePnPRansac, tools/trainer.py:423-427).
//
//   projection   x, y, w of a point under a hypothesis are three dot products of (X, Y, Z, 1) with rows of K[R|t]:
//                a GEMM with K = 4.  Every FP32 operand is split into two FP16 (v ~ h + m, 22 bits; the constant
//                column into three), the four cross products of every coordinate are laid out along K = 16, so ONE
//                tcgen05.mma (kind::f16, M = 128 points, N = 192 = 64 hypotheses x {x, y, w}, FP32 accumulate in
//                TMEM) gives 8192 projections.  Point rows are scaled by a power of two (largest coordinate in
//                [0.5, 1)), hypothesis rows are centred on one of the frame's pixels (x - u0 w, y - v0 w) so that
//                the accumulators hold small numbers; both leave the sign of  s = dx^2 + dy^2 - w|w|  unchanged.
//   epilogue     thread = point (TMEM lane), columns = hypotheses: dx = x' - (u - u0) w, dy likewise, s as in the
//                contract (5 FMA-pipe instructions instead of 14), verdict = sign bit pushed into a 32-bit word by a
//                funnel shift.  Two more FMAs give  |s| - (c2 w^2 + c1 |w|): the band inside which the tensor-core
//                value and the contract's FP32 chain may disagree (bound derived below, per point).  A group of
//                4 hypotheses x 32 points with any value inside the band is re-evaluated with inlier_s itself from
//                the FP32 matrices kept in shared memory -- so the counts equal the contract's, not approximately.
//   counts       32 verdict words of a warp (32 points x 32 hypotheses) are reduced by a 5-step butterfly that
//                doubles the field width at every step: lane L ends with the count of bit L.
//   pipeline     one persistent CTA per SM (all 512 TMEM columns: two accumulator sets of 192), 19 warps:
//                16 epilogue warps (TMEM lane quadrant x 16-hypothesis slice), the MMA issuer, a producer of point
//                tiles and a producer of hypothesis operands (one frame ahead); hand-offs are mbarriers only.
//                CTAs take equal ranges of the batch's 128-point tiles (prefix by the hypothesis kernel).
//
// Error bound (u = 2^-24; S = sum of the absolute values of the terms of a dot product, in the point's scaled units):
//   contract     |x_c - x| <= 3.1 u S_x  (three roundings of the FMA chain), likewise y, w;
//   tensor core  |x_tc - x| <= 25 u S_x' : 8 u for the split residuals (2^-22 per operand), 1 u for the centring FMA,
//                16 u allowed for the accumulation of the 16 exact products (the hardware's internal alignment is
//                not documented; measured on C2 with the calibration pass, see profiles/r02_k3_tc.md);
//   FP16 subnormal parts add at most 2^-25 per product (absolute).
// With S_x <= (|us| + 1)|w| + 2 A1 near the threshold (A1 = largest absolute row sum of the rotation part over the
// frame's hypotheses, |P'| <= 1), the difference of the two s values is at most  c2 w^2 + c1 |w| + c0  with the
// per-point constants computed in band_consts() below.
#include <cuda_fp16.h>
#include "ransac_plan.h"
#include "tc.cuh"

namespace {

constexpr int TP = 128;                       // points per tile = TMEM lanes
constexpr int HG = 32;                        // hypotheses per MMA
constexpr int SETC = 3 * HG;                  // accumulator columns of one set: x | y | w
constexpr int NSET = 4;                       // accumulator sets in flight (384 of the 512 TMEM columns)
constexpr int NG_MAX = 8;                     // H <= 256
constexpr int NSTG = 4;                       // point-tile stages
constexpr int EPI_WARPS = 16;
constexpr int W_MMA = 16, W_PA = 17, W_PB = 18;
constexpr int TC_THREADS = 19 * 32;
constexpr int A_TILE_BYTES = TP * 32;         // [kc 0..1][row][8 halves]
constexpr int B_GROUP_BYTES = SETC * 32;
constexpr int B_FRAME_BYTES = NG_MAX * B_GROUP_BYTES;
constexpr int HP = NG_MAX * HG;               // hypothesis slots per frame buffer
constexpr size_t TC_SMEM = (size_t)NSTG * A_TILE_BYTES + 2 * B_FRAME_BYTES + 2 * HP * 12 * sizeof(float) + 2 * 8 * sizeof(float);

__device__ __forceinline__ uint32_t pkh(__half a, __half b) {
    return (uint32_t)__half_as_ushort(a) | ((uint32_t)__half_as_ushort(b) << 16);
}
__device__ __forceinline__ void split2h(float v, __half& h, __half& m) {
    h = __float2half_rn(v);
    m = __float2half_rn(v - __half2float(h));      // the difference is exact
}
// power-of-two scale of a point: largest |coordinate| * sc in [0.5, 1); clamped to what FP16 holds.
// big = coordinates beyond 2^14 (or not finite): the point is always evaluated exactly.
__device__ __forceinline__ float point_scale(float X, float Y, float Z, bool& big) {
    const float m = fmaxf(fabsf(X), fmaxf(fabsf(Y), fabsf(Z)));
    int eb = (__float_as_int(m) >> 23) & 0xff;
    big = eb > 140 || !(X == X) || !(Y == Y) || !(Z == Z);
    eb = min(max(eb, 113), 140);
    return __int_as_float((253 - eb) << 23);       // 2^(126 - eb)
}

// bounded mbarrier wait that gives up quickly once any thread of the CTA has timed out.  try_wait carries a suspend-time
// hint: the thread sleeps in hardware until the phase completes (or the hint expires), so waiting warps do not spend
// issue slots on polling -- a third of all executed instructions were polling loops without it.
__device__ __forceinline__ bool wait_bar(uint64_t* bar, uint32_t parity, volatile int* bad) {
    const uint32_t a = tc::smem_u32(bar);
#pragma unroll 1
    for (uint32_t spin = 0; spin < (1u << 16); ++spin) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(a), "r"(parity), "r"(200000u) : "memory");
        if (ok) return true;
        if (*bad) return false;
    }
    *bad = 1;
    return false;
}

__device__ __forceinline__ int frame_of(const int* __restrict__ tp, int F, int t) {   // last f with tp[f] <= t
    int lo = 0, hi = F;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (tp[mid] <= t) lo = mid; else hi = mid;
    }
    return lo;
}
__device__ __forceinline__ void advance(const int* __restrict__ tp, int F, int t, int& f) {
    while (f + 1 < F && tp[f + 1] <= t) ++f;
}

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
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ float min3nan(float a, float b, float c) {   // NaN if any input is NaN
    float d;
    asm("min.NaN.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// counts of the 32 bit positions over the warp's 32 words: lane L returns the number of lanes whose bit L is set
__device__ __forceinline__ int bit_column_counts(uint32_t v, int lane) {
    {
        const uint32_t lo = v & 0x55555555u, hi = (v >> 1) & 0x55555555u;
        const bool up = lane & 1;
        v = (up ? hi : lo) + __shfl_xor_sync(0xffffffffu, up ? lo : hi, 1);
    }
    {
        const uint32_t lo = v & 0x33333333u, hi = (v >> 2) & 0x33333333u;
        const bool up = lane & 2;
        v = (up ? hi : lo) + __shfl_xor_sync(0xffffffffu, up ? lo : hi, 2);
    }
    {
        const uint32_t lo = v & 0x0F0F0F0Fu, hi = (v >> 4) & 0x0F0F0F0Fu;
        const bool up = lane & 4;
        v = (up ? hi : lo) + __shfl_xor_sync(0xffffffffu, up ? lo : hi, 4);
    }
    {
        const uint32_t lo = v & 0x00FF00FFu, hi = (v >> 8) & 0x00FF00FFu;
        const bool up = lane & 8;
        v = (up ? hi : lo) + __shfl_xor_sync(0xffffffffu, up ? lo : hi, 8);
    }
    {
        const uint32_t lo = v & 0x0000FFFFu, hi = v >> 16;
        const bool up = lane & 16;
        v = (up ? hi : lo) + __shfl_xor_sync(0xffffffffu, up ? lo : hi, 16);
    }
    return (int)v;
}

// frame constants written by the hypothesis-operand producer: sBand[buf][0..7]
//   0 u0, 1 v0 (centre pixel in threshold units), 2 A1 of the x / y rows as the contract uses them, 3 A1 of the
//   centred rows, 4 A1 of the w row, 5 largest absolute matrix entry (FP16 range guard)
struct BandC { float c2, c1, c0; };
__device__ __forceinline__ BandC band_consts(float us, float vs, float du, float dv, float sc, float A1, float A1c,
                                             float A1w, float Amax, bool big, float kappa) {
    const float u = 5.9604645e-8f;                  // 2^-24
    const float aus = fabsf(us), avs = fabsf(vs), adu = fabsf(du), adv = fabsf(dv);
    const float ax = u * (6.2f * aus + 51.0f * adu + 30.1f);
    const float ay = u * (6.2f * avs + 51.0f * adv + 30.1f);
    const float sub = 2.9802322e-8f;                // 2^-25: FP16 subnormal quantum / 2
    const float bx = u * (6.2f * A1 + 50.0f * A1c + (6.2f * aus + 50.0f * adu) * A1w) + sub * ((A1c + 3.0f + sc) + adu * (A1w + 3.0f + sc));
    const float by = u * (6.2f * A1 + 50.0f * A1c + (6.2f * avs + 50.0f * adv) * A1w) + sub * ((A1c + 3.0f + sc) + adv * (A1w + 3.0f + sc));
    const float ew = 28.1f * u;                     // w: |w_c - w| + |w_tc - w| <= ew (|w| + 2 A1w) + sub (A1w + 3 + sc)
    BandC b;
    b.c2 = kappa * 1.02f * (2.0f * (ax + ay) + 2.0f * ew + 12.0f * u);
    b.c1 = kappa * 1.02f * (2.0f * (bx + by) + 4.0f * ew * A1w + 2.0f * sub * (A1w + 3.0f + sc));
    const float e0 = 2.0f * ew * A1w + sub * (A1w + 3.0f + sc);
    b.c0 = kappa * kappa * 4.0f * (bx * bx + by * by + e0 * e0);
    // outside what the FP16 operands can hold, or a bound that is not a number: everything is re-evaluated
    if (big || !(Amax < 32768.0f) || !(b.c1 == b.c1) || !(b.c2 == b.c2) || !(b.c0 == b.c0)) b.c0 = __int_as_float(0x7f800000);
    return b;
}

// MODE 0: product; 1: calibration (every test also evaluated exactly; fills tc_stats[2..3]); 2: clock64 breakdown of
// the roles' waits (tc_stats + 4 .. as twelve 64-bit sums, see scratch/k3_tc_check.py)
#define K3CLK(v) long long v = 0; if (PROF) v = clock64()
#define K3ACC(k, t0, t1) do { if (PROF) prof[k] += (t1) - (t0); } while (0)
template <int MODE>
__global__ void __launch_bounds__(TC_THREADS, 1) ransac_score_tc_kernel(ScoreArgs a) {
    constexpr bool CALIB = MODE == 1, PROF = MODE == 2;
    long long prof[4] = {0, 0, 0, 0};
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* sA = smem_raw;                                              // [NSTG] point tiles
    unsigned char* sB = sA + NSTG * A_TILE_BYTES;                              // [2 frames][NG_MAX groups] hypothesis operands
    float* sProj = reinterpret_cast<float*>(sB + 2 * B_FRAME_BYTES);           // [2][HP][12] the contract's FP32 matrices
    float* sBand = sProj + 2 * HP * 12;                                        // [2][8]
    __shared__ __align__(8) uint64_t a_full[NSTG], a_free[NSTG], b_full[2], b_free[2], acc_full[NSET], acc_free[NSET];
    __shared__ uint32_t tmem_base;
    __shared__ int s_bad;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int F = a.F, H = a.H, cap = a.cap;
    const int* __restrict__ tp = a.tprefix;
    const int T = tp[F];
    const int U = (T + (int)gridDim.x - 1) / (int)gridDim.x;
    const int t0 = (int)blockIdx.x * U, t1 = min(T, t0 + U);
    if (t0 >= t1) return;                                                      // uniform per CTA
    const int NG = (H + HG - 1) / HG;

    if (tid == 0) {
        for (int i = 0; i < NSTG; ++i) { tc::mbar_init(&a_full[i], 1); tc::mbar_init(&a_free[i], 1); }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&b_full[i], 1); tc::mbar_init(&b_free[i], EPI_WARPS); }
        for (int i = 0; i < NSET; ++i) { tc::mbar_init(&acc_full[i], 1); tc::mbar_init(&acc_free[i], EPI_WARPS); }
        s_bad = 0;
        tc::fence_mbar_init();
    }
    if (warp == W_MMA) tc::tmem_alloc<512>(&tmem_base);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = tmem_base;
    volatile int* bad = &s_bad;
    bool ok = true;

    if (warp == W_PA) {
        // ===== point tiles: FP32 (X, Y, Z) -> scaled FP16 pairs, rows (Xh,Xm,Xh,Xm, Yh,Ym,Yh,Ym | Zh,Zm,Zh,Zm, sc,sc,sc,0)
        int f = frame_of(tp, F, t0);
        int tbeg = tp[f], tend = tp[f + 1], P = max(0, min(a.count[f], cap));
        for (int t = t0, tl = 0; t < t1 && ok; ++t, ++tl) {
            if (t >= tend) { advance(tp, F, t, f); tbeg = tp[f]; tend = tp[f + 1]; P = max(0, min(a.count[f], cap)); }
            const int s = tl % NSTG, use = tl / NSTG;
            const int p0 = (t - tbeg) * TP;
            const float* X = a.pts + (size_t)f * 5 * cap;
            const float* Y = X + cap;
            const float* Z = Y + cap;
            float px[4], py[4], pz[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int p = p0 + 32 * i + lane;
                const bool v = p < P;
                px[i] = v ? X[p] : 0.0f; py[i] = v ? Y[p] : 0.0f; pz[i] = v ? Z[p] : 0.0f;
            }
            K3CLK(c0);
            if (use > 0) ok = __all_sync(0xffffffffu, wait_bar(&a_free[s], (uint32_t)(use - 1) & 1u, bad));
            K3CLK(c1);
            K3ACC(0, c0, c1);
            uint4* dst = reinterpret_cast<uint4*>(sA + s * A_TILE_BYTES);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = 32 * i + lane;
                uint4 k0 = make_uint4(0u, 0u, 0u, 0u), k1 = k0;
                if (p0 + r < P) {
                    bool big;
                    const float sc = point_scale(px[i], py[i], pz[i], big);
                    __half xh, xm, yh, ym, zh, zm;
                    split2h(px[i] * sc, xh, xm); split2h(py[i] * sc, yh, ym); split2h(pz[i] * sc, zh, zm);
                    const __half sh = __float2half_rn(sc), z0 = __float2half_rn(0.0f);
                    k0 = make_uint4(pkh(xh, xm), pkh(xh, xm), pkh(yh, ym), pkh(yh, ym));
                    k1 = make_uint4(pkh(zh, zm), pkh(zh, zm), pkh(sh, sh), pkh(sh, z0));
                }
                dst[r] = k0; dst[TP + r] = k1;
            }
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&a_full[s]);
            K3CLK(c2);
            K3ACC(1, c1, c2);
        }
        if (PROF && lane == 0) {      // [8] waiting for a free stage, [9] building tiles
            atomicAdd(reinterpret_cast<unsigned long long*>(a.tc_stats + 4) + 8, (unsigned long long)prof[0]);
            atomicAdd(reinterpret_cast<unsigned long long*>(a.tc_stats + 4) + 9, (unsigned long long)prof[1]);
        }
    } else if (warp == W_PB) {
        // ===== hypothesis operands of a frame: rows (m0h,m0h,m0m,m0m, m1h,m1h,m1m,m1m | m2h,m2h,m2m,m2m, m3h,m3m,m3l,0),
        // x and y rows centred on the pixel (u0, v0); the contract's FP32 matrices beside them
        int f = frame_of(tp, F, t0);
        int fl = 0;
        for (int t = t0; t < t1 && ok; ++fl) {
            advance(tp, F, t, f);
            const int fb = fl & 1, use = fl >> 1;
            if (use > 0) ok = __all_sync(0xffffffffu, wait_bar(&b_free[fb], (uint32_t)(use - 1) & 1u, bad));
            const int P = max(0, min(a.count[f], cap));
            const float* Up = a.pts + (size_t)f * 5 * cap + 3 * (size_t)cap;
            const float u0 = P > 0 ? Up[P >> 1] * a.inv_thr : 0.0f;
            const float v0 = P > 0 ? Up[cap + (P >> 1)] * a.inv_thr : 0.0f;
            const float fx = a.K4[4 * f], fy = a.K4[4 * f + 1], cx = a.K4[4 * f + 2], cy = a.K4[4 * f + 3];
            float A1 = 0.0f, A1c = 0.0f, A1w = 0.0f, Amax = 0.0f;
            unsigned char* sBf = sB + fb * B_FRAME_BYTES;
            float* sPf = sProj + fb * HP * 12;
            const float qnan = __int_as_float(0x7fffffff);
            float4 n0 = make_float4(0.0f, 0.0f, 0.0f, 0.0f), n1 = n0, n2 = make_float4(0.0f, 0.0f, 0.0f, qnan);
            if (lane < H) {
                const float4* o = reinterpret_cast<const float4*>(a.hypRt + ((size_t)f * H + lane) * 12);
                n0 = o[0]; n1 = o[1]; n2 = o[2];
            }
            for (int h = lane; h < NG * HG; h += 32) {
                float mj[12];
                bool valid = false;
                const float4 r0 = n0, r1 = n1, r2 = n2;
                n2.w = qnan;
                if (h + 32 < H) {                                            // the next pose is in flight during this one's conversion
                    const float4* o = reinterpret_cast<const float4*>(a.hypRt + ((size_t)f * H + h + 32) * 12);
                    n0 = o[0]; n1 = o[1]; n2 = o[2];
                }
                if (h < H) {
                    valid = (r2.w == r2.w);
                    if (valid) {
                        const float R[9] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x};
                        const float tt[3] = {r2.y, r2.z, r2.w};
                        make_proj_s(fx, fy, cx, cy, R, tt, a.inv_thr, mj);
                    }
                }
                if (!valid) {
#pragma unroll
                    for (int k = 0; k < 12; ++k) mj[k] = 0.0f;
                    mj[11] = -1.0f;                                            // w < 0 for every point: never an inlier
                }
                float4* pd = reinterpret_cast<float4*>(sPf + h * 12);
                pd[0] = make_float4(mj[0], mj[1], mj[2], mj[3]);
                pd[1] = make_float4(mj[4], mj[5], mj[6], mj[7]);
                pd[2] = make_float4(mj[8], mj[9], mj[10], mj[11]);
                float row[3][4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    row[0][k] = fmaf(-u0, mj[8 + k], mj[k]);
                    row[1][k] = fmaf(-v0, mj[8 + k], mj[4 + k]);
                    row[2][k] = mj[8 + k];
                }
                A1 = fmaxf(A1, fmaxf(fabsf(mj[0]) + fabsf(mj[1]) + fabsf(mj[2]), fabsf(mj[4]) + fabsf(mj[5]) + fabsf(mj[6])));
                A1c = fmaxf(A1c, fmaxf(fabsf(row[0][0]) + fabsf(row[0][1]) + fabsf(row[0][2]),
                                       fabsf(row[1][0]) + fabsf(row[1][1]) + fabsf(row[1][2])));
                A1w = fmaxf(A1w, fabsf(mj[8]) + fabsf(mj[9]) + fabsf(mj[10]));
                const int g = h / HG, hl = h % HG;
                uint4* grp = reinterpret_cast<uint4*>(sBf + g * B_GROUP_BYTES);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    __half eh[3], em[3], h3, m3, l3;
#pragma unroll
                    for (int k = 0; k < 3; ++k) split2h(row[c][k], eh[k], em[k]);
                    {
                        const float v = row[c][3];
                        h3 = __float2half_rn(v);
                        const float r1 = v - __half2float(h3);
                        m3 = __float2half_rn(r1);
                        l3 = __float2half_rn(r1 - __half2float(m3));
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float av = fabsf(row[c][k]);
                        Amax = (av > Amax || !(av == av)) ? (av == av ? av : __int_as_float(0x7f800000)) : Amax;
                    }
                    const int rr = c * HG + hl;
                    grp[rr] = make_uint4(pkh(eh[0], eh[0]), pkh(em[0], em[0]), pkh(eh[1], eh[1]), pkh(em[1], em[1]));
                    grp[SETC + rr] = make_uint4(pkh(eh[2], eh[2]), pkh(em[2], em[2]), pkh(h3, m3), pkh(l3, __float2half_rn(0.0f)));
                }
            }
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                A1 = fmaxf(A1, __shfl_xor_sync(0xffffffffu, A1, off));
                A1c = fmaxf(A1c, __shfl_xor_sync(0xffffffffu, A1c, off));
                A1w = fmaxf(A1w, __shfl_xor_sync(0xffffffffu, A1w, off));
                Amax = fmaxf(Amax, __shfl_xor_sync(0xffffffffu, Amax, off));
            }
            if (lane == 0) {
                float* bd = sBand + fb * 8;
                bd[0] = u0; bd[1] = v0; bd[2] = A1; bd[3] = A1c; bd[4] = A1w; bd[5] = Amax;
            }
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&b_full[fb]);
            t = min(t1, tp[f + 1]);                                            // first tile of the next frame
        }
    } else if (warp == W_MMA) {
        // ===== MMA issuer: per tile NG MMAs (one per 32-hypothesis group) into the NSET accumulator sets in turn
        const uint32_t idesc = tc::idesc_f16_f32(TP, SETC);
        int f = frame_of(tp, F, t0), cur_f = -1, fl = -1, n = 0;
        int tend = tp[f + 1];
        for (int t = t0, tl = 0; t < t1 && ok; ++t, ++tl) {
            if (t >= tend) { advance(tp, F, t, f); tend = tp[f + 1]; }
            K3CLK(c0);
            if (f != cur_f) {
                cur_f = f; ++fl;
                ok = wait_bar(&b_full[fl & 1], (uint32_t)(fl >> 1) & 1u, bad);
            }
            K3CLK(c1);
            const int s = tl % NSTG;
            ok = __all_sync(0xffffffffu, wait_bar(&a_full[s], (uint32_t)(tl / NSTG) & 1u, bad) && ok);
            K3CLK(c2);
            K3ACC(0, c0, c1); K3ACC(1, c1, c2);
            tc::fence_after_sync();
            const uint64_t ad = tc::smem_desc(tc::smem_u32(sA + s * A_TILE_BYTES), TP * 16, 128);
            for (int g = 0; g < NG && ok; ++g, ++n) {
                const int set = n % NSET, use = n / NSET;
                K3CLK(c3);
                if (use > 0) ok = __all_sync(0xffffffffu, wait_bar(&acc_free[set], (uint32_t)(use - 1) & 1u, bad));
                K3CLK(c4);
                K3ACC(2, c3, c4);
                tc::fence_after_sync();
                const uint64_t bd = tc::smem_desc(tc::smem_u32(sB + (fl & 1) * B_FRAME_BYTES + g * B_GROUP_BYTES), SETC * 16, 128);
                if (lane == 0 && ok) {
                    tc::mma_bf16(tmem + (uint32_t)(set * SETC), ad, bd, idesc, 0);
                    tc::mma_commit(&acc_full[set]);
                }
                __syncwarp();
            }
            if (lane == 0) tc::mma_commit(&a_free[s]);
            __syncwarp();
        }
        if (PROF && lane == 0)        // [4] waiting for hypothesis operands, [5] for point tiles, [6] for a free accumulator set
            for (int k = 0; k < 3; ++k) atomicAdd(reinterpret_cast<unsigned long long*>(a.tc_stats + 4) + 4 + k, (unsigned long long)prof[k]);
    } else {
        // ===== epilogue: warp = (TMEM lane quadrant q, 8-hypothesis slice j of every 32-hypothesis set); thread = point.
        // Software pipeline over the CTA's accumulator sets: the loads of set n + 1 (x | y | w: three tcgen05.ld.x8)
        // are in flight while set n is evaluated; a set is handed back to the MMA warp as soon as it is in registers.
        const int q = warp & 3, j = warp >> 2;
        const uint32_t tlane = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(8 * j);
        const int NS = (t1 - t0) * NG;                       // accumulator sets of this CTA
        int f = frame_of(tp, F, t0), cur_f = -1, fl = -1;
        int tbeg = tp[f], tend = tp[f + 1];                  // tiles of frame f (register copies: no load per tile)
        int cnt[2] = {0, 0};
        int nslow = 0, nmis = 0;
        float calmax = 0.0f;
        float u0 = 0.0f, v0 = 0.0f, A1 = 0.0f, A1c = 0.0f, A1w = 0.0f, Amax = 0.0f;
        int P = 0;
        auto flush = [&](int ff) {
#pragma unroll
            for (int wi = 0; wi < 2; ++wi) {
                const int k = (31 - lane) >> 3, c = (31 - lane) & 7;       // bit position L <- set 4 wi + k of the tile, test c
                const int h = (4 * wi + k) * HG + 8 * j + c;
                if (h < H && cnt[wi] > 0) atomicAdd(&a.cnt_all[(size_t)ff * H + h], cnt[wi]);
                cnt[wi] = 0;
            }
        };
        // this lane's point of the next tile: raw values fetched one tile ahead; rows beyond cap are not read
        float nX = 0.0f, nY = 0.0f, nZ = 0.0f, nU = 0.0f, nV = 0.0f;
        auto fetch_point = [&](int tile_in_frame, int ff) {
            const int p = tile_in_frame * TP + 32 * q + lane;
            const float* base = a.pts + (size_t)ff * 5 * cap;
            if (p < cap) { nX = base[p]; nY = base[cap + p]; nZ = base[2 * (size_t)cap + p]; nU = base[3 * (size_t)cap + p]; nV = base[4 * (size_t)cap + p]; }
        };
        fetch_point(t0 - tbeg, f);
        float X = 0.0f, Y = 0.0f, Z = 0.0f, us = 0.0f, vs = 0.0f, du = 0.0f, dv = 0.0f;
        f32x2 ndu2 = 0ull, ndv2 = 0ull;
        BandC bc = {0.0f, 0.0f, -1.0f};
        bool valid = false;
        const float* sPf = sProj;
        uint32_t bits = 0u;
        uint32_t ra[24], rb[24];
        // evaluate the warp's 8 columns of a set held in r = x[8] | y[8] | w[8]; returns a 2-bit mask of the groups of
        // 4 hypotheses that hold a value inside the band
        auto eval8 = [&](const uint32_t (&r)[24]) -> uint32_t {
            uint32_t fm = 0u;
#pragma unroll
            for (int gq = 0; gq < 2; ++gq) {
                float smin = __int_as_float(0x7f800000), wmax = 0.0f;
#pragma unroll
                for (int pr = 0; pr < 2; ++pr) {
                    const int c = 4 * gq + 2 * pr;
                    const float w0 = __uint_as_float(r[16 + c]), w1 = __uint_as_float(r[16 + c + 1]);
                    const f32x2 x2 = pack2u(r[c], r[c + 1]), y2 = pack2u(r[8 + c], r[8 + c + 1]), w2 = pack2u(r[16 + c], r[16 + c + 1]);
                    const f32x2 dx2 = ffma2(ndu2, w2, x2), dy2 = ffma2(ndv2, w2, y2);
                    const f32x2 nq2 = pack2(w0 * -fabsf(w0), w1 * -fabsf(w1));
                    const f32x2 s2 = ffma2(dx2, dx2, ffma2(dy2, dy2, nq2));
                    const uint32_t s0 = (uint32_t)s2, s1 = (uint32_t)(s2 >> 32);
                    bits = __funnelshift_l(s0, bits, 1);
                    bits = __funnelshift_l(s1, bits, 1);
                    smin = min3nan(smin, fabsf(__uint_as_float(s0)), fabsf(__uint_as_float(s1)));
                    wmax = fmax3(wmax, fabsf(w0), fabsf(w1));
                }
                // band at the group's largest |w| against the group's smallest |s|: covers every one of its tests
                const float band = fmaf(fmaf(bc.c2, wmax, bc.c1), wmax, bc.c0);
                fm |= (!(smin > band)) ? (1u << gq) : 0u;
            }
            return fm;
        };
        // the rest of a set's work once its 8 columns are evaluated: exact re-evaluation inside the band, word completion
        auto finish_set = [&](uint32_t fm, int g, const uint32_t (&r)[24]) {
            if (CALIB) {
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int h = g * HG + 8 * j + c;
                    const float4* mp = reinterpret_cast<const float4*>(sPf + h * 12);
                    const float4 m0 = mp[0], m1 = mp[1], m2 = mp[2];
                    const float M[12] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w, m2.x, m2.y, m2.z, m2.w};
                    bool big2;
                    const float sc = point_scale(X, Y, Z, big2);
                    const float se = inlier_s(M, X, Y, Z, us, vs) * (sc * sc);
                    const float xw = __uint_as_float(r[c]), yw = __uint_as_float(r[8 + c]), ww = __uint_as_float(r[16 + c]);
                    const float dx = fmaf(-du, ww, xw), dy = fmaf(-dv, ww, yw);
                    const float nq = ww * -fabsf(ww);
                    const float sv = fmaf(dx, dx, fmaf(dy, dy, nq));
                    const float band = fmaf(fmaf(bc.c2, fabsf(ww), bc.c1), fabsf(ww), bc.c0);
                    const float ratio = fabsf(sv - se) / band;
                    if (valid && ratio == ratio && band < 1e30f && fabsf(se) <= 0.25f * ww * ww) calmax = fmaxf(calmax, ratio);
                    const bool inband = !(fabsf(sv) > band);
                    if (valid && !inband && ((__float_as_uint(sv) ^ __float_as_uint(se)) >> 31)) ++nmis;
                }
            }
            if (__any_sync(0xffffffffu, fm != 0u)) {
#pragma unroll
                for (int qd = 0; qd < 2; ++qd) {
                    if (__any_sync(0xffffffffu, (fm >> qd) & 1u)) {
                        ++nslow;
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const int c = 4 * qd + k, h = g * HG + 8 * j + c;
                            const float4* mp = reinterpret_cast<const float4*>(sPf + h * 12);
                            const float4 m0 = mp[0], m1 = mp[1], m2 = mp[2];
                            const float M[12] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w, m2.x, m2.y, m2.z, m2.w};
                            const uint32_t bit = __float_as_uint(inlier_s(M, X, Y, Z, us, vs)) >> 31;
                            bits = (bits & ~(1u << (7 - c))) | (bit << (7 - c));
                        }
                    }
                }
            }
            if ((g & 3) == 3 || g == NG - 1) {                             // a verdict word is complete
                bits <<= 8 * (3 - (g & 3));
                const int c = bit_column_counts(valid ? bits : 0u, lane);
                if (g >> 2) cnt[1] += c; else cnt[0] += c;
                bits = 0u;
            }
        };
        if (NS > 0) {
            ok = __all_sync(0xffffffffu, wait_bar(&acc_full[0], 0u, bad));
            tc::fence_after_sync();
            tc::tmem_ld8(tlane, ra); tc::tmem_ld8(tlane + HG, ra + 8); tc::tmem_ld8(tlane + 2 * HG, ra + 16);
        }
        int g = 0, t = t0;
        for (int n = 0; n < NS && ok; ++n) {
            if (g == 0) {
                // ---- a new tile: frame constants, this lane's point, its band
                K3CLK(e0);
                if (t >= tend) {                                           // (uniform) the next frame that has tiles
                    advance(tp, F, t, f);
                    tbeg = tp[f]; tend = tp[f + 1];
                }
                if (f != cur_f) {
                    if (cur_f >= 0) {
                        flush(cur_f);
                        __syncwarp();
                        if (lane == 0) tc::mbar_arrive(&b_free[fl & 1]);
                    }
                    cur_f = f; ++fl;
                    ok = __all_sync(0xffffffffu, wait_bar(&b_full[fl & 1], (uint32_t)(fl >> 1) & 1u, bad));
                    const float* bd = sBand + (fl & 1) * 8;
                    u0 = bd[0]; v0 = bd[1]; A1 = bd[2]; A1c = bd[3]; A1w = bd[4]; Amax = bd[5];
                    P = max(0, min(a.count[f], cap));
                    sPf = sProj + (fl & 1) * HP * 12;
                }
                valid = (t - tbeg) * TP + 32 * q + lane < P;
                X = valid ? nX : 0.0f; Y = valid ? nY : 0.0f; Z = valid ? nZ : 0.0f;
                us = valid ? nU * a.inv_thr : 0.0f; vs = valid ? nV * a.inv_thr : 0.0f;
                if (t + 1 < t1) {
                    if (t + 1 < tend) fetch_point(t + 1 - tbeg, f);
                    else { int f2 = f; advance(tp, F, t + 1, f2); fetch_point(t + 1 - tp[f2], f2); }
                }
                bool big;
                const float sc = point_scale(X, Y, Z, big);
                du = us - u0; dv = vs - v0;
                ndu2 = pack2(-du, -du); ndv2 = pack2(-dv, -dv);
                bc = band_consts(us, vs, du, dv, sc, A1, A1c, A1w, Amax, big, a.tc_kappa);
                if (!valid) { bc.c2 = 0.0f; bc.c1 = 0.0f; bc.c0 = -1.0f; }     // padding lanes (all-zero rows) never ask for a re-evaluation
                K3CLK(e1);
                K3ACC(0, e0, e1);
            }
            K3CLK(e2);
            tc::tmem_ld_wait();                                            // this set is in registers
            tc::fence_before_sync();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&acc_free[n % NSET]);
            K3CLK(e3);
            const bool more = n + 1 < NS;
            if (more) {
                const int sn = (n + 1) % NSET;
                ok = __all_sync(0xffffffffu, wait_bar(&acc_full[sn], (uint32_t)((n + 1) / NSET) & 1u, bad));
                tc::fence_after_sync();
            }
            K3CLK(e4);
            const uint32_t tn = tlane + (uint32_t)(((n + 1) % NSET) * SETC);
            if (n & 1) {
                if (more) { tc::tmem_ld8(tn, ra); tc::tmem_ld8(tn + HG, ra + 8); tc::tmem_ld8(tn + 2 * HG, ra + 16); }
                const uint32_t fm = eval8(rb);
                finish_set(fm, g, rb);
            } else {
                if (more) { tc::tmem_ld8(tn, rb); tc::tmem_ld8(tn + HG, rb + 8); tc::tmem_ld8(tn + 2 * HG, rb + 16); }
                const uint32_t fm = eval8(ra);
                finish_set(fm, g, ra);
            }
            K3CLK(e5);
            K3ACC(2, e2, e3); K3ACC(1, e3, e4); K3ACC(3, e4, e5);
            if (++g == NG) { g = 0; ++t; }
        }
        if (cur_f >= 0) flush(cur_f);
        if (PROF && warp == 0 && lane == 0)   // epilogue warp 0: [0] per-tile prologue, [1] waiting for the MMA, [2] TMEM loads, [3] arithmetic
            for (int k = 0; k < 4; ++k) atomicAdd(reinterpret_cast<unsigned long long*>(a.tc_stats + 4) + k, (unsigned long long)prof[k]);
        nslow = __shfl_sync(0xffffffffu, nslow, 0);
        if (lane == 0 && nslow) atomicAdd(&a.tc_stats[1], nslow);
        if (CALIB) {
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                calmax = fmaxf(calmax, __shfl_xor_sync(0xffffffffu, calmax, off));
                nmis += __shfl_xor_sync(0xffffffffu, nmis, off);
            }
            if (lane == 0) {
                atomicMax(reinterpret_cast<unsigned int*>(&a.tc_stats[2]), __float_as_uint(calmax));
                if (nmis) atomicAdd(&a.tc_stats[3], nmis);
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    if (warp == W_MMA) tc::tmem_dealloc<512>(tmem);
    if (tid == 0 && s_bad) a.tc_stats[0] = 1;
}

}  // namespace

int pb_launch_score_tc(const ScoreArgs& a, int mode, cudaStream_t st) {
    static_assert(TC_SMEM <= 200 * 1024, "shared-memory plan");
    cudaMemsetAsync(a.tc_stats, 0, PB_TC_STATS_BYTES, st);
    const int grid = pb_sm_count();
    if (mode == 1) {
        cudaFuncSetAttribute(ransac_score_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM);
        ransac_score_tc_kernel<1><<<grid, TC_THREADS, TC_SMEM, st>>>(a);
    } else if (mode == 2) {
        cudaFuncSetAttribute(ransac_score_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM);
        ransac_score_tc_kernel<2><<<grid, TC_THREADS, TC_SMEM, st>>>(a);
    } else {
        cudaFuncSetAttribute(ransac_score_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM);
        ransac_score_tc_kernel<0><<<grid, TC_THREADS, TC_SMEM, st>>>(a);
    }
    return pb_check_launch("ransac_score_tc_kernel");
}
