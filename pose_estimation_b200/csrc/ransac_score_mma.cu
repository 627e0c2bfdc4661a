// ransac_score_mma.cu -- K3 (score every correspondence against every hypothesis) with the projection on the tensor
// cores through the warp-level path (mma.sync.m16n8k16, FP16 operands, FP32 accumulators in registers), sm_100a.
// Same result as ransac_score_bal_kernel: integer inlier counts cnt_all[f][h] bit-identical to the FP32 contract
// common.cuh::inlier_s (cv2's PnPRansacCallback::computeError + findInliers inside cv2.solvePnPRansac,
// tools/trainer.py:423-427).
//
// Why not tcgen05 here: the contraction is only K = 16 deep and every accumulator needs five more FMA-pipe
// instructions, so the work is in the epilogue.  With tcgen05 (ransac_score_tc.cu) every accumulator set costs a
// ~1000-cycle hand-off between the issuing warp and the epilogue warps and the values come back through TMEM loads;
// it measures 0.17 - 0.19 ms against 0.150 ms for the FP32-pipe kernel (profiles/r02_k3_tc.md).  mma.sync delivers the
// accumulators straight into the registers of the warp that evaluates them: no hand-off, no TMEM, no specialised
// warps, and its rate (8.0 cycles per m16n8k16 per scheduler = 594 TFLOP/s, scratch/hmma_lab.cu) is 0.75 x 8 = 6
// cycles per warp-level test, below the FMA-pipe cost of the rest.
//
//   operands    as in ransac_score_tc.cu: every FP32 value split into two FP16 (the constant column into three), the
//               four cross products of a coordinate along K = 16; points scaled by a power of two per frame, x / y rows
//               centred on one of the frame's pixels.  A = 16 hypotheses x K (fragments in registers for the whole
//               frame segment: 2 row blocks x {x, y, w} per warp = 32 hypotheses, 8 warps = the frame's 256), B = K x 8
//               points (two 32-bit shared-memory loads per thread and group of 8 points).
//   epilogue    a thread holds rows g, g + 8 and columns 2t, 2t + 1 of every 16 x 8 tile: per row the two points go
//               through packed FP32x2 FMAs (dx, dy, s as in the contract), the sign bits are pushed into the row's
//               verdict word; a (row, group) whose smaller |s| is inside the hypothesis' band
//               (score_tc_common.cuh::hyp_band) is re-evaluated with inlier_s itself after the block of 4 groups.
//   measured    0.162 ms with the derived band, 0.148 ms with 1/16 of it, against 0.150 ms for the FP32-pipe kernel
//               (profiles/r02_k3_tc.md): level, not ahead, so it is opt-in.
//   work        CTAs take equal ranges of the batch's concatenated point list exactly as the balanced FP32 kernel does
//               (same prefix, same asynchronous staging of <= 256 points of one frame into two buffers); a conversion
//               pass per sub-chunk writes the FP16 rows and the negated centred pixel coordinates.
#include "score_tc_common.cuh"

namespace {
using namespace k3tc;

constexpr int SUBM = PB_BAL_SUB;     // points per staging buffer

__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(void* dst_smem, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// D = A (16 x 16, row) * B (16 x 8, col), FP16 in, FP32 out, C = 0
__device__ __forceinline__ void hmma16816(uint32_t (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    float d0, d1, d2, d3;
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
        : "=f"(d0), "=f"(d1), "=f"(d2), "=f"(d3)
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(0.0f));
    d[0] = __float_as_uint(d0); d[1] = __float_as_uint(d1); d[2] = __float_as_uint(d2); d[3] = __float_as_uint(d3);
}

// One warp per frame: the constants the scoring kernel needs of a frame -> fstats[f][0..7] (FR_* of score_tc_common.cuh):
// centre pixel, power-of-two scale (largest |coordinate| * sc in [0.5, 1)), maxima of |us|, |vs|, |us - u0|, |vs - v0|.
__global__ void __launch_bounds__(128) frame_stats_kernel(ScoreArgs a) {
    const int f = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (f >= a.F) return;
    const int cap = a.cap, P = max(0, min(a.count[f], cap));
    const float* Xp = a.pts + (size_t)f * 5 * cap;
    const float u0 = P > 0 ? Xp[3 * (size_t)cap + (P >> 1)] * a.inv_thr : 0.0f;
    const float v0 = P > 0 ? Xp[4 * (size_t)cap + (P >> 1)] * a.inv_thr : 0.0f;
    float cmax = 0.0f, mus = 0.0f, mvs = 0.0f, mdu = 0.0f, mdv = 0.0f;
    bool nan = false;
    for (int p = lane; p < P; p += 32) {
        const float X = Xp[p], Y = Xp[cap + p], Z = Xp[2 * (size_t)cap + p];
        const float us = Xp[3 * (size_t)cap + p] * a.inv_thr, vs = Xp[4 * (size_t)cap + p] * a.inv_thr;
        nan = nan || !(X == X) || !(Y == Y) || !(Z == Z) || !(us == us) || !(vs == vs);
        cmax = fmaxf(cmax, fmaxf(fabsf(X), fmaxf(fabsf(Y), fabsf(Z))));
        mus = fmaxf(mus, fabsf(us)); mvs = fmaxf(mvs, fabsf(vs));
        mdu = fmaxf(mdu, fabsf(us - u0)); mdv = fmaxf(mdv, fabsf(vs - v0));
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        cmax = fmaxf(cmax, __shfl_xor_sync(0xffffffffu, cmax, off));
        mus = fmaxf(mus, __shfl_xor_sync(0xffffffffu, mus, off));
        mvs = fmaxf(mvs, __shfl_xor_sync(0xffffffffu, mvs, off));
        mdu = fmaxf(mdu, __shfl_xor_sync(0xffffffffu, mdu, off));
        mdv = fmaxf(mdv, __shfl_xor_sync(0xffffffffu, mdv, off));
    }
    nan = __any_sync(0xffffffffu, nan);
    int eb = (__float_as_int(cmax) >> 23) & 0xff;
    const bool big = nan || eb > 140 || !(mus < 3.0e37f) || !(mvs < 3.0e37f);
    eb = min(max(eb, 113), 140);
    if (lane == 0) {
        float* fr = a.fstats + (size_t)f * 8;
        fr[FR_U0] = u0; fr[FR_V0] = v0; fr[FR_SC] = __int_as_float((253 - eb) << 23); fr[FR_MAXUS] = mus; fr[FR_MAXVS] = mvs;
        fr[FR_MAXDU] = mdu; fr[FR_MAXDV] = mdv; fr[FR_BIG] = big ? 1.0f : 0.0f;
    }
}

// One warp = 32 hypotheses (two row blocks of 16), one CTA = every hypothesis of the frame: 32 * NW threads, NW = ceil(H / 32)
// <= 8.  CALIB: every test also evaluated exactly (tc_stats[2..3]).
constexpr int NRB = 2;
constexpr int MMA_OFF_ROW = 2 * 5 * SUBM * 4, MMA_OFF_ND = MMA_OFF_ROW + 2 * SUBM * 32, MMA_OFF_PROJ = MMA_OFF_ND + 2 * 2 * SUBM * 4;
constexpr int mma_smem_bytes(int nw) { return MMA_OFF_PROJ + 32 * nw * (48 + 96 + 4); }
template <int NW, bool CALIB>
__global__ void __launch_bounds__(32 * NW, (NW >= 8 ? 3 : NW >= 4 ? 6 : 12)) ransac_score_mma_kernel(ScoreArgs a) {
    constexpr int NT = 32 * NW;
    extern __shared__ __align__(16) unsigned char mma_smem[];
    float (*sbuf)[5][SUBM] = reinterpret_cast<float (*)[5][SUBM]>(mma_smem);                       // [2] X | Y | Z | U | V as in the list
    uint32_t (*srow)[SUBM][8] = reinterpret_cast<uint32_t (*)[SUBM][8]>(mma_smem + MMA_OFF_ROW);   // [2] FP16 operand row of a point
    float (*snd)[2][SUBM] = reinterpret_cast<float (*)[2][SUBM]>(mma_smem + MMA_OFF_ND);           // [2] -(us - u0) | -(vs - v0)
    float (*sproj)[12] = reinterpret_cast<float (*)[12]>(mma_smem + MMA_OFF_PROJ);                 // [NT] the contract's matrices
    uint32_t (*sfrag)[3][8] = reinterpret_cast<uint32_t (*)[3][8]>(mma_smem + MMA_OFF_PROJ + NT * 48);   // [NT] FP16 rows x' | y' | w
    float* sband = reinterpret_cast<float*>(mma_smem + MMA_OFF_PROJ + NT * (48 + 96));             // [NT]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int F = a.F, H = a.H;
    const int* __restrict__ prefix = a.prefix;
    const int T = prefix[F];
    int U = (T + (int)gridDim.x - 1) / (int)gridDim.x;
    U = (U + 7) & ~7;
    const long long g0l = (long long)blockIdx.x * U;
    if (g0l >= T) return;
    const int g0 = (int)g0l, g1 = min(T, g0 + U);
    int f = 0;
    {   // last frame with prefix[f] <= g0
        int lo = 0, hi = F;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (prefix[mid] <= g0) lo = mid; else hi = mid;
        }
        f = lo;
    }
    const float pinf = __int_as_float(0x7f800000);

    auto describe = [&](int gg, int& fr, int& sf, int& sp0, int& sn) -> bool {
        if (gg >= g1) return false;
        while (fr + 1 < F && prefix[fr + 1] <= gg) ++fr;
        sf = fr; sp0 = gg - prefix[fr];
        sn = min(min(SUBM, prefix[fr + 1] - gg), g1 - gg);
        return true;
    };
    auto stage = [&](int b, int sf, int sp0, int sn) {
        const int P = max(0, min(a.count[sf], a.cap));
        const float* base = a.pts + (size_t)sf * 5 * a.cap + sp0;
        for (int e = tid; e < (sn >> 2) * 5; e += NT) {
            const int arr = e / (sn >> 2), i = (e - arr * (sn >> 2)) << 2;
            float* dst = &sbuf[b][arr][i];
            const float* src = base + (size_t)arr * a.cap + i;
            if (sp0 + i + 4 <= P) cp_async16(dst, src);
            else {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (sp0 + i + k < P) cp_async4(dst + k, src + k);
                    else dst[k] = arr >= 3 ? pinf : 0.0f;          // u = v = +inf: the contract gives +inf or NaN, never an inlier
                }
            }
        }
        cp_async_commit();
    };

    // per-thread state of the current frame: this thread's four rows are hypotheses hbase + 16 rb + 8 hr + g
    const int hbase = warp * 32;
    uint32_t A[NRB][3][4];
    float bandv[NRB][2];
    int cnt[NRB][2];
#pragma unroll
    for (int rb = 0; rb < NRB; ++rb) { cnt[rb][0] = cnt[rb][1] = 0; bandv[rb][0] = bandv[rb][1] = 0.0f; }
    float u0 = 0.0f, v0 = 0.0f, sc = 1.0f;
    int nslow = 0, nmis = 0;
    float calmax = 0.0f;
    int cur_f = -1;

    auto flush = [&](int ff) {
#pragma unroll
        for (int rb = 0; rb < NRB; ++rb)
#pragma unroll
            for (int hr = 0; hr < 2; ++hr) {
                int c = cnt[rb][hr];
                c += __shfl_xor_sync(0xffffffffu, c, 1);
                c += __shfl_xor_sync(0xffffffffu, c, 2);
                const int h = hbase + 16 * rb + 8 * hr + g;
                if (t == 0 && h < H && c > 0) atomicAdd(&a.cnt_all[(size_t)ff * H + h], c);
                cnt[rb][hr] = 0;
            }
    };
    // thread tid converts hypothesis tid of the frame (matrix, band, FP16 rows) into shared memory; then every thread
    // picks the fragment words of its four rows: K columns (2t, 2t+1) and (2t+8, 2t+9) of
    // (m0h,m0h,m0m,m0m, m1h,m1h,m1m,m1m | m2h,m2h,m2m,m2m, m3h,m3m,m3l,0)
    auto load_frame = [&](int ff) {
        const float* fr = a.fstats + (size_t)ff * 8;
        u0 = fr[FR_U0]; v0 = fr[FR_V0]; sc = fr[FR_SC];
        __syncthreads();                                           // nobody is still reading the previous frame's rows
        {
            const HypM hm = hyp_matrix(a, ff, tid, u0, v0);
            sband[tid] = hyp_band(hm, fr, a.tc_kappa);
            float4* pd = reinterpret_cast<float4*>(sproj[tid]);
            pd[0] = make_float4(hm.m[0], hm.m[1], hm.m[2], hm.m[3]);
            pd[1] = make_float4(hm.m[4], hm.m[5], hm.m[6], hm.m[7]);
            pd[2] = make_float4(hm.m[8], hm.m[9], hm.m[10], hm.m[11]);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float* rw = c == 0 ? hm.c : c == 1 ? hm.c + 4 : hm.m + 8;
                __half h0, m0, h1, m1, h2, m2, h3, m3, l3;
                split2h(rw[0], h0, m0); split2h(rw[1], h1, m1); split2h(rw[2], h2, m2);
                {
                    const float v = rw[3];
                    h3 = __float2half_rn(v);
                    const float r1 = v - __half2float(h3);
                    m3 = __float2half_rn(r1);
                    l3 = __float2half_rn(r1 - __half2float(m3));
                }
                uint4* d = reinterpret_cast<uint4*>(sfrag[tid][c]);
                d[0] = make_uint4(pkh(h0, h0), pkh(m0, m0), pkh(h1, h1), pkh(m1, m1));
                d[1] = make_uint4(pkh(h2, h2), pkh(m2, m2), pkh(h3, m3), pkh(l3, __float2half_rn(0.0f)));
            }
        }
        __syncthreads();
#pragma unroll
        for (int rb = 0; rb < NRB; ++rb)
#pragma unroll
            for (int hr = 0; hr < 2; ++hr) {
                const int hl = hbase + 16 * rb + 8 * hr + g;
                bandv[rb][hr] = sband[hl];
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    A[rb][c][hr] = sfrag[hl][c][t];                // a0,a1 (row g) / a2,a3 (row g + 8): K 2t, 2t+1
                    A[rb][c][2 + hr] = sfrag[hl][c][4 + t];        // a4,a5 / a6,a7: K 2t+8, 2t+9
                }
            }
    };

    int sf, sp0, sn, nf = 0, np0 = 0, nn = 0;
    int gg = g0;
    bool have = describe(gg, f, sf, sp0, sn);
    if (have) stage(0, sf, sp0, sn);
    for (int kk = 0; have; ++kk) {
        const int b = kk & 1;
        gg += sn;
        const bool more = describe(gg, f, nf, np0, nn);
        if (more) stage(b ^ 1, nf, np0, nn);                       // in flight under this sub-chunk's scoring
        if (sf != cur_f) {                                         // (uniform) another frame: flush, load its hypotheses
            if (cur_f >= 0) flush(cur_f);
            cur_f = sf;
            load_frame(sf);
        }
        if (more) cp_async_wait<1>(); else cp_async_wait<0>();     // this sub-chunk's copies have landed (own thread)
        __syncthreads();                                           // ... and everybody else's
        // ---- conversion pass: FP16 operand rows and negated centred pixel coordinates of the sub-chunk's points
        {
            const int P = max(0, min(a.count[sf], a.cap));
            const __half sh = __float2half_rn(sc), z = __float2half_rn(0.0f);
            for (int i = tid; i < sn; i += NT) {
                const bool in = sp0 + i < P;
                __half xh, xm, yh, ym, zh, zm;
                split2h(sbuf[b][0][i] * sc, xh, xm); split2h(sbuf[b][1][i] * sc, yh, ym); split2h(sbuf[b][2][i] * sc, zh, zm);
                uint4* d = reinterpret_cast<uint4*>(srow[b][i]);
                d[0] = make_uint4(pkh(xh, xm), pkh(xh, xm), pkh(yh, ym), pkh(yh, ym));
                d[1] = make_uint4(pkh(zh, zm), pkh(zh, zm), pkh(sh, sh), pkh(sh, z));
                // padding: the origin with a pixel far away -> s overflows to +inf, never an inlier, never in the band
                snd[b][0][i] = in ? -(sbuf[b][3][i] * a.inv_thr - u0) : 1.0e18f;
                snd[b][1][i] = in ? -(sbuf[b][4][i] * a.inv_thr - v0) : 1.0e18f;
            }
        }
        __syncthreads();
        // ---- blocks of 4 groups of 8 points: 8 tests per hypothesis row and thread
        const uint32_t* rowp = &srow[b][g][t];                     // + 64 words per group
        const float* ndup = &snd[b][0][2 * t];                     // + 8 floats per group
        const float* ndvp = &snd[b][1][2 * t];
#pragma unroll 1
        for (int i0 = 0; i0 < sn; i0 += 32, rowp += 256, ndup += 32, ndvp += 32) {
            const int ngrp = min(4, (sn - i0) >> 3);
            uint32_t bits[NRB][2];
#pragma unroll
            for (int rb = 0; rb < NRB; ++rb) bits[rb][0] = bits[rb][1] = 0u;
            uint32_t fm = 0u;                                      // bit 4 * (2 rb + hr) + grp: that row's two tests of the group
#pragma unroll
            for (int grp = 0; grp < 4; ++grp) {
                if (grp < ngrp) {
                    const uint32_t b0 = rowp[64 * grp], b1 = rowp[64 * grp + 4];
                    const f32x2 ndu2 = *reinterpret_cast<const f32x2*>(ndup + 8 * grp);
                    const f32x2 ndv2 = *reinterpret_cast<const f32x2*>(ndvp + 8 * grp);
#pragma unroll
                    for (int rb = 0; rb < NRB; ++rb) {
                        uint32_t cx[4], cy[4], cw[4];
                        hmma16816(cx, A[rb][0], b0, b1);
                        hmma16816(cy, A[rb][1], b0, b1);
                        hmma16816(cw, A[rb][2], b0, b1);
#pragma unroll
                        for (int hr = 0; hr < 2; ++hr) {
                            const float w0 = __uint_as_float(cw[2 * hr]), w1 = __uint_as_float(cw[2 * hr + 1]);
                            const f32x2 x2 = pack2u(cx[2 * hr], cx[2 * hr + 1]), y2 = pack2u(cy[2 * hr], cy[2 * hr + 1]), w2 = pack2u(cw[2 * hr], cw[2 * hr + 1]);
                            const f32x2 dx2 = ffma2(ndu2, w2, x2), dy2 = ffma2(ndv2, w2, y2);
                            const f32x2 nq2 = pack2(w0 * -fabsf(w0), w1 * -fabsf(w1));
                            const f32x2 s2 = ffma2(dx2, dx2, ffma2(dy2, dy2, nq2));
                            const uint32_t s0 = (uint32_t)s2, s1 = (uint32_t)(s2 >> 32);
                            bits[rb][hr] = __funnelshift_l(s0, bits[rb][hr], 1);
                            bits[rb][hr] = __funnelshift_l(s1, bits[rb][hr], 1);
                            const float sm = min3nan(fabsf(__uint_as_float(s0)), fabsf(__uint_as_float(s1)), pinf);
                            if (!(sm > bandv[rb][hr])) fm |= 1u << (4 * (2 * rb + hr) + grp);
                            if (CALIB) {
#pragma unroll
                                for (int cc = 0; cc < 2; ++cc) {
                                    const int p = i0 + 8 * grp + 2 * t + cc;
                                    const float* M = sproj[hbase + 16 * rb + 8 * hr + g];
                                    const float se = inlier_s(M, sbuf[b][0][p], sbuf[b][1][p], sbuf[b][2][p], sbuf[b][3][p] * a.inv_thr,
                                                              sbuf[b][4][p] * a.inv_thr) * (sc * sc);
                                    const float sv = __uint_as_float(cc ? s1 : s0), ww = cc ? w1 : w0;
                                    const float ratio = fabsf(sv - se) / bandv[rb][hr];
                                    const bool real = snd[b][0][p] < 1.0e17f;
                                    if (real && ratio == ratio && bandv[rb][hr] < 1e30f && fabsf(se) <= 0.25f * ww * ww) calmax = fmaxf(calmax, ratio);
                                    if (real && fabsf(sv) > bandv[rb][hr] && ((__float_as_uint(sv) ^ __float_as_uint(se)) >> 31)) ++nmis;
                                }
                            }
                        }
                    }
                }
            }
            // ---- (row, group) pairs with a value inside the row's band: the contract's own arithmetic decides their two tests
            uint32_t any = __reduce_or_sync(0xffffffffu, fm);
#pragma unroll 1
            while (any) {
                const int bit = __ffs(any) - 1;
                any &= any - 1;
                const int r = bit >> 2, grp = bit & 3;             // r = 2 rb + hr
                ++nslow;
                const float* Mp = sproj[hbase + 16 * (r >> 1) + 8 * (r & 1) + g];
                const float4 m0 = reinterpret_cast<const float4*>(Mp)[0], m1 = reinterpret_cast<const float4*>(Mp)[1],
                             m2 = reinterpret_cast<const float4*>(Mp)[2];
                const float M[12] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w, m2.x, m2.y, m2.z, m2.w};
                const int p = i0 + 8 * grp + 2 * t;
                const uint32_t e0 = __float_as_uint(inlier_s(M, sbuf[b][0][p], sbuf[b][1][p], sbuf[b][2][p], sbuf[b][3][p] * a.inv_thr,
                                                             sbuf[b][4][p] * a.inv_thr)) >> 31;
                const uint32_t e1 = __float_as_uint(inlier_s(M, sbuf[b][0][p + 1], sbuf[b][1][p + 1], sbuf[b][2][p + 1],
                                                             sbuf[b][3][p + 1] * a.inv_thr, sbuf[b][4][p + 1] * a.inv_thr)) >> 31;
                // the group's two verdicts sit at bits 2 (ngrp - 1 - grp) + 1 and + 0 of the row's word
                const int sh = 2 * (ngrp - 1 - grp);
                const uint32_t patch = ((e0 << 1) | e1) << sh, keep = ~(3u << sh);
#pragma unroll
                for (int rb = 0; rb < NRB; ++rb)
#pragma unroll
                    for (int hr = 0; hr < 2; ++hr)
                        if (r == 2 * rb + hr) bits[rb][hr] = (bits[rb][hr] & keep) | patch;
            }
#pragma unroll
            for (int rb = 0; rb < NRB; ++rb) { cnt[rb][0] += __popc(bits[rb][0]); cnt[rb][1] += __popc(bits[rb][1]); }
        }
        __syncthreads();                                           // buffer b is free for the sub-chunk after next
        have = more; sf = nf; sp0 = np0; sn = nn;
    }
    if (cur_f >= 0) flush(cur_f);
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

template <int NW>
int launch_mma(const ScoreArgs& a, int calib, cudaStream_t st) {
    const int grid = pb_sm_count() * (NW >= 8 ? 3 : NW >= 4 ? 6 : 12);
    constexpr int smem = mma_smem_bytes(NW);
    if (calib) {
        cudaFuncSetAttribute(ransac_score_mma_kernel<NW, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        ransac_score_mma_kernel<NW, true><<<grid, 32 * NW, smem, st>>>(a);
    } else {
        cudaFuncSetAttribute(ransac_score_mma_kernel<NW, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        ransac_score_mma_kernel<NW, false><<<grid, 32 * NW, smem, st>>>(a);
    }
    return pb_check_launch("ransac_score_mma_kernel");
}

}  // namespace

int pb_launch_score_mma(const ScoreArgs& a, int calib, cudaStream_t st) {
    cudaMemsetAsync(a.tc_stats, 0, PB_TC_STATS_BYTES, st);
    frame_stats_kernel<<<(a.F + 3) / 4, 128, 0, st>>>(a);
    if (pb_check_launch("frame_stats_kernel")) return 1;
    if (a.H <= 64) return launch_mma<2>(a, calib, st);
    if (a.H <= 128) return launch_mma<4>(a, calib, st);
    return launch_mma<8>(a, calib, st);
}
