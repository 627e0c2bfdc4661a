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
//               frame segment: 4 row blocks x {x, y, w} per warp = 64 hypotheses), B = K x 8 points (two 32-bit
//               shared-memory loads per thread and group of 8 points).
//   epilogue    a thread holds rows g, g + 8 and columns 2t, 2t + 1 of every 16 x 8 tile: per row the two points go
//               through packed FP32x2 FMAs (dx, dy, s as in the contract), the sign bits are pushed into the row's
//               verdict word, min |s| is kept per row.  After 4 groups (8 tests per row) a row whose min |s| is inside
//               the hypothesis' band (score_tc_common.cuh::hyp_band) is re-evaluated with inlier_s itself.
//   work        CTAs take equal ranges of the batch's concatenated point list exactly as the balanced FP32 kernel does
//               (same prefix, same asynchronous staging of <= 256 points of one frame into two buffers); a conversion
//               pass per sub-chunk writes the FP16 rows and the negated centred pixel coordinates.
#include "score_tc_common.cuh"

namespace {
using namespace k3tc;

constexpr int MT_ = 128;             // threads per CTA: 4 warps
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
__device__ __forceinline__ void hmma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(0.0f));
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

// NRB row blocks of 16 hypotheses per warp: H <= 64 * NRB.  CALIB: every test also evaluated exactly (tc_stats[2..3]).
template <int NRB, bool CALIB>
__global__ void __launch_bounds__(MT_, 4) ransac_score_mma_kernel(ScoreArgs a) {
    __shared__ __align__(16) float sbuf[2][5][SUBM];             // X | Y | Z | U | V as they lie in the correspondence list
    __shared__ __align__(16) uint32_t srow[2][SUBM][8];          // FP16 operand row of a point: 16 halves
    __shared__ __align__(16) float snd[2][2][SUBM];              // -(us - u0) | -(vs - v0)
    __shared__ __align__(16) float sproj[64 * NRB * 4 / 4][12];  // the contract's matrices of the CTA's hypotheses (one frame)
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
        for (int e = tid; e < (sn >> 2) * 5; e += MT_) {
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

    // per-thread state of the current frame: this thread's rows are hypotheses hbase + 16 rb + g (+ 8)
    const int hbase = warp * 16 * NRB;
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
    auto load_frame = [&](int ff) {
        const float* fr = a.fstats + (size_t)ff * 8;
        u0 = fr[FR_U0]; v0 = fr[FR_V0]; sc = fr[FR_SC];
        __syncthreads();                                           // nobody is still reading the previous frame's matrices
#pragma unroll
        for (int rb = 0; rb < NRB; ++rb)
#pragma unroll
            for (int hr = 0; hr < 2; ++hr) {
                const int hl = 16 * rb + 8 * hr + g;
                const HypM hm = hyp_matrix(a, ff, hbase + hl, u0, v0);
                bandv[rb][hr] = hyp_band(hm, fr, a.tc_kappa);
                if (t == 0) {
                    float4* pd = reinterpret_cast<float4*>(sproj[hbase + hl]);
                    pd[0] = make_float4(hm.m[0], hm.m[1], hm.m[2], hm.m[3]);
                    pd[1] = make_float4(hm.m[4], hm.m[5], hm.m[6], hm.m[7]);
                    pd[2] = make_float4(hm.m[8], hm.m[9], hm.m[10], hm.m[11]);
                }
                // fragment registers of this row: K columns (2t, 2t+1) and (2t+8, 2t+9) of
                // (m0h,m0h,m0m,m0m, m1h,m1h,m1m,m1m | m2h,m2h,m2m,m2m, m3h,m3m,m3l,0)
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
                    const __half z = __float2half_rn(0.0f);
                    const uint32_t lo = t == 0 ? pkh(h0, h0) : t == 1 ? pkh(m0, m0) : t == 2 ? pkh(h1, h1) : pkh(m1, m1);
                    const uint32_t hi = t == 0 ? pkh(h2, h2) : t == 1 ? pkh(m2, m2) : t == 2 ? pkh(h3, m3) : pkh(l3, z);
                    A[rb][c][hr] = lo;                             // a0,a1 (row g) / a2,a3 (row g + 8): K 2t, 2t+1
                    A[rb][c][2 + hr] = hi;                         // a4,a5 / a6,a7: K 2t+8, 2t+9
                }
            }
        __syncthreads();
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
            for (int i = tid; i < sn; i += MT_) {
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
#pragma unroll 1
        for (int i0 = 0; i0 < sn; i0 += 32) {
            const int ngrp = min(4, (sn - i0) >> 3);
            uint32_t bits[NRB][2];
            float smin[NRB][2];
#pragma unroll
            for (int rb = 0; rb < NRB; ++rb) { bits[rb][0] = bits[rb][1] = 0u; smin[rb][0] = smin[rb][1] = pinf; }
#pragma unroll
            for (int grp = 0; grp < 4; ++grp) {
                if (grp < ngrp) {
                    const int ib = i0 + 8 * grp;
                    const uint32_t b0 = srow[b][ib + g][t], b1 = srow[b][ib + g][4 + t];
                    const f32x2 ndu2 = *reinterpret_cast<const f32x2*>(&snd[b][0][ib + 2 * t]);
                    const f32x2 ndv2 = *reinterpret_cast<const f32x2*>(&snd[b][1][ib + 2 * t]);
#pragma unroll
                    for (int rb = 0; rb < NRB; ++rb) {
                        float cx[4], cy[4], cw[4];
                        hmma16816(cx, A[rb][0], b0, b1);
                        hmma16816(cy, A[rb][1], b0, b1);
                        hmma16816(cw, A[rb][2], b0, b1);
#pragma unroll
                        for (int hr = 0; hr < 2; ++hr) {
                            const float w0 = cw[2 * hr], w1 = cw[2 * hr + 1];
                            const f32x2 x2 = pack2(cx[2 * hr], cx[2 * hr + 1]), y2 = pack2(cy[2 * hr], cy[2 * hr + 1]), w2 = pack2(w0, w1);
                            const f32x2 dx2 = ffma2(ndu2, w2, x2), dy2 = ffma2(ndv2, w2, y2);
                            const f32x2 nq2 = pack2(w0 * -fabsf(w0), w1 * -fabsf(w1));
                            const f32x2 s2 = ffma2(dx2, dx2, ffma2(dy2, dy2, nq2));
                            const uint32_t s0 = (uint32_t)s2, s1 = (uint32_t)(s2 >> 32);
                            bits[rb][hr] = __funnelshift_l(s0, bits[rb][hr], 1);
                            bits[rb][hr] = __funnelshift_l(s1, bits[rb][hr], 1);
                            smin[rb][hr] = min3nan(smin[rb][hr], fabsf(__uint_as_float(s0)), fabsf(__uint_as_float(s1)));
                            if (CALIB) {
#pragma unroll
                                for (int cc = 0; cc < 2; ++cc) {
                                    const int p = ib + 2 * t + cc;
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
            // ---- rows with a value inside their band: the contract's own arithmetic decides their 2 * ngrp tests
            uint32_t fm = 0u;
#pragma unroll
            for (int rb = 0; rb < NRB; ++rb)
#pragma unroll
                for (int hr = 0; hr < 2; ++hr) fm |= (!(smin[rb][hr] > bandv[rb][hr])) ? (1u << (2 * rb + hr)) : 0u;
            if (__any_sync(0xffffffffu, fm != 0u)) {
#pragma unroll
                for (int rb = 0; rb < NRB; ++rb)
#pragma unroll
                    for (int hr = 0; hr < 2; ++hr) {
                        if (__any_sync(0xffffffffu, (fm >> (2 * rb + hr)) & 1u)) {
                            ++nslow;
                            const float* Mp = sproj[hbase + 16 * rb + 8 * hr + g];
                            const float4 m0 = reinterpret_cast<const float4*>(Mp)[0], m1 = reinterpret_cast<const float4*>(Mp)[1],
                                         m2 = reinterpret_cast<const float4*>(Mp)[2];
                            const float M[12] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w, m2.x, m2.y, m2.z, m2.w};
                            uint32_t w = 0u;
#pragma unroll 1
                            for (int grp = 0; grp < ngrp; ++grp)
#pragma unroll
                                for (int cc = 0; cc < 2; ++cc) {
                                    const int p = i0 + 8 * grp + 2 * t + cc;
                                    const float s = inlier_s(M, sbuf[b][0][p], sbuf[b][1][p], sbuf[b][2][p], sbuf[b][3][p] * a.inv_thr,
                                                             sbuf[b][4][p] * a.inv_thr);
                                    w = (w << 1) | (__float_as_uint(s) >> 31);
                                }
                            bits[rb][hr] = w;
                        }
                    }
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

template <int NRB>
int launch_mma(const ScoreArgs& a, int calib, int grid, cudaStream_t st) {
    if (calib) ransac_score_mma_kernel<NRB, true><<<grid, MT_, 0, st>>>(a);
    else ransac_score_mma_kernel<NRB, false><<<grid, MT_, 0, st>>>(a);
    return pb_check_launch("ransac_score_mma_kernel");
}

}  // namespace

int pb_launch_score_mma(const ScoreArgs& a, int calib, cudaStream_t st) {
    cudaMemsetAsync(a.tc_stats, 0, PB_TC_STATS_BYTES, st);
    frame_stats_kernel<<<(a.F + 3) / 4, 128, 0, st>>>(a);
    if (pb_check_launch("frame_stats_kernel")) return 1;
    const int grid = pb_sm_count() * 4;
    if (a.H <= 64) return launch_mma<1>(a, calib, grid, st);
    if (a.H <= 128) return launch_mma<2>(a, calib, grid, st);
    return launch_mma<4>(a, calib, grid, st);
}
