// ransac_score_tc.cu -- K3 (score every correspondence against every hypothesis) on the 5th-generation tensor
// cores (tcgen05 + TMEM), sm_100a.  Same result as ransac_score_bal_kernel: integer inlier counts cnt_all[f][h] that
// are bit-identical to the FP32 contract common.cuh::inlier_s (cv2's PnPRansacCallback::computeError +
// findInliers inside cv2.solvePnPRansac, tools/trainer.py:423-427).
//
//   projection   x, y, w of a point under a hypothesis are three dot products of (X, Y, Z, 1) with rows of K[R|t]:
//                a GEMM with K = 4.  Every FP32 operand is split into two FP16 (v ~ h + m, 22 bits; the constant
//                column into three), the four cross products of every coordinate lie along K = 16, so ONE
//                tcgen05.mma (kind::f16, M = 128 hypotheses, N = 64 points, FP32 accumulate in TMEM) gives 8192
//                values of one component; three of them (x | y | w) fill an accumulator set of 192 columns.
//                Points are scaled by a power of two per frame (largest coordinate in [0.5, 1)), the x / y rows are
//                centred on one of the frame's pixels (x - u0 w, y - v0 w) so that the accumulators hold small
//                numbers; both leave the sign of  s = dx^2 + dy^2 - w|w|  unchanged.
//   epilogue     thread = hypothesis (TMEM lane), columns = points: dx = x' - (u - u0) w, dy likewise, s as in the
//                contract, two points per packed FP32x2 instruction (5 instructions per hypothesis*point instead of
//                14); verdict = sign bit pushed into a 32-bit word by a funnel shift, one popcount per 32 points.
//                The pixel coordinates of the tile's points come from shared memory (one broadcast LDS.128 per
//                two points).  A group of 8 points x 32 hypotheses whose smallest |s| is inside the band in which
//                the tensor-core value and the contract's FP32 chain may disagree (per-hypothesis bound below) is
//                re-evaluated with inlier_s itself -- so the counts equal the contract's, not approximately.
//   pipeline     one persistent CTA per SM (two accumulator sets of 192 TMEM columns), 19 warps: two groups of 8 epilogue
//                warps (TMEM lane quadrant x 32-column half; 32 tests per thread and hand-off, loads of the next 8
//                columns in flight under the evaluation of the current 8), each group bound to one accumulator set so
//                that one group evaluates while the other's MMA is in flight; the MMA issuer, a producer of point tiles
//                and a producer of hypothesis operands (one frame ahead); hand-offs are mbarriers only.  CTAs take
//                equal ranges of the batch's 64-point tiles (prefix by the hypothesis kernel).
//
// Error bound (u = 2^-24; S = sum of the absolute values of the terms of a dot product, in the frame's scaled units):
//   contract     |x_c - x| <= 3.1 u S_x  (three roundings of the FMA chain), likewise y, w;
//   tensor core  |x_tc - x| <= 25 u S_x' : 8 u for the split residuals (2^-22 per operand), 1 u for the centring FMA,
//                16 u allowed for the accumulation of the 16 exact products (the hardware's internal alignment is
//                not documented; measured with the calibration pass, see profiles/r02_k3_tc.md);
//   FP16 subnormal parts add at most 2^-25 per product (absolute).
// With S_x <= (|us| + 1)|w| + 2 A1 near the threshold (A1 = absolute row sum of the hypothesis' rotation part,
// |P'| <= 1), the two s values differ by at most  c2 w^2 + c1 |w| + c0 ; |w| <= Wb = A1w + |t_z| sc for every point of
// the frame, so one number per hypothesis -- hyp_band() -- bounds it.
#include "score_tc_common.cuh"
#include "tc.cuh"

namespace {
using namespace k3tc;

constexpr int TH = 128;                       // hypotheses per M tile = TMEM lanes
constexpr int TPN = PB_TC_TILE;               // points per tile = columns per component
constexpr int SETC = 3 * TPN;                 // accumulator columns of one set: x | y | w
constexpr int NSET = 2;
constexpr int MT_MAX = 2;                     // H <= 256
constexpr int NSTG = 6;                       // point-tile stages
constexpr int EPI_WARPS = 16;                 // two groups of 8: group g reads accumulator set g (M tile g when H > 128)
constexpr int GRP_WARPS = 8;
constexpr int W_MMA = 16, W_PP = 18, W_PH = 19;   // warps 16 and 17 issue the MMAs of accumulator set 0 and 1
constexpr int TC_THREADS = 20 * 32;
constexpr int A_COMP_BYTES = TH * 32;         // [kc 0..1][row][8 halves] of one (M tile, component)
constexpr int A_FRAME_BYTES = MT_MAX * 3 * A_COMP_BYTES;
constexpr int B_TILE_BYTES = TPN * 32;
constexpr int PINFO_BYTES = (TPN / 2) * 16;   // (-du0, -du1, -dv0, -dv1) per pair of points
constexpr int PEXACT_BYTES = 5 * TPN * 4;     // X | Y | Z | us | vs as the contract reads them
constexpr int STAGE_BYTES = B_TILE_BYTES + PINFO_BYTES + PEXACT_BYTES;
constexpr int FR_FLOATS = 16;                 // frame constants: see the hypothesis-operand producer
constexpr int HP = MT_MAX * TH;               // hypothesis slots per frame buffer
constexpr size_t TC_SMEM = (size_t)2 * A_FRAME_BYTES + (size_t)NSTG * STAGE_BYTES + 2 * FR_FLOATS * sizeof(float) + 2 * HP * 12 * sizeof(float);

// bounded mbarrier wait that gives up quickly once any thread of the CTA has timed out.  try_wait carries a suspend-time
// hint: the thread sleeps in hardware until the phase completes (or the hint expires).
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

// MODE 0: product; 1: calibration (every test also evaluated exactly; fills tc_stats[2..3]); 2: clock64 breakdown of
// the roles' waits (tc_stats + 4 .. as twelve 64-bit sums, see scratch/k3_tc_check.py)
#define K3CLK(v) long long v = 0; if (PROF) v = clock64()
#define K3ACC(k, t0, t1) do { if (PROF) prof[k] += (t1) - (t0); } while (0)
template <int MODE>
__global__ void __launch_bounds__(TC_THREADS, 1) ransac_score_tc_kernel(ScoreArgs a) {
    constexpr bool CALIB = MODE == 1, PROF = MODE == 2;
    long long prof[4] = {0, 0, 0, 0};
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* sA = smem_raw;                                              // [2 frames][MT][3 comps] hypothesis operands
    unsigned char* sP = sA + 2 * A_FRAME_BYTES;                                // [NSTG] point tile | pixel pairs | exact values
    float* sFr = reinterpret_cast<float*>(sP + NSTG * STAGE_BYTES);            // [2][FR_FLOATS]
    float* sProj = sFr + 2 * FR_FLOATS;                                        // [2][HP][12] the contract's FP32 matrices
    __shared__ __align__(8) uint64_t p_full[NSTG], p_free[NSTG], h_full[2], h_free[2], acc_full[NSET], acc_free[NSET];
    __shared__ uint32_t tmem_base;
    __shared__ int s_bad;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int F = a.F, H = a.H, cap = a.cap;
    const int* __restrict__ tp = a.tprefix;
    const int T = tp[F];
    const int U = (T + (int)gridDim.x - 1) / (int)gridDim.x;
    const int t0 = (int)blockIdx.x * U, t1 = min(T, t0 + U);
    if (t0 >= t1) return;                                                      // uniform per CTA
    const int MT = (H + TH - 1) / TH;

    if (tid == 0) {
        for (int i = 0; i < NSTG; ++i) { tc::mbar_init(&p_full[i], 1); tc::mbar_init(&p_free[i], 2 + EPI_WARPS); }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&h_full[i], 1); tc::mbar_init(&h_free[i], EPI_WARPS + 1); }
        for (int i = 0; i < NSET; ++i) { tc::mbar_init(&acc_full[i], 1); tc::mbar_init(&acc_free[i], GRP_WARPS); }
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

    if (warp == W_PH) {
        // ===== hypothesis operands of a frame: rows (m0h,m0h,m0m,m0m, m1h,m1h,m1m,m1m | m2h,m2h,m2m,m2m, m3h,m3m,m3l,0) of the
        // centred x / y rows and the w row; before that the frame's constants (scale, centre pixel, coordinate maxima)
        int f = frame_of(tp, F, t0);
        int fl = 0;
        for (int t = t0; t < t1 && ok; ++fl) {
            advance(tp, F, t, f);
            const int fb = fl & 1, use = fl >> 1;
            if (use > 0) ok = __all_sync(0xffffffffu, wait_bar(&h_free[fb], (uint32_t)(use - 1) & 1u, bad));
            const int P = max(0, min(a.count[f], cap));
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
            // power-of-two scale: largest |coordinate| * sc in [0.5, 1), clamped to what FP16 holds
            int eb = (__float_as_int(cmax) >> 23) & 0xff;
            const bool big = nan || eb > 140 || !(mus < 3.0e37f) || !(mvs < 3.0e37f);
            eb = min(max(eb, 113), 140);
            const float sc = __int_as_float((253 - eb) << 23);       // 2^(126 - eb)
            if (lane == 0) {
                float* fr = sFr + fb * FR_FLOATS;
                fr[FR_U0] = u0; fr[FR_V0] = v0; fr[FR_SC] = sc; fr[FR_MAXUS] = mus; fr[FR_MAXVS] = mvs;
                fr[FR_MAXDU] = mdu; fr[FR_MAXDV] = mdv; fr[FR_BIG] = big ? 1.0f : 0.0f;
            }
            unsigned char* sAf = sA + fb * A_FRAME_BYTES;
            for (int h = lane; h < MT * TH; h += 32) {
                const HypM hm = hyp_matrix(a, f, h, u0, v0);
                const int mt = h / TH, row = h % TH;
                float4* pd = reinterpret_cast<float4*>(sProj + ((size_t)fb * HP + h) * 12);
                pd[0] = make_float4(hm.m[0], hm.m[1], hm.m[2], hm.m[3]);
                pd[1] = make_float4(hm.m[4], hm.m[5], hm.m[6], hm.m[7]);
                pd[2] = make_float4(hm.m[8], hm.m[9], hm.m[10], hm.m[11]);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float* rw = c == 0 ? hm.c : c == 1 ? hm.c + 4 : hm.m + 8;
                    __half eh[3], em[3], h3, m3, l3;
#pragma unroll
                    for (int k = 0; k < 3; ++k) split2h(rw[k], eh[k], em[k]);
                    {
                        const float v = rw[3];
                        h3 = __float2half_rn(v);
                        const float r1 = v - __half2float(h3);
                        m3 = __float2half_rn(r1);
                        l3 = __float2half_rn(r1 - __half2float(m3));
                    }
                    uint4* dst = reinterpret_cast<uint4*>(sAf + (mt * 3 + c) * A_COMP_BYTES);
                    dst[row] = make_uint4(pkh(eh[0], eh[0]), pkh(em[0], em[0]), pkh(eh[1], eh[1]), pkh(em[1], em[1]));
                    dst[TH + row] = make_uint4(pkh(eh[2], eh[2]), pkh(em[2], em[2]), pkh(h3, m3), pkh(l3, __float2half_rn(0.0f)));
                }
            }
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&h_full[fb]);
            t = min(t1, tp[f + 1]);                                            // first tile of the next frame
        }
    } else if (warp == W_PP) {
        // ===== point tiles: FP32 (X, Y, Z) -> scaled FP16 pairs, rows (Xh,Xm,Xh,Xm, Yh,Ym,Yh,Ym | Zh,Zm,Zh,Zm, sc,sc,sc,0);
        // beside them the negated centred pixel coordinates (pairs of points) and the values the contract reads
        int f = frame_of(tp, F, t0), cur_f = -1, fl = -1;
        int tbeg = tp[f], tend = tp[f + 1], P = max(0, min(a.count[f], cap));
        float u0 = 0.0f, v0 = 0.0f, sc = 1.0f;
        for (int t = t0, tl = 0; t < t1 && ok; ++t, ++tl) {
            if (t >= tend) { advance(tp, F, t, f); tbeg = tp[f]; tend = tp[f + 1]; P = max(0, min(a.count[f], cap)); }
            const int s = tl % NSTG, use = tl / NSTG;
            const int p0 = (t - tbeg) * TPN;
            const float* Xp = a.pts + (size_t)f * 5 * cap;
            float v[TPN / 32][5];
#pragma unroll
            for (int i = 0; i < TPN / 32; ++i) {
                const int p = p0 + 32 * i + lane;
                const bool in = p < P;
#pragma unroll
                for (int k = 0; k < 5; ++k) v[i][k] = in ? Xp[(size_t)k * cap + p] : 0.0f;
            }
            if (f != cur_f) {                                                  // the frame's constants (same order as the other roles)
                __syncwarp();
                if (lane == 0 && fl >= 0) tc::mbar_arrive(&h_free[fl & 1]);    // done with the previous frame's constants
                cur_f = f; ++fl;
                ok = __all_sync(0xffffffffu, wait_bar(&h_full[fl & 1], (uint32_t)(fl >> 1) & 1u, bad));
                const float* fr = sFr + (fl & 1) * FR_FLOATS;
                u0 = fr[FR_U0]; v0 = fr[FR_V0]; sc = fr[FR_SC];
            }
            K3CLK(c0);
            if (use > 0) ok = __all_sync(0xffffffffu, wait_bar(&p_free[s], (uint32_t)(use - 1) & 1u, bad) && ok);
            K3CLK(c1);
            K3ACC(0, c0, c1);
            unsigned char* st = sP + s * STAGE_BYTES;
            uint4* dst = reinterpret_cast<uint4*>(st);
            float* pinfo = reinterpret_cast<float*>(st + B_TILE_BYTES);
            float* pex = reinterpret_cast<float*>(st + B_TILE_BYTES + PINFO_BYTES);
            const __half sh = __float2half_rn(sc), z0 = __float2half_rn(0.0f);
#pragma unroll
            for (int i = 0; i < TPN / 32; ++i) {
                const int r = 32 * i + lane;
                const bool in = p0 + r < P;
                const float X = v[i][0], Y = v[i][1], Z = v[i][2];
                const float us = v[i][3] * a.inv_thr, vs = v[i][4] * a.inv_thr;
                __half xh, xm, yh, ym, zh, zm;
                split2h(X * sc, xh, xm); split2h(Y * sc, yh, ym); split2h(Z * sc, zh, zm);
                dst[r] = make_uint4(pkh(xh, xm), pkh(xh, xm), pkh(yh, ym), pkh(yh, ym));
                dst[TPN + r] = make_uint4(pkh(zh, zm), pkh(zh, zm), pkh(sh, sh), pkh(sh, z0));
                // padding columns: the origin with a pixel far away -> s overflows to +inf, never an inlier, never in the band
                pinfo[(r >> 1) * 4 + (r & 1)] = in ? -(us - u0) : 1.0e18f;
                pinfo[(r >> 1) * 4 + 2 + (r & 1)] = in ? -(vs - v0) : 1.0e18f;
                // (the contract's formula gives +inf or NaN for a pixel at infinity: a padding column is never counted)
                const float pinf = __int_as_float(0x7f800000);
                pex[r] = X; pex[TPN + r] = Y; pex[2 * TPN + r] = Z; pex[3 * TPN + r] = in ? us : pinf; pex[4 * TPN + r] = in ? vs : pinf;
            }
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&p_full[s]);
            K3CLK(c2);
            K3ACC(1, c1, c2);
        }
        __syncwarp();
        if (lane == 0 && fl >= 0) tc::mbar_arrive(&h_free[fl & 1]);
        if (PROF && lane == 0) {      // [8] waiting for a free stage, [9] building tiles
            atomicAdd(reinterpret_cast<unsigned long long*>(a.tc_stats + 4) + 8, (unsigned long long)prof[0]);
            atomicAdd(reinterpret_cast<unsigned long long*>(a.tc_stats + 4) + 9, (unsigned long long)prof[1]);
        }
    } else if (warp == W_MMA || warp == W_MMA + 1) {
        // ===== MMA issuers: warp W_MMA + g issues the three MMAs (x | y | w) of accumulator set g -- M tile g of every point
        // tile when H > 128, every second point tile otherwise.  Two issuers because a tcgen05.mma occupies its issuing
        // thread for ~100 cycles whatever its size.
        const int g = warp - W_MMA;
        const int mt = MT == 2 ? g : 0;
        const uint32_t idesc = tc::idesc_f16_f32(TH, TPN);
        int f = frame_of(tp, F, t0), cur_f = -1, fl = -1, k = 0;
        int tend = tp[f + 1];
        for (int t = t0, tl = 0; t < t1 && ok; ++t, ++tl) {
            if (t >= tend) { advance(tp, F, t, f); tend = tp[f + 1]; }
            K3CLK(c0);
            if (f != cur_f) {
                cur_f = f; ++fl;
                ok = __all_sync(0xffffffffu, wait_bar(&h_full[fl & 1], (uint32_t)(fl >> 1) & 1u, bad));
            }
            K3CLK(c1);
            const int s = tl % NSTG;
            const bool mine = MT == 2 || (tl & 1) == g;
            if (mine) {
                ok = __all_sync(0xffffffffu, wait_bar(&p_full[s], (uint32_t)(tl / NSTG) & 1u, bad) && ok);
                K3CLK(c2);
                K3ACC(0, c0, c1); K3ACC(1, c1, c2);
                if (k > 0) ok = __all_sync(0xffffffffu, wait_bar(&acc_free[g], (uint32_t)(k - 1) & 1u, bad) && ok);
                ++k;
                K3CLK(c4);
                K3ACC(2, c2, c4);
                tc::fence_after_sync();
                if (lane == 0 && ok) {
                    const uint64_t bd = tc::smem_desc(tc::smem_u32(sP + s * STAGE_BYTES), TPN * 16, 128);
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const uint64_t ad = tc::smem_desc(tc::smem_u32(sA + (fl & 1) * A_FRAME_BYTES + (mt * 3 + c) * A_COMP_BYTES), TH * 16, 128);
                        tc::mma_bf16(tmem + (uint32_t)(g * SETC + c * TPN), ad, bd, idesc, 0);
                    }
                    tc::mma_commit(&acc_full[g]);
                    tc::mma_commit(&p_free[s]);
                }
            } else if (lane == 0) {
                tc::mbar_arrive(&p_free[s]);
            }
            __syncwarp();
        }
        if (PROF && lane == 0 && g == 0)  // [4] waiting for hypothesis operands, [5] for point tiles, [6] for a free accumulator set
            for (int i = 0; i < 3; ++i) atomicAdd(reinterpret_cast<unsigned long long*>(a.tc_stats + 4) + 4 + i, (unsigned long long)prof[i]);
    } else {
        // ===== epilogue: warp = (group g, TMEM lane quadrant q, column half j); thread = hypothesis 128 mt + 32 q + lane.
        // Group g owns accumulator set g: with two M tiles (H > 128) it is M tile g of every point tile, with one it is every
        // second point tile -- so while a group waits for its next MMA the other one keeps the multiprocessor busy.
        // Software pipeline over the warp's 32 columns of a set: four chunks of 8 (x | y | w: three tcgen05.ld.x8), the next
        // chunk's loads in flight under the evaluation of the current one.
        constexpr int CPW = TPN / 2;                         // columns per warp and component
        constexpr int NCH = CPW / 8;                         // chunks per set
        const int g = warp >> 3, q = warp & 3, j = (warp >> 2) & 1;
        const int mt = MT == 2 ? g : 0;
        const int hyp = mt * TH + 32 * q + lane;
        const uint32_t tb = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(g * SETC + CPW * j);
        int f = frame_of(tp, F, t0), cur_f = -1, fl = -1;
        int tend = tp[f + 1];
        int cnt = 0, nslow = 0, nmis = 0, k = 0;             // k: uses of this group's accumulator set so far
        float calmax = 0.0f, bnd = 0.0f, scf = 1.0f;
        const float* sPf = sProj;
        uint32_t ra[24], rb[24];
        auto flush = [&](int ff) {
            if (hyp < H && cnt > 0) atomicAdd(&a.cnt_all[(size_t)ff * H + hyp], cnt);
            cnt = 0;
        };
        for (int t = t0, tl = 0; t < t1 && ok; ++t, ++tl) {
            K3CLK(e0);
            if (t >= tend) { advance(tp, F, t, f); tend = tp[f + 1]; }
            if (f != cur_f) {
                if (cur_f >= 0) {
                    flush(cur_f);
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(&h_free[fl & 1]);
                }
                cur_f = f; ++fl;
                ok = __all_sync(0xffffffffu, wait_bar(&h_full[fl & 1], (uint32_t)(fl >> 1) & 1u, bad));
                const float* fr = sFr + (fl & 1) * FR_FLOATS;
                scf = fr[FR_SC];
                sPf = sProj + ((size_t)(fl & 1) * HP + hyp) * 12;
                const HypM hm = hyp_matrix(a, f, hyp, fr[FR_U0], fr[FR_V0]);
                bnd = hyp_band(hm, fr, a.tc_kappa);
            }
            const int s = tl % NSTG;
            const unsigned char* st = sP + s * STAGE_BYTES;
            const ulonglong2* pinfo = reinterpret_cast<const ulonglong2*>(st + B_TILE_BYTES) + (CPW / 2) * j;
            const float* pex = reinterpret_cast<const float*>(st + B_TILE_BYTES + PINFO_BYTES) + CPW * j;
            K3CLK(e1);
            K3ACC(0, e0, e1);
            if ((MT == 2 || (tl & 1) == g) && ok) {
                K3CLK(e2);
                ok = __all_sync(0xffffffffu, wait_bar(&acc_full[g], (uint32_t)k & 1u, bad));
                ++k;
                K3CLK(e3);
                K3ACC(1, e2, e3);
                tc::fence_after_sync();
                uint32_t bits = 0u;
                // evaluate 8 columns held in r = x[8] | y[8] | w[8]; returns true when a value is inside the band
                auto eval8 = [&](const uint32_t (&r)[24], int ch) -> bool {
                    float smin = __int_as_float(0x7f800000);
#pragma unroll
                    for (int pr = 0; pr < 4; ++pr) {
                        const int c = 2 * pr;
                        const ulonglong2 pi = pinfo[4 * ch + pr];               // (-du0, -du1), (-dv0, -dv1)
                        const float w0 = __uint_as_float(r[16 + c]), w1 = __uint_as_float(r[16 + c + 1]);
                        const f32x2 x2 = pack2u(r[c], r[c + 1]), y2 = pack2u(r[8 + c], r[8 + c + 1]), w2 = pack2u(r[16 + c], r[16 + c + 1]);
                        const f32x2 dx2 = ffma2(pi.x, w2, x2), dy2 = ffma2(pi.y, w2, y2);
                        const f32x2 nq2 = pack2(w0 * -fabsf(w0), w1 * -fabsf(w1));
                        const f32x2 s2 = ffma2(dx2, dx2, ffma2(dy2, dy2, nq2));
                        const uint32_t s0 = (uint32_t)s2, s1 = (uint32_t)(s2 >> 32);
                        bits = __funnelshift_l(s0, bits, 1);
                        bits = __funnelshift_l(s1, bits, 1);
                        smin = min3nan(smin, fabsf(__uint_as_float(s0)), fabsf(__uint_as_float(s1)));
                    }
                    return !(smin > bnd);
                };
                // chunk ch (its verdicts are the low 8 bits of `bits`) holds a value inside the band for some hypothesis of the
                // warp: find the columns, let the contract's own arithmetic decide them
                auto exact8 = [&](const uint32_t (&r)[24], int ch) {
                    ++nslow;
                    uint32_t inb = 0u;
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const float* pf = reinterpret_cast<const float*>(pinfo + 4 * ch + (c >> 1));
                        const float ww = __uint_as_float(r[16 + c]);
                        const float dx = fmaf(pf[c & 1], ww, __uint_as_float(r[c])), dy = fmaf(pf[2 + (c & 1)], ww, __uint_as_float(r[8 + c]));
                        const float sv = fmaf(dx, dx, fmaf(dy, dy, ww * -fabsf(ww)));
                        inb |= (!(fabsf(sv) > bnd)) ? (1u << c) : 0u;
                    }
                    inb = __reduce_or_sync(0xffffffffu, inb);
                    const float4 m0 = reinterpret_cast<const float4*>(sPf)[0], m1 = reinterpret_cast<const float4*>(sPf)[1],
                                 m2 = reinterpret_cast<const float4*>(sPf)[2];
                    const float M[12] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w, m2.x, m2.y, m2.z, m2.w};
#pragma unroll 1
                    while (inb) {
                        const int c = __ffs(inb) - 1;
                        inb &= inb - 1;
                        const int col = 8 * ch + c;
                        const uint32_t bit = __float_as_uint(inlier_s(M, pex[col], pex[TPN + col], pex[2 * TPN + col],
                                                                      pex[3 * TPN + col], pex[4 * TPN + col])) >> 31;
                        bits = (bits & ~(1u << (7 - c))) | (bit << (7 - c));
                    }
                };
                auto calib8 = [&](const uint32_t (&r)[24], int ch) {
                    const float4 m0 = reinterpret_cast<const float4*>(sPf)[0], m1 = reinterpret_cast<const float4*>(sPf)[1],
                                 m2 = reinterpret_cast<const float4*>(sPf)[2];
                    const float M[12] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w, m2.x, m2.y, m2.z, m2.w};
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const int col = 8 * ch + c;
                        const float se = inlier_s(M, pex[col], pex[TPN + col], pex[2 * TPN + col], pex[3 * TPN + col], pex[4 * TPN + col]) * (scf * scf);
                        const float* pf = reinterpret_cast<const float*>(pinfo + 4 * ch + (c >> 1));
                        const float ndu = pf[c & 1], ndv = pf[2 + (c & 1)];
                        const float xw = __uint_as_float(r[c]), yw = __uint_as_float(r[8 + c]), ww = __uint_as_float(r[16 + c]);
                        const float dx = fmaf(ndu, ww, xw), dy = fmaf(ndv, ww, yw);
                        const float sv = fmaf(dx, dx, fmaf(dy, dy, ww * -fabsf(ww)));
                        const float ratio = fabsf(sv - se) / bnd;
                        const bool real = ndu < 1.0e17f;
                        if (real && ratio == ratio && bnd < 1e30f && fabsf(se) <= 0.25f * ww * ww) calmax = fmaxf(calmax, ratio);
                        if (real && fabsf(sv) > bnd && ((__float_as_uint(sv) ^ __float_as_uint(se)) >> 31)) ++nmis;
                    }
                };
                K3CLK(e4);
                tc::tmem_ld8(tb, ra); tc::tmem_ld8(tb + TPN, ra + 8); tc::tmem_ld8(tb + 2 * TPN, ra + 16);
#pragma unroll
                for (int ch = 0; ch < NCH; ++ch) {
                    tc::tmem_ld_wait();
                    if (ch + 1 < NCH) {
                        const uint32_t tn = tb + 8u * (ch + 1);
                        if (ch & 1) { tc::tmem_ld8(tn, ra); tc::tmem_ld8(tn + TPN, ra + 8); tc::tmem_ld8(tn + 2 * TPN, ra + 16); }
                        else { tc::tmem_ld8(tn, rb); tc::tmem_ld8(tn + TPN, rb + 8); tc::tmem_ld8(tn + 2 * TPN, rb + 16); }
                    } else {
                        tc::fence_before_sync();                               // the whole set is in registers
                        __syncwarp();
                        if (lane == 0) tc::mbar_arrive(&acc_free[g]);
                    }
                    const bool flag = (ch & 1) ? eval8(rb, ch) : eval8(ra, ch);
                    if (CALIB) { if (ch & 1) calib8(rb, ch); else calib8(ra, ch); }
                    if (__any_sync(0xffffffffu, flag)) { if (ch & 1) exact8(rb, ch); else exact8(ra, ch); }
                }
                cnt += __popc(bits);
                K3CLK(e5);
                K3ACC(3, e4, e5);
            }
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&p_free[s]);                            // this warp is done with the stage's pixel / exact arrays
        }
        if (cur_f >= 0) {
            flush(cur_f);
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&h_free[fl & 1]);
        }
        if (PROF && warp == 0 && lane == 0)   // epilogue warp 0: [0] per-tile prologue, [1] waiting for the MMA, [3] loads + arithmetic
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
    static_assert(TPN == 64, "the epilogue's column split assumes 64-point tiles");
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
