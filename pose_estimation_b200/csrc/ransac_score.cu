// ransac_score.cu -- K1 + K2 (sample, solve) and K3 (score) of PnP-RANSAC (sm_100a).
//
// Replaces the iteration loop inside cv2.solvePnPRansac (tools/trainer.py:423-427): cv::RNG subset
// draw -> Philox4x32-10 keyed (seed, frame, hyp); 5-point EPnP -> P3P + 4th-point disambiguation
// (p3p.cuh); PnPRansacCallback::computeError + findInliers -> division-free reprojection test of
// EVERY correspondence.
//
//   ransac_hypothesis_kernel  one THREAD per (frame, hypothesis): Philox sample, P3P, R|t -> hypRt.
//       Register-heavy and divergent, but only F*H threads of ~4k instructions; kept apart so that
//       the scoring kernel stays lean (high occupancy, no solver registers).
//   ransac_score_kernel       one CTA per (frame, hypothesis slice): thread <-> hypothesis (3x4
//       projection matrix in 12 registers, NHL hypotheses per thread), correspondences staged per
//       chunk in shared memory as float4(X,Y,Z,U') + float(V') and read as warp-broadcast LDS.128,
//       so one shared-memory wavefront feeds 32*NHL tests.  The arithmetic is PACKED FP32x2
//       (Blackwell FFMA2/FMUL2, fma.rn.f32x2): two points per instruction, IEEE-rn per half, so
//       the bits equal the scalar contract in common.cuh::inlier_s.  Per hypothesis*point: 4.5 packed
//       + 5 scalar FMA-pipe instructions (the tail reads three 64-bit operands when packed, which costs
//       register-file bandwidth, so it is scalar) and one funnel shift that pushes the verdict's sign
//       bit into a 32-bit window (popcount every 8 points) -- no compare, no select, no reduction
//       in the loop.  Chunks are padded to a multiple of 8 with never-inlier sentinels so the loop
//       has no tail test.  FP32-pipe bound.
#include "p3p.cuh"
#include "ransac_plan.h"

namespace {

constexpr int HT = 128;  // threads per CTA of the hypothesis kernel

// exclusive prefix over the frames of  v(P) = P rounded up to 8 (TILES = false: the balanced scoring kernel cuts this
// padded index space into equal ranges, so every range and every frame segment starts at a multiple of 8) or of the
// number of PB_TC_TILE-point tiles (TILES = true: the tensor-core scoring kernel's work units) -> out[0..F]; one CTA
template <bool TILES>
__device__ void frame_prefix_scan(const ScoreArgs& a, int32_t* out) {
    __shared__ int wsum[HT / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (a.F + HT - 1) / HT;
    auto val = [&](int f) {
        const int P = max(0, min(a.count[f], a.cap));
        return TILES ? (P + PB_TC_TILE - 1) / PB_TC_TILE : (P + 7) & ~7;
    };
    int mine = 0;
    for (int k = 0; k < per; ++k) {
        const int f = tid * per + k;
        if (f < a.F) mine += val(f);
    }
    int incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += v;
    }
    __syncthreads();                                   // wsum may still be read by the previous scan
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int base = incl - mine;
    for (int w = 0; w < warp; ++w) base += wsum[w];
    for (int k = 0; k < per; ++k) {
        const int f = tid * per + k;
        if (f < a.F) { out[f] = base; base += val(f); }
    }
    if (tid == HT - 1) out[a.F] = base;
}

__global__ void __launch_bounds__(HT) ransac_hypothesis_kernel(ScoreArgs a) {
    if (blockIdx.x == 0 && a.prefix) {                                             // uniform per CTA
        frame_prefix_scan<false>(a, a.prefix);
        if (threadIdx.x == 0) a.prefix[a.F + 1] = 0;                               // the balanced scorer's unit counter
    }
    if (blockIdx.x == 0 && a.tprefix) frame_prefix_scan<true>(a, a.tprefix);
    const long long gid = (long long)blockIdx.x * HT + threadIdx.x;
    if (gid >= (long long)a.F * a.H) return;
    const int f = (int)(gid / a.H), h = (int)(gid - (long long)f * a.H);
    const int P = min(a.count[f], a.cap);
    const float* X = a.pts + (size_t)f * 5 * a.cap;
    const float* Y = X + a.cap;
    const float* Z = Y + a.cap;
    const float* U = Z + a.cap;
    const float* V = U + a.cap;
    float* o = a.hypRt + (size_t)gid * 12;
    bool valid = false;
    float R[9], t[3];
    if (P >= 4) {
        int idx[4];
        bool have;
        if (a.sample_idx) {
            const int32_t* si = a.sample_idx + (size_t)gid * 4;
            have = true;
#pragma unroll
            for (int k = 0; k < 4; ++k) { idx[k] = si[k]; have = have && idx[k] >= 0 && idx[k] < P; }
        } else {
            have = sample4(a.seed_lo, a.seed_hi, a.frame0 + (uint32_t)f, (uint32_t)h, P, idx);
        }
        if (have) {
            const float fx = a.K4[4 * f], fy = a.K4[4 * f + 1], cx = a.K4[4 * f + 2], cy = a.K4[4 * f + 3];
            float x[4][3], uv[4][2];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                x[k][0] = X[idx[k]]; x[k][1] = Y[idx[k]]; x[k][2] = Z[idx[k]];
                uv[k][0] = U[idx[k]]; uv[k][1] = V[idx[k]];
            }
            valid = p3p::solve(x, uv, fx, fy, cx, cy, R, t);
        }
    }
    if (valid) {
#pragma unroll
        for (int k = 0; k < 9; ++k) o[k] = R[k];
        o[9] = t[0]; o[10] = t[1]; o[11] = t[2];
    } else {
        o[11] = __int_as_float(0x7fffffff);  // t_z = NaN marks "no model"
    }
    if (a.cnt_all) a.cnt_all[gid] = valid ? 0 : -1;   // balanced scoring accumulates into these
}

typedef unsigned long long f32x2;   // two packed binary32 values (lo = even point, hi = odd point)
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
// opaque to the optimiser on purpose: a (v,v) pair must live in its own aligned register pair,
// otherwise the compiler keeps one copy and rebuilds the pair with moves inside the hot loop
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
    return d;
}
// inlier_s (common.cuh) for two points at once; nus/nvs are the NEGATED scaled pixel coordinates.
// q = w*|w| enters as  w * (-|w|) = -q  (sign flips are exact), so every half has inlier_s' bits.
__device__ __forceinline__ float lo32(f32x2 v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float hi32(f32x2 v) { return __uint_as_float((uint32_t)(v >> 32)); }
// Returns the two verdict words (sign bit = inlier).  The nine matrix FMAs are packed (scalar-broadcast matrix
// entry x packed point coordinate: 2-3 register reads per FFMA2); the five tail operations read three full
// 64-bit operands as FFMA2 (6 register reads, 3 issue cycles), so they are issued as scalar FFMA instead.
// Measured per warp instruction and SMSP on B200 (scratch/fma_lab.cu): FFMA 1.06 / 1.14 / 1.7 cycles with
// 1 / 2 / 3 operands not in the reuse cache; FFMA2 2.1 cycles with <= 128 fresh operand bits, 2.2 as used here
// (broadcast entry, point pair shared by the six rows in a row, accumulator pair), 2.8 for (32, 64, 64) fresh,
// 3.1 for three fresh pairs.  The register file, not the FMA lanes, bounds this loop: in isolation
// (scratch/k3_lab.cu) it runs at 18.5 cycles per warp-level hypothesis*point against the ideal 14; an all-packed
// tail reaches 17.7 there but is no faster inside the kernel (one more live pair per chain), scalar orders 19.7.
__device__ __forceinline__ void inlier_s2(const f32x2 m[12], f32x2 X, f32x2 Y, f32x2 Z, f32x2 nus, f32x2 nvs,
                                          uint32_t& s0, uint32_t& s1) {
    const f32x2 x = ffma2(m[0], X, ffma2(m[1], Y, ffma2(m[2], Z, m[3])));
    const f32x2 y = ffma2(m[4], X, ffma2(m[5], Y, ffma2(m[6], Z, m[7])));
    const f32x2 w = ffma2(m[8], X, ffma2(m[9], Y, ffma2(m[10], Z, m[11])));
    {
        const float w0 = lo32(w), dx0 = fmaf(lo32(nus), w0, lo32(x)), dy0 = fmaf(lo32(nvs), w0, lo32(y));
        const float nq0 = w0 * -fabsf(w0);
        s0 = __float_as_uint(fmaf(dx0, dx0, fmaf(dy0, dy0, nq0)));
        const float w1 = hi32(w), dx1 = fmaf(hi32(nus), w1, hi32(x)), dy1 = fmaf(hi32(nvs), w1, hi32(y));
        const float nq1 = w1 * -fabsf(w1);
        s1 = __float_as_uint(fmaf(dx1, dx1, fmaf(dy1, dy1, nq1)));
    }
}

constexpr int ST = 128;  // threads per scoring CTA: 128 x 64 regs x ~26 KB smem -> 8 CTAs per SM, so the
                         // 1024 CTAs of a 1024-frame batch are all resident (one wave, no tail)

template <int NHL>
__global__ void __launch_bounds__(ST) ransac_score_kernel(ScoreArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int CH = a.CH, Hs = a.Hs;
    // points are staged as PAIRS (2i, 2i+1): sA[i] = (X0,X1,Y0,Y1), sB[i] = (Z0,Z1,-U0',-U1'), sC[i] = (-V0',-V1')
    float4* sA = reinterpret_cast<float4*>(smem_raw);            // [CH/2]
    float4* sB = sA + CH / 2;                                    // [CH/2]
    float2* sC = reinterpret_cast<float2*>(sB + CH / 2);         // [CH/2]
    int* cnt = reinterpret_cast<int*>(sC + CH / 2);              // [Hs]
    __shared__ unsigned long long wkey[ST / 32];

    const int f = blockIdx.x / a.HS, s = blockIdx.x % a.HS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int P = min(a.count[f], a.cap);
    const float* X = a.pts + (size_t)f * 5 * a.cap;
    const float* Y = X + a.cap;
    const float* Z = Y + a.cap;
    const float* U = Z + a.cap;
    const float* V = U + a.cap;
    const float fx = a.K4[4 * f], fy = a.K4[4 * f + 1], cx = a.K4[4 * f + 2], cy = a.K4[4 * f + 3];
    const int h0 = s * Hs;
    const int PP = a.PP;                       // >1 only when Hs < ST (then NHL == 1)
    const int part = (PP > 1) ? tid / Hs : 0;  // warp-uniform: Hs is a multiple of 32

    // ---- this thread's hypotheses: K[R|t] in threshold units, each entry duplicated into both halves
    f32x2 m[NHL][12];
    int c[NHL];
#pragma unroll
    for (int j = 0; j < NHL; ++j) {
        const int hl = (PP > 1) ? (tid % Hs) : (tid + j * ST);
        const int h = h0 + hl;
        c[j] = 0;
        bool valid = false;
        float mj[12];
        if (h < a.H) {
            const float4* o = reinterpret_cast<const float4*>(a.hypRt + ((size_t)f * a.H + h) * 12);
            const float4 r0 = o[0], r1 = o[1], r2 = o[2];
            valid = (r2.w == r2.w);
            if (valid) {
                const float R[9] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x};
                const float t[3] = {r2.y, r2.z, r2.w};
                make_proj_s(fx, fy, cx, cy, R, t, a.inv_thr, mj);
            }
        }
        if (!valid) {
#pragma unroll
            for (int k = 0; k < 12; ++k) mj[k] = 0.0f;
            mj[11] = -1.0f;  // w < 0 for every point: never an inlier
        }
#pragma unroll
        for (int k = 0; k < 12; ++k) m[j][k] = pack2(mj[k], mj[k]);
        if (part == 0 && hl < Hs) cnt[hl] = valid ? 0 : -1;
    }

    // ---- K3: every correspondence against this thread's hypotheses
    const float ninf = __int_as_float(0xff800000);
    float* sAf = reinterpret_cast<float*>(sA);
    float* sBf = reinterpret_cast<float*>(sB);
    float* sCf = reinterpret_cast<float*>(sC);
    for (int c0 = 0; c0 < P; c0 += CH) {
        const int n = min(CH, P - c0);
        const int npp = ((n + 8 * PP - 1) / (8 * PP)) * 8;   // points per partition, multiple of 8
        __syncthreads();                                     // previous chunk consumed / cnt initialised
        for (int i = tid; i < npp * PP; i += ST) {
            const int o4 = (i >> 1) * 4 + (i & 1);
            if (i < n) {
                sAf[o4] = X[c0 + i]; sAf[o4 + 2] = Y[c0 + i];
                sBf[o4] = Z[c0 + i]; sBf[o4 + 2] = -(U[c0 + i] * a.inv_thr);
                sCf[i] = -(V[c0 + i] * a.inv_thr);
            } else {                                         // sentinel: s = +inf or NaN, never an inlier
                sAf[o4] = 0.0f; sAf[o4 + 2] = 0.0f;
                sBf[o4] = 0.0f; sBf[o4 + 2] = ninf;
                sCf[i] = ninf;
            }
        }
        __syncthreads();
        const int pb = (part * npp) >> 1;                    // first pair of this partition
        const ulonglong2* a2 = reinterpret_cast<const ulonglong2*>(sA) + pb;
        const ulonglong2* b2 = reinterpret_cast<const ulonglong2*>(sB) + pb;
        const ulonglong2* c2 = reinterpret_cast<const ulonglong2*>(sC + pb);   // two pairs per 16 bytes
#pragma unroll 1
        for (int i = 0; i < (npp >> 1); i += 4) {            // 4 pairs = 8 points per iteration
            const ulonglong2 v01 = c2[i >> 1], v23 = c2[(i >> 1) + 1];
            const f32x2 nvs[4] = {v01.x, v01.y, v23.x, v23.y};
            uint32_t acc[NHL];
#pragma unroll
            for (int j = 0; j < NHL; ++j) acc[j] = 0u;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const ulonglong2 pa = a2[i + k], pbv = b2[i + k];
#pragma unroll
                for (int j = 0; j < NHL; ++j) {
                    uint32_t s0, s1;
                    inlier_s2(m[j], pa.x, pa.y, pbv.x, pbv.y, nvs[k], s0, s1);
                    acc[j] = __funnelshift_l(s0, acc[j], 1);
                    acc[j] = __funnelshift_l(s1, acc[j], 1);
                }
            }
#pragma unroll
            for (int j = 0; j < NHL; ++j) c[j] += __popc(acc[j]);
        }
    }
    if (P <= 0) __syncthreads();               // cnt initialised even when there was no chunk
#pragma unroll
    for (int j = 0; j < NHL; ++j) {
        const int hl = (PP > 1) ? (tid % Hs) : (tid + j * ST);
        if (hl < Hs && c[j] > 0) atomicAdd(&cnt[hl], c[j]);   // c > 0 only for valid hypotheses
    }
    __syncthreads();

    // ---- slice argmax: max count, lowest hypothesis index on ties
    unsigned long long key = 0ull;
    for (int hl = tid; hl < Hs; hl += ST) {
        const int h = h0 + hl;
        if (h < a.H) {
            const int cv = cnt[hl];
            if (a.hyp_counts) a.hyp_counts[(size_t)f * a.H + h] = cv;
            const unsigned long long k2 =
                ((unsigned long long)(uint32_t)(cv + 1) << 32) | (0xFFFFFFFFu - (uint32_t)h);
            key = k2 > key ? k2 : key;
        }
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, key, off);
        key = o > key ? o : key;
    }
    if (lane == 0) wkey[warp] = key;
    __syncthreads();
    if (tid == 0) {
        unsigned long long k = wkey[0];
        for (int w = 1; w < ST / 32; ++w) k = wkey[w] > k ? wkey[w] : k;
        a.slice_best[(size_t)f * a.HS + s] = k;
    }
}

// inlier_s2 with the pixel coordinates as they are in the correspondence list (already in threshold units): the
// negation is an operand modifier of the FMA, so the bits equal inlier_s2's / common.cuh::inlier_s.
__device__ __forceinline__ void inlier_s2u(const f32x2 m[12], f32x2 X, f32x2 Y, f32x2 Z, f32x2 us, f32x2 vs, uint32_t& s0,
                                           uint32_t& s1) {
    const f32x2 x = ffma2(m[0], X, ffma2(m[1], Y, ffma2(m[2], Z, m[3])));
    const f32x2 y = ffma2(m[4], X, ffma2(m[5], Y, ffma2(m[6], Z, m[7])));
    const f32x2 w = ffma2(m[8], X, ffma2(m[9], Y, ffma2(m[10], Z, m[11])));
    const float w0 = lo32(w), dx0 = fmaf(-lo32(us), w0, lo32(x)), dy0 = fmaf(-lo32(vs), w0, lo32(y));
    const float nq0 = w0 * -fabsf(w0);
    s0 = __float_as_uint(fmaf(dx0, dx0, fmaf(dy0, dy0, nq0)));
    const float w1 = hi32(w), dx1 = fmaf(-hi32(us), w1, hi32(x)), dy1 = fmaf(-hi32(vs), w1, hi32(y));
    const float nq1 = w1 * -fabsf(w1);
    s1 = __float_as_uint(fmaf(dx1, dx1, fmaf(dy1, dy1, nq1)));
}

__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(void* dst_smem, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

constexpr int SUB = PB_BAL_SUB;      // points per staging buffer

// Balanced scoring: gridDim.x CTAs, all resident, each takes an EQUAL range of the batch's concatenated point list
// (frames padded to a multiple of 8 points; a range usually touches two frames).  The range is worked off in
// sub-chunks of <= 256 points of one frame, staged as they lie in the correspondence list (SoA) by asynchronous
// 16-byte copies into one of two shared-memory buffers while the previous sub-chunk is being scored; the frames'
// prefix comes from the hypothesis kernel (one scan per launch, not one per CTA).  For every frame segment the CTA
// holds that frame's hypotheses (two per thread) and adds its counts to cnt_all[f][h]; inlier counts are integer sums,
// so the result does not depend on the cut.  The arg-max is taken by the finalize kernel.
#ifndef PB_K3_TIMELINE
#define PB_K3_TIMELINE 0     // 1 (measurement builds only, scratch/k3_timeline.py): per-CTA start / end times and SM ids
#endif
#if PB_K3_TIMELINE
__device__ unsigned long long g_k3_timeline[4096][4];
__device__ __forceinline__ unsigned long long k3_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#endif

template <int NHL>
__global__ void __launch_bounds__(ST, 6) ransac_score_bal_kernel(ScoreArgs a) {
    __shared__ __align__(16) float sbuf[2][5][SUB];
    const int tid = threadIdx.x;
#if PB_K3_TIMELINE
    if (tid == 0 && blockIdx.x < 4096) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        g_k3_timeline[blockIdx.x][0] = k3_now();
        g_k3_timeline[blockIdx.x][1] = 0;
        g_k3_timeline[blockIdx.x][2] = smid;
        g_k3_timeline[blockIdx.x][3] = clock64();
    }
#endif
    const int F = a.F;
    const int* __restrict__ prefix = a.prefix;
    const int T = prefix[F];
#if PB_BAL_DYN_PCT > 0
    // every CTA takes a fixed range of (100 - PB_BAL_DYN_PCT) % of an equal share, then units of SUB points from the rest
    // through a counter (zeroed by the hypothesis kernel): the warp schedulers favour the oldest CTA of an SM, so CTAs of
    // equal ranges end one after the other and the SM runs its last ranges with few warps
    __shared__ int s_unit;
    int U = (int)((long long)T * (100 - PB_BAL_DYN_PCT) / 100 / (int)gridDim.x) & ~7;
    const int pool0 = U * (int)gridDim.x;
    int g0 = blockIdx.x * U, g1 = g0 + U;
#else
    int U = (T + (int)gridDim.x - 1) / (int)gridDim.x;
    U = (U + 7) & ~7;
    const long long g0l = (long long)blockIdx.x * U;
    if (g0l >= T) return;
    const int g0 = (int)g0l, g1 = min(T, g0 + U);
#endif
    int f = 0;
    // last frame with prefix[f] <= g (frames without points share a prefix value: take the last of them)
    auto find_frame = [&](int g) {
        int lo = 0, hi = F;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (prefix[mid] <= g) lo = mid; else hi = mid;
        }
        return lo;
    };
    const bool scale = a.inv_thr != 1.0f;
    const float ninf = __int_as_float(0x7f800000);

    // sub-chunk descriptor: frame sf, first point sp0 (multiple of 8), padded length sn (multiple of 8, <= SUB)
    auto describe = [&](int g, int& fr, int& sf, int& sp0, int& sn) -> bool {
        if (g >= g1) return false;
        while (fr + 1 < F && prefix[fr + 1] <= g) ++fr;
        sf = fr; sp0 = g - prefix[fr];
        sn = min(min(SUB, prefix[fr + 1] - g), g1 - g);
        return true;
    };
    // asynchronous staging of a sub-chunk: groups of 4 points, 16 bytes per array; beyond the frame's count: sentinels
    auto stage = [&](int b, int sf, int sp0, int sn) {
        const int P = max(0, min(a.count[sf], a.cap));
        const float* base = a.pts + (size_t)sf * 5 * a.cap + sp0;
        for (int e = tid; e < (sn >> 2) * 5; e += ST) {
            const int arr = e / (sn >> 2), i = (e - arr * (sn >> 2)) << 2;
            float* dst = &sbuf[b][arr][i];
            const float* src = base + (size_t)arr * a.cap + i;
            if (sp0 + i + 4 <= P) cp_async16(dst, src);
            else {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (sp0 + i + k < P) cp_async4(dst + k, src + k);
                    else dst[k] = arr >= 3 ? ninf : 0.0f;          // u = v = +inf: s = +inf or NaN, never an inlier
                }
            }
        }
        cp_async_commit();
    };
    // pixel coordinates into threshold units (only when thr != 1): every thread scales what it copied itself
    auto scale_uv = [&](int b, int sf, int sp0, int sn) {
        const int P = max(0, min(a.count[sf], a.cap));
        for (int e = tid; e < (sn >> 2) * 5; e += ST) {
            const int arr = e / (sn >> 2), i = (e - arr * (sn >> 2)) << 2;
            if (arr < 3) continue;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (sp0 + i + k < P) sbuf[b][arr][i + k] = sbuf[b][arr][i + k] * a.inv_thr;
        }
    };

    f32x2 m[NHL][12];
    int c[NHL];
#pragma unroll
    for (int j = 0; j < NHL; ++j) c[j] = 0;
    int cur_f = -1;
    int sf, sp0, sn, nf = 0, np0 = 0, nn = 0;
#if PB_BAL_DYN_PCT > 0
    for (bool first = true;; first = false) {
    if (!first) {
        if (tid == 0) s_unit = atomicAdd(a.prefix + F + 1, 1);
        __syncthreads();
        const long long gl = (long long)pool0 + (long long)s_unit * SUB;
        __syncthreads();                                           // s_unit is read before the next fetch overwrites it
        if (gl >= T) break;
        g0 = (int)gl; g1 = min(T, g0 + SUB);
    }
    if (g1 <= g0) continue;
#endif
    f = find_frame(g0);
    int g = g0;
    bool have = describe(g, f, sf, sp0, sn);
    if (have) stage(0, sf, sp0, sn);
    for (int k = 0; have; ++k) {
        const int b = k & 1;
        g += sn;
        const bool more = describe(g, f, nf, np0, nn);
        if (more) stage(b ^ 1, nf, np0, nn);                       // in flight under this sub-chunk's scoring
        if (sf != cur_f) {                                         // (uniform) another frame: flush, load its hypotheses
            if (cur_f >= 0) {
#pragma unroll
                for (int j = 0; j < NHL; ++j) {
                    const int h = tid + j * ST;
                    if (h < a.H && c[j] > 0) atomicAdd(&a.cnt_all[(size_t)cur_f * a.H + h], c[j]);
                    c[j] = 0;
                }
            }
            cur_f = sf;
            const float fx = a.K4[4 * sf], fy = a.K4[4 * sf + 1], cx = a.K4[4 * sf + 2], cy = a.K4[4 * sf + 3];
#pragma unroll
            for (int j = 0; j < NHL; ++j) {
                const int h = tid + j * ST;
                bool valid = false;
                float mj[12];
                if (h < a.H) {
                    const float4* o = reinterpret_cast<const float4*>(a.hypRt + ((size_t)sf * a.H + h) * 12);
                    const float4 r0 = o[0], r1 = o[1], r2 = o[2];
                    valid = (r2.w == r2.w);
                    if (valid) {
                        const float R[9] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x};
                        const float t[3] = {r2.y, r2.z, r2.w};
                        make_proj_s(fx, fy, cx, cy, R, t, a.inv_thr, mj);
                    }
                }
                if (!valid) {
#pragma unroll
                    for (int q = 0; q < 12; ++q) mj[q] = 0.0f;
                    mj[11] = -1.0f;                                // w < 0 for every point: never an inlier
                }
#pragma unroll
                for (int q = 0; q < 12; ++q) m[j][q] = pack2(mj[q], mj[q]);
            }
        }
        if (more) cp_async_wait<1>(); else cp_async_wait<0>();     // this sub-chunk's copies have landed (own thread)
        if (scale) scale_uv(b, sf, sp0, sn);
        __syncthreads();                                           // ... and everybody else's
        const ulonglong2* X2 = reinterpret_cast<const ulonglong2*>(sbuf[b][0]);
        const ulonglong2* Y2 = reinterpret_cast<const ulonglong2*>(sbuf[b][1]);
        const ulonglong2* Z2 = reinterpret_cast<const ulonglong2*>(sbuf[b][2]);
        const ulonglong2* U2 = reinterpret_cast<const ulonglong2*>(sbuf[b][3]);
        const ulonglong2* V2 = reinterpret_cast<const ulonglong2*>(sbuf[b][4]);
#pragma unroll 1
        for (int i = 0; i < (sn >> 2); i += 2) {                   // 8 points per iteration: two 128-bit loads per array
            uint32_t acc[NHL];
#pragma unroll
            for (int j = 0; j < NHL; ++j) acc[j] = 0u;
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const ulonglong2 xv = X2[i + q], yv = Y2[i + q], zv = Z2[i + q], uv = U2[i + q], vv = V2[i + q];
#pragma unroll
                for (int j = 0; j < NHL; ++j) {
                    uint32_t s0, s1;
                    inlier_s2u(m[j], xv.x, yv.x, zv.x, uv.x, vv.x, s0, s1);
                    acc[j] = __funnelshift_l(s0, acc[j], 1);
                    acc[j] = __funnelshift_l(s1, acc[j], 1);
                    inlier_s2u(m[j], xv.y, yv.y, zv.y, uv.y, vv.y, s0, s1);
                    acc[j] = __funnelshift_l(s0, acc[j], 1);
                    acc[j] = __funnelshift_l(s1, acc[j], 1);
                }
            }
#pragma unroll
            for (int j = 0; j < NHL; ++j) c[j] += __popc(acc[j]);
        }
        __syncthreads();                                           // buffer b is free for the sub-chunk after next
        have = more; sf = nf; sp0 = np0; sn = nn;
    }
#if PB_BAL_DYN_PCT > 0
    }
#endif
    if (cur_f >= 0) {
#pragma unroll
        for (int j = 0; j < NHL; ++j) {
            const int h = tid + j * ST;
            if (h < a.H && c[j] > 0) atomicAdd(&a.cnt_all[(size_t)cur_f * a.H + h], c[j]);   // c > 0 only for valid hypotheses
        }
    }
#if PB_K3_TIMELINE
    __syncthreads();
    if (tid == 0 && blockIdx.x < 4096) {
        g_k3_timeline[blockIdx.x][1] = k3_now();
        g_k3_timeline[blockIdx.x][3] = clock64() - g_k3_timeline[blockIdx.x][3];
    }
#endif
}

#if PB_K3_TIMELINE
}  // namespace
extern "C" int pb_debug_k3_timeline(unsigned long long* out, int n) {
    return cudaMemcpyFromSymbol(out, g_k3_timeline, (size_t)n * 4 * sizeof(unsigned long long)) == cudaSuccess ? 0 : -1;
}
namespace {
#endif

template <int NHL>
int launch_bal(const ScoreArgs& a, const Plan& p, cudaStream_t st) {
    ransac_score_bal_kernel<NHL><<<p.bal_grid, ST, 0, st>>>(a);
    return pb_check_launch("ransac_score_bal_kernel");
}

template <int NHL>
int launch(const ScoreArgs& a, const Plan& p, int F, cudaStream_t st) {
    cudaFuncSetAttribute(ransac_score_kernel<NHL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem_score);
    ransac_score_kernel<NHL><<<F * p.HS, ST, p.smem_score, st>>>(a);
    return pb_check_launch("ransac_score_kernel");
}

}  // namespace

int pb_launch_score(const ScoreArgs& a, const Plan& p, int F, cudaStream_t st) {
    const long long nh = (long long)F * a.H;
    pb_stage_mark(0, st);
    ransac_hypothesis_kernel<<<(unsigned)((nh + HT - 1) / HT), HT, 0, st>>>(a);
    if (pb_check_launch("ransac_hypothesis_kernel")) return 1;
    pb_stage_mark(1, st);
    if (a.fstats && a.cnt_all) return pb_launch_score_mma(a, a.tc_mode == 1, st);
    if (a.tprefix && a.cnt_all) return pb_launch_score_tc(a, a.tc_mode, st);
    if (p.bal && a.cnt_all) return p.bal_nhl == 2 ? launch_bal<2>(a, p, st) : launch_bal<1>(a, p, st);
    switch (p.NHL) {
        case 1: return launch<1>(a, p, F, st);
        case 2: return launch<2>(a, p, F, st);
        default: pb_set_error("ransac_score: unsupported NHL=%d", p.NHL); return 1;
    }
}
