// ransac_score.cu -- K1 + K2 + K3 of PnP-RANSAC (sm_100a): sample, solve, score.
//
// One CTA per (frame, hypothesis slice).  Replaces the iteration loop inside cv2.solvePnPRansac
// (tools/trainer.py:423-427): cv::RNG subset draw -> Philox4x32-10 keyed (seed, frame, hyp);
// 5-point EPnP -> P3P + 4th-point disambiguation (p3p.cuh); PnPRansacCallback::computeError +
// findInliers -> division-free reprojection test of EVERY correspondence.
//
// K3 layout: thread <-> hypothesis (its 3x4 projection matrix lives in 12 registers per
// hypothesis, NHL hypotheses per thread), correspondences staged once per chunk in shared memory
// as float4(X,Y,Z,U)+float(V) and read as warp-broadcast LDS.128 + LDS.32, so one shared-memory
// wavefront feeds 32*NHL hypothesis*point tests and the inlier counter is a private register (no
// ballot, no reduction).  FP32-pipe bound: 15 FMA-pipe + 2 ALU ops per hypothesis*point.
#include "p3p.cuh"
#include "ransac_plan.h"

namespace {

template <int NHL>
__global__ void __launch_bounds__(PB_THREADS) ransac_score_kernel(ScoreArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int CH = a.CH, Hs = a.Hs;
    float4* sp4 = reinterpret_cast<float4*>(smem_raw);          // [CH] X,Y,Z,U
    float* spv = reinterpret_cast<float*>(sp4 + CH);            // [CH] V
    float* Msm = spv + CH;                                      // [Hs][12]
    int* cnt = reinterpret_cast<int*>(Msm + Hs * 12);           // [Hs]
    __shared__ unsigned long long wkey[PB_THREADS / 32];

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

    // ---- K1 + K2: one hypothesis per thread
    for (int hl = tid; hl < Hs; hl += PB_THREADS) {
        const int h = h0 + hl;
        float M[12];
#pragma unroll
        for (int k = 0; k < 12; ++k) M[k] = 0.0f;
        M[11] = -1.0f;  // w < 0 for every point: never an inlier
        bool valid = false;
        if (h < a.H && P >= 4) {
            int idx[4];
            bool have;
            if (a.sample_idx) {
                const int32_t* si = a.sample_idx + ((size_t)f * a.H + h) * 4;
                have = true;
#pragma unroll
                for (int k = 0; k < 4; ++k) { idx[k] = si[k]; have = have && idx[k] >= 0 && idx[k] < P; }
            } else {
                have = sample4(a.seed_lo, a.seed_hi, a.frame0 + (uint32_t)f, (uint32_t)h, P, idx);
            }
            if (have) {
                float x[4][3], uv[4][2];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    x[k][0] = X[idx[k]]; x[k][1] = Y[idx[k]]; x[k][2] = Z[idx[k]];
                    uv[k][0] = U[idx[k]]; uv[k][1] = V[idx[k]];
                }
                float R[9], t[3];
                valid = p3p::solve(x, uv, fx, fy, cx, cy, R, t);
                if (valid) {
                    make_proj(fx, fy, cx, cy, R, t, M);
                    float* o = a.hypRt + ((size_t)f * a.H + h) * 12;
#pragma unroll
                    for (int k = 0; k < 9; ++k) o[k] = R[k];
                    o[9] = t[0]; o[10] = t[1]; o[11] = t[2];
                }
            }
        }
#pragma unroll
        for (int k = 0; k < 12; ++k) Msm[hl * 12 + k] = M[k];
        cnt[hl] = valid ? 0 : -1;
    }
    __syncthreads();

    // ---- K3: every correspondence against this thread's hypotheses
    const int PP = a.PP;                       // >1 only when Hs < PB_THREADS (then NHL == 1)
    const int part = (PP > 1) ? tid / Hs : 0;  // warp-uniform: Hs is a multiple of 32
    float m[NHL][12];
    int c[NHL];
#pragma unroll
    for (int j = 0; j < NHL; ++j) {
        const int hl = (PP > 1) ? (tid % Hs) : (tid + j * PB_THREADS);
        c[j] = 0;
#pragma unroll
        for (int k = 0; k < 12; ++k) m[j][k] = Msm[hl * 12 + k];
    }
    for (int c0 = 0; c0 < P; c0 += CH) {
        const int n = min(CH, P - c0);
        if (c0 > 0) __syncthreads();           // previous chunk fully consumed
        for (int i = tid; i < n; i += PB_THREADS) {
            sp4[i] = make_float4(X[c0 + i], Y[c0 + i], Z[c0 + i], U[c0 + i]);
            spv[i] = V[c0 + i];
        }
        __syncthreads();
#pragma unroll 4
        for (int i = part; i < n; i += PP) {
            const float4 p = sp4[i];
            const float pv = spv[i];
#pragma unroll
            for (int j = 0; j < NHL; ++j) c[j] += inlier_test(m[j], p.x, p.y, p.z, p.w, pv, a.thr2) ? 1 : 0;
        }
    }
#pragma unroll
    for (int j = 0; j < NHL; ++j) {
        const int hl = (PP > 1) ? (tid % Hs) : (tid + j * PB_THREADS);
        if (hl < Hs && cnt[hl] >= 0 && c[j] > 0) atomicAdd(&cnt[hl], c[j]);
    }
    __syncthreads();

    // ---- slice argmax: max count, lowest hypothesis index on ties
    unsigned long long key = 0ull;
    for (int hl = tid; hl < Hs; hl += PB_THREADS) {
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
        for (int w = 1; w < PB_THREADS / 32; ++w) k = wkey[w] > k ? wkey[w] : k;
        a.slice_best[(size_t)f * a.HS + s] = k;
    }
}

template <int NHL>
int launch(const ScoreArgs& a, const Plan& p, int F, cudaStream_t st) {
    cudaFuncSetAttribute(ransac_score_kernel<NHL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem_score);
    ransac_score_kernel<NHL><<<F * p.HS, PB_THREADS, p.smem_score, st>>>(a);
    return pb_check_launch("ransac_score_kernel");
}

}  // namespace

int pb_launch_score(const ScoreArgs& a, const Plan& p, int F, cudaStream_t st) {
    switch (p.NHL) {
        case 1: return launch<1>(a, p, F, st);
        case 2: return launch<2>(a, p, F, st);
        case 4: return launch<4>(a, p, F, st);
        default: pb_set_error("ransac_score: unsupported NHL=%d", p.NHL); return 1;
    }
}
