// sim3.cu -- 3D-3D RANSAC + Umeyama similarity refit, sm_100a.
//
// Replaces lib/transform/umeyama.py:8-62 (estimateSimilarityTransform: 128 x {5 random
// correspondences -> Umeyama -> residual test of every pair}, early stop by confidence, Umeyama refit
// on the inliers) and :65-97 (estimateSimilarityUmeyama).  One launch sequence handles B independent problems:
//
//   sim3_prepare_kernel     CTA per problem: source centroid, InlierT = 2 max|s - c| / 10 (:17-20)
//   sim3_hypothesis_kernel  THREAD per (problem, iteration): 5 indices with replacement from a counter RNG
//                           (np.random.randint(n, size=5), :32) or the caller's index sets; Umeyama in FP64
//   sim3_score_kernel       CTA per (problem, 128 iterations): thread <-> hypothesis, all pairs broadcast
//                           from shared memory, |t - (sR s + T)|^2 < (scale InlierT)^2 in FP32 (:34-38)
//   sim3_finalize_kernel    CTA per problem: replays the reference's sequential loop over the iteration
//                           counts (strict '>' update :41, confidence break :48), then the refit (:54-57)
//
// Umeyama without an SVD: the proper rotation maximising trace(R^T Cov) (what U diag(1,1,det) Vh is, :83-89)
// is the top eigenvector of Horn's 4x4 quaternion matrix, found with a fixed number of cyclic Jacobi sweeps
// in FP64; the eigenvalue is sum(D) of :92, so scale = lambda / var(source).  Every reduction runs in a
// fixed order (256 virtual threads, xor butterfly per 32, groups in order), so the CPU oracle of the tests
// reproduces every output bit.
//
// rule = 1 is the OLDER selection rule of lib/transform/trans.py:9-123 (imported by tools/script/eval.py): 100
// iterations, PassThreshold = max(|t|/|s|, |s|/|t|) of the mean point norms, a hypothesis is kept when its TOTAL
// residual sqrt(sum r_i^2) is strictly smaller than the best so far (not when it has more inliers), the loop stops
// once that residual is below PassThreshold / 100, and InlierRatio counts the non-zero inlier INDICES
// (np.count_nonzero of np.where's tuple, :87-88).  trans.py's Umeyama returns the TRANSPOSED rotation and
// puts it into OutTransform (:104, :114) while the translation is the one of the untransposed rotation, so its
// hypotheses are scored with  s R^T x + T ; that is reproduced as it stands (FP64, like numpy).
//
// pb_kabsch_fit / pb_icp_refine: the batched rigid fit of version/transparent/lib/declarative/leastsquares.py:10-48
// (same Horn eigen-solve, no scale) and the KNN -> gather -> fit -> compose loop of
// version/transparent/scripts/train_declarative.py:80-100.
#include "common.cuh"

namespace {

constexpr int S3_T = 256;          // threads of the prepare / finalize CTAs (= virtual threads of the sums)
constexpr int S3_ST = 128;         // hypotheses per scoring CTA
constexpr int S3_CH = 1024;        // pairs staged per chunk
constexpr uint32_t S3_TAG = 0x53494D33u;   // Philox counter word: "SIM3"

struct Sim3Args {
    const float* src; const float* dst; const int32_t* count;
    int B, cap, H;
    uint32_t seed_lo, seed_hi, prob0;
    float confidence, min_ratio;
    const int32_t* sample_idx;
    double* inlier_t;              // [B]
    float* hyp;                    // [B,H,16]: sR (9), T (3), thr^2, valid, -, -
    int32_t* hyp_count;            // [B,H]
    float* scale; float* R; float* t; int32_t* n_inl; int32_t* best_hyp; int32_t* status; uint32_t* bitmap;
    int fit_all;                   // pb_umeyama_fit: no hypotheses, every pair is an inlier
    int rule;                      // 0 umeyama.py, 1 trans.py
    double* hypd;                  // rule 1: [B,H,16] doubles: s R^T (9), T (3), valid, -, -, -
    double* residual;              // rule 1: [B,H] total residual sqrt(sum r^2)
    double* pass_t;                // rule 1: [B] PassThreshold
    int32_t* hyp_inl;              // rule 1: [B,H] number of inliers (every index)
};

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v = v + __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}
// canonical sum of the 256 per-thread partials: butterfly per warp, then the 8 warp sums in order
__device__ __forceinline__ double block_sum_d(double v, double* red) {
    v = warp_sum_d(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = red[0];
#pragma unroll
    for (int w = 1; w < S3_T / 32; ++w) s = s + red[w];
    return s;
}

// ---- Umeyama from moments ---------------------------------------------------------------------
// ms, mt: means; C[a][b] = mean (t_a - mt_a)(s_b - ms_b); varp = mean |s - ms|^2.
// Returns false when the fit is undefined (zero variance / non-finite input).
__device__ bool umeyama_core(const double ms[3], const double mt[3], const double C[3][3], double varp,
                             double& scale, double R[9], double T[3]) {
    // Horn: S_ab = mean s_a t_b = C[b][a]
    const double Sxx = C[0][0], Sxy = C[1][0], Sxz = C[2][0];
    const double Syx = C[0][1], Syy = C[1][1], Syz = C[2][1];
    const double Szx = C[0][2], Szy = C[1][2], Szz = C[2][2];
    double a[4][4], v[4][4];
    a[0][0] = Sxx + Syy + Szz; a[0][1] = Syz - Szy;        a[0][2] = Szx - Sxz;         a[0][3] = Sxy - Syx;
    a[1][1] = Sxx - Syy - Szz; a[1][2] = Sxy + Syx;        a[1][3] = Szx + Sxz;
    a[2][2] = -Sxx + Syy - Szz; a[2][3] = Syz + Szy;
    a[3][3] = -Sxx - Syy + Szz;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (j < i) a[i][j] = a[j][i];
            v[i][j] = (i == j) ? 1.0 : 0.0;
        }
#pragma unroll 1
    for (int sweep = 0; sweep < 10; ++sweep) {
#pragma unroll
        for (int p = 0; p < 3; ++p)
#pragma unroll
            for (int q = p + 1; q < 4; ++q) {
                const double apq = a[p][q];
                if (apq != 0.0) {
                    const double theta = (a[q][q] - a[p][p]) / (2.0 * apq);
                    const double tt = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                    const double c = 1.0 / sqrt(tt * tt + 1.0), s = tt * c;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {          // A <- A J
                        const double akp = a[k][p], akq = a[k][q];
                        a[k][p] = c * akp - s * akq;
                        a[k][q] = s * akp + c * akq;
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k) {          // A <- J^T A
                        const double apk = a[p][k], aqk = a[q][k];
                        a[p][k] = c * apk - s * aqk;
                        a[q][k] = s * apk + c * aqk;
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k) {          // V <- V J
                        const double vkp = v[k][p], vkq = v[k][q];
                        v[k][p] = c * vkp - s * vkq;
                        v[k][q] = s * vkp + c * vkq;
                    }
                }
            }
    }
    int best = 0;
    double lam = a[0][0];
#pragma unroll
    for (int i = 1; i < 4; ++i)
        if (a[i][i] > lam) { lam = a[i][i]; best = i; }
    double q[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) q[k] = best == 0 ? v[k][0] : best == 1 ? v[k][1] : best == 2 ? v[k][2] : v[k][3];
    const double nn = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    if (!(nn > 0.0) || !(varp > 0.0) || !isfinite(lam) || !isfinite(varp)) return false;
    const double sg = q[0] < 0.0 ? -1.0 : 1.0;
    const double w = sg * q[0] / nn, x = sg * q[1] / nn, y = sg * q[2] / nn, z = sg * q[3] / nn;
    R[0] = 1.0 - 2.0 * (y * y + z * z); R[1] = 2.0 * (x * y - w * z);       R[2] = 2.0 * (x * z + w * y);
    R[3] = 2.0 * (x * y + w * z);       R[4] = 1.0 - 2.0 * (x * x + z * z); R[5] = 2.0 * (y * z - w * x);
    R[6] = 2.0 * (x * z - w * y);       R[7] = 2.0 * (y * z + w * x);       R[8] = 1.0 - 2.0 * (x * x + y * y);
    scale = lam / varp;
#pragma unroll
    for (int i = 0; i < 3; ++i)
        T[i] = mt[i] - scale * (R[3 * i] * ms[0] + R[3 * i + 1] * ms[1] + R[3 * i + 2] * ms[2]);
    return isfinite(scale) && isfinite(T[0]) && isfinite(T[1]) && isfinite(T[2]);
}

// FP32 residual test shared by the scoring and finalize kernels: |t - (A s + T)|^2 < thr2
__device__ __forceinline__ bool sim3_inlier(const float A[9], const float T[3], float thr2, float sx, float sy, float sz,
                                            float tx, float ty, float tz) {
    const float dx = tx - fmaf(A[0], sx, fmaf(A[1], sy, fmaf(A[2], sz, T[0])));
    const float dy = ty - fmaf(A[3], sx, fmaf(A[4], sy, fmaf(A[5], sz, T[1])));
    const float dz = tz - fmaf(A[6], sx, fmaf(A[7], sy, fmaf(A[8], sz, T[2])));
    return fmaf(dx, dx, fmaf(dy, dy, dz * dz)) < thr2;
}

__global__ void __launch_bounds__(S3_T) sim3_prepare_kernel(Sim3Args a) {
    __shared__ double red[S3_T / 32];
    const int b = blockIdx.x, tid = threadIdx.x;
    const int n = min(a.count[b], a.cap);
    const float* s = a.src + (size_t)b * a.cap * 3;
    double p[3] = {0.0, 0.0, 0.0};
    for (int i = tid; i < n; i += S3_T)
#pragma unroll
        for (int c = 0; c < 3; ++c) p[c] = p[c] + (double)s[3 * (size_t)i + c];
    double m[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) m[c] = block_sum_d(p[c], red) / (double)max(n, 1);
    double d2 = 0.0;
    for (int i = tid; i < n; i += S3_T) {
        const double x = (double)s[3 * (size_t)i] - m[0], y = (double)s[3 * (size_t)i + 1] - m[1],
                     z = (double)s[3 * (size_t)i + 2] - m[2];
        d2 = fmax(d2, x * x + y * y + z * z);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) d2 = fmax(d2, __shfl_xor_sync(0xffffffffu, d2, off));
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = d2;
    __syncthreads();
    if (tid == 0) {
        double mx = red[0];
        for (int w = 1; w < S3_T / 32; ++w) mx = fmax(mx, red[w]);
        a.inlier_t[b] = 2.0 * sqrt(mx) / 10.0;           // SourceDiameter / 10 (umeyama.py:19-20)
    }
    if (a.rule == 1) {                                   // trans.py:15-21: ratio of the mean point norms
        const float* t = a.dst + (size_t)b * a.cap * 3;
        double ns = 0.0, nt = 0.0;
        for (int i = tid; i < n; i += S3_T) {
            const size_t g = 3 * (size_t)i;
            ns = ns + sqrt((double)s[g] * (double)s[g] + (double)s[g + 1] * (double)s[g + 1] + (double)s[g + 2] * (double)s[g + 2]);
            nt = nt + sqrt((double)t[g] * (double)t[g] + (double)t[g + 1] * (double)t[g + 1] + (double)t[g + 2] * (double)t[g + 2]);
        }
        const double sn = block_sum_d(ns, red) / (double)max(n, 1), tn = block_sum_d(nt, red) / (double)max(n, 1);
        if (tid == 0) {
            const double rts = tn / sn, rst = sn / tn;
            a.pass_t[b] = rst > rts ? rst : rts;
        }
    }
}

__global__ void __launch_bounds__(128) sim3_hypothesis_kernel(Sim3Args a) {
    const long long gid = (long long)blockIdx.x * 128 + threadIdx.x;
    if (gid >= (long long)a.B * a.H) return;
    const int b = (int)(gid / a.H), h = (int)(gid - (long long)b * a.H);
    const int n = min(a.count[b], a.cap);
    float* o = a.hyp + (size_t)gid * 16;
    bool valid = n >= 1;
    int idx[5] = {0, 0, 0, 0, 0};
    if (valid) {
        if (a.sample_idx) {
#pragma unroll
            for (int k = 0; k < 5; ++k) { idx[k] = a.sample_idx[(size_t)gid * 5 + k]; valid = valid && idx[k] >= 0 && idx[k] < n; }
        } else {
            uint32_t r0[4], r1[4];
            philox4x32_10((uint32_t)h, a.prob0 + (uint32_t)b, 0u, S3_TAG, a.seed_lo, a.seed_hi, r0);
            philox4x32_10((uint32_t)h, a.prob0 + (uint32_t)b, 1u, S3_TAG, a.seed_lo, a.seed_hi, r1);
            idx[0] = (int)__umulhi(r0[0], (uint32_t)n); idx[1] = (int)__umulhi(r0[1], (uint32_t)n);
            idx[2] = (int)__umulhi(r0[2], (uint32_t)n); idx[3] = (int)__umulhi(r0[3], (uint32_t)n);
            idx[4] = (int)__umulhi(r1[0], (uint32_t)n);
        }
    }
    double scale = 0.0, R[9], T[3];
    if (valid) {
        const float* s = a.src + (size_t)b * a.cap * 3;
        const float* t = a.dst + (size_t)b * a.cap * 3;
        double ps[5][3], pt[5][3], ms[3] = {0.0, 0.0, 0.0}, mt[3] = {0.0, 0.0, 0.0};
#pragma unroll
        for (int k = 0; k < 5; ++k)
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                ps[k][c] = (double)s[3 * (size_t)idx[k] + c];
                pt[k][c] = (double)t[3 * (size_t)idx[k] + c];
                ms[c] = ms[c] + ps[k][c];
                mt[c] = mt[c] + pt[k][c];
            }
#pragma unroll
        for (int c = 0; c < 3; ++c) { ms[c] = ms[c] / 5.0; mt[c] = mt[c] / 5.0; }
        double C[3][3], varp = 0.0;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) C[i][j] = 0.0;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            double ds[3], dt[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) { ds[c] = ps[k][c] - ms[c]; dt[c] = pt[k][c] - mt[c]; }
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) C[i][j] = C[i][j] + dt[i] * ds[j];
            varp = varp + (ds[0] * ds[0] + ds[1] * ds[1] + ds[2] * ds[2]);
        }
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) C[i][j] = C[i][j] / 5.0;
        varp = varp / 5.0;
        valid = umeyama_core(ms, mt, C, varp, scale, R, T);
    }
    if (a.rule == 1) {                                     // trans.py:104,114: OutTransform holds s * (U Vh)^T
        double* od = a.hypd + (size_t)gid * 16;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) od[3 * i + j] = valid ? scale * R[3 * j + i] : 0.0;
#pragma unroll
        for (int k = 0; k < 3; ++k) od[9 + k] = valid ? T[k] : 0.0;
        od[12] = valid ? 1.0 : 0.0;
    }
    if (valid) {
#pragma unroll
        for (int k = 0; k < 9; ++k) o[k] = (float)(scale * R[k]);
        o[9] = (float)T[0]; o[10] = (float)T[1]; o[11] = (float)T[2];
        const double thr = scale * a.inlier_t[b];          // PassThreshold = Scale * InlierT (umeyama.py:34)
        o[12] = thr > 0.0 ? (float)(thr * thr) : 0.0f;
        o[13] = 1.0f;
    } else {
#pragma unroll
        for (int k = 0; k < 13; ++k) o[k] = 0.0f;          // thr^2 = 0: nothing is an inlier
        o[13] = 0.0f;
    }
}

__global__ void __launch_bounds__(S3_ST) sim3_score_kernel(Sim3Args a) {
    __shared__ float4 sS[S3_CH];       // sx, sy, sz, tx
    __shared__ float2 sT[S3_CH];       // ty, tz
    const int nslice = (a.H + S3_ST - 1) / S3_ST;
    const int b = blockIdx.x / nslice, h = (blockIdx.x % nslice) * S3_ST + threadIdx.x;
    const int n = min(a.count[b], a.cap);
    const float* s = a.src + (size_t)b * a.cap * 3;
    const float* t = a.dst + (size_t)b * a.cap * 3;
    float A[9], T[3], thr2 = 0.0f;
#pragma unroll
    for (int k = 0; k < 9; ++k) A[k] = 0.0f;
    T[0] = T[1] = T[2] = 0.0f;
    if (h < a.H) {
        const float4* o = reinterpret_cast<const float4*>(a.hyp + ((size_t)b * a.H + h) * 16);
        const float4 r0 = o[0], r1 = o[1], r2 = o[2], r3 = o[3];
        A[0] = r0.x; A[1] = r0.y; A[2] = r0.z; A[3] = r0.w; A[4] = r1.x; A[5] = r1.y; A[6] = r1.z; A[7] = r1.w;
        A[8] = r2.x; T[0] = r2.y; T[1] = r2.z; T[2] = r2.w; thr2 = r3.x;
    }
    int c = 0;
    for (int c0 = 0; c0 < n; c0 += S3_CH) {
        const int m = min(S3_CH, n - c0);
        __syncthreads();
        for (int i = threadIdx.x; i < m; i += S3_ST) {
            const size_t g = 3 * (size_t)(c0 + i);
            sS[i] = make_float4(s[g], s[g + 1], s[g + 2], t[g]);
            sT[i] = make_float2(t[g + 1], t[g + 2]);
        }
        __syncthreads();
#pragma unroll 4
        for (int i = 0; i < m; ++i) {
            const float4 p = sS[i];
            const float2 q = sT[i];
            c += sim3_inlier(A, T, thr2, p.x, p.y, p.z, p.w, q.x, q.y) ? 1 : 0;
        }
    }
    if (h < a.H) a.hyp_count[(size_t)b * a.H + h] = c;
}

// trans.py's evaluateModel (:83-90) in FP64: residual of every pair under  OutTransform = [s R^T | T]
__device__ __forceinline__ double trans_res2(const double A[9], const double T[3], float sx, float sy, float sz,
                                             float tx, float ty, float tz) {
    const double dx = (double)tx - (A[0] * sx + A[1] * sy + A[2] * sz + T[0]);
    const double dy = (double)ty - (A[3] * sx + A[4] * sy + A[5] * sz + T[1]);
    const double dz = (double)tz - (A[6] * sx + A[7] * sy + A[8] * sz + T[2]);
    return dx * dx + dy * dy + dz * dz;
}

__global__ void __launch_bounds__(S3_ST) sim3_trans_score_kernel(Sim3Args a) {
    __shared__ float4 sS[S3_CH];       // sx, sy, sz, tx
    __shared__ float2 sT[S3_CH];       // ty, tz
    const int nslice = (a.H + S3_ST - 1) / S3_ST;
    const int b = blockIdx.x / nslice, h = (blockIdx.x % nslice) * S3_ST + threadIdx.x;
    const int n = min(a.count[b], a.cap);
    const float* s = a.src + (size_t)b * a.cap * 3;
    const float* t = a.dst + (size_t)b * a.cap * 3;
    double A[9], T[3];
    bool valid = false;
#pragma unroll
    for (int k = 0; k < 9; ++k) A[k] = 0.0;
    T[0] = T[1] = T[2] = 0.0;
    if (h < a.H) {
        const double* o = a.hypd + ((size_t)b * a.H + h) * 16;
#pragma unroll
        for (int k = 0; k < 9; ++k) A[k] = o[k];
        T[0] = o[9]; T[1] = o[10]; T[2] = o[11]; valid = o[12] != 0.0;
    }
    const double pass = a.pass_t[b];
    double res2 = 0.0;
    int cnz = 0, call = 0;             // inliers with a non-zero index (the reference's nInliers) / all inliers
    for (int c0 = 0; c0 < n; c0 += S3_CH) {
        const int m = min(S3_CH, n - c0);
        __syncthreads();
        for (int i = threadIdx.x; i < m; i += S3_ST) {
            const size_t g = 3 * (size_t)(c0 + i);
            sS[i] = make_float4(s[g], s[g + 1], s[g + 2], t[g]);
            sT[i] = make_float2(t[g + 1], t[g + 2]);
        }
        __syncthreads();
        for (int i = 0; i < m; ++i) {
            const float4 p = sS[i];
            const float2 q = sT[i];
            const double r2 = trans_res2(A, T, p.x, p.y, p.z, p.w, q.x, q.y);
            res2 = res2 + r2;
            const bool in = sqrt(r2) < pass;
            call += in ? 1 : 0;
            cnz += (in && (c0 + i) != 0) ? 1 : 0;
        }
    }
    if (h < a.H) {
        a.residual[(size_t)b * a.H + h] = valid ? sqrt(res2) : INFINITY;
        a.hyp_count[(size_t)b * a.H + h] = valid ? cnz : 0;
        a.hyp_inl[(size_t)b * a.H + h] = valid ? call : 0;
    }
}

__global__ void __launch_bounds__(S3_T) sim3_finalize_kernel(Sim3Args a) {
    __shared__ double red[S3_T / 32];
    __shared__ int s_best;
    const int b = blockIdx.x, tid = threadIdx.x;
    const int n = min(a.count[b], a.cap);
    const float* s = a.src + (size_t)b * a.cap * 3;
    const float* t = a.dst + (size_t)b * a.cap * 3;
    const int nw = (a.cap + 31) / 32;
    if (tid == 0 && a.fit_all) s_best = n >= 1 ? 0 : -1;
    if (tid == 0 && !a.fit_all && a.rule == 1) {
        // trans.py:47-72, iteration by iteration: strictly smaller total residual wins, stop below PassT / 100
        int best = -1;
        double best_res = 1e10, best_ratio = 0.0;
        const double stop_t = a.pass_t[b] / 100.0;
        for (int i = 0; i < a.H && n > 0; ++i) {
            const double r = a.residual[(size_t)b * a.H + i];
            if (r < best_res) { best_res = r; best = i; best_ratio = (double)a.hyp_count[(size_t)b * a.H + i] / (double)n; }
            if (best_res < stop_t) break;
        }
        if (best >= 0 && best_ratio < (double)a.min_ratio) best = -2 - best;
        s_best = best;
    }
    if (tid == 0 && !a.fit_all && a.rule != 1) {
        // the reference's loop, iteration by iteration (umeyama.py:31-49)
        int best = -1;
        double best_ratio = 0.0, base = 1.0, pw = 1.0;             // pw = base^i, base = 1 - best_ratio^5
        for (int i = 0; i < a.H && n > 0; ++i) {
            const double ratio = (double)a.hyp_count[(size_t)b * a.H + i] / (double)n;
            if (ratio > best_ratio) {
                best_ratio = ratio; best = i;
                base = 1.0 - best_ratio * best_ratio * best_ratio * best_ratio * best_ratio;
                pw = 1.0;
                for (int k = 0; k < i; ++k) pw = pw * base;
            } else if (i > 0) {
                pw = pw * base;
            }
            if (1.0 - pw > (double)a.confidence) break;            // (1 - (1 - ratio^5)^i) > confidence
        }
        if (best >= 0 && best_ratio < (double)a.min_ratio) best = -2 - best;   // keep the index, mark the failure
        s_best = best;
    }
    __syncthreads();
    const int best = s_best;
    const int bh = best >= 0 ? best : (best <= -2 ? -2 - best : -1);
    int status = n < 1 ? PB_STATUS_TOO_FEW_POINTS : (best >= 0 ? PB_STATUS_OK : PB_STATUS_NO_MODEL);
    float A[9], T[3], thr2 = 0.0f;
    double Ad[9], Td[3], pass = 0.0;
#pragma unroll
    for (int k = 0; k < 9; ++k) { A[k] = 0.0f; Ad[k] = 0.0; }
    T[0] = T[1] = T[2] = 0.0f; Td[0] = Td[1] = Td[2] = 0.0;
    const bool dmode = a.rule == 1 && !a.fit_all;
    if (a.fit_all) {
        thr2 = INFINITY;                       // A = 0, T = 0: |t|^2 < inf for every finite pair
    } else if (bh >= 0) {
        const float* o = a.hyp + ((size_t)b * a.H + bh) * 16;
#pragma unroll
        for (int k = 0; k < 9; ++k) A[k] = o[k];
        T[0] = o[9]; T[1] = o[10]; T[2] = o[11]; thr2 = o[12];
        if (dmode) {
            const double* od = a.hypd + ((size_t)b * a.H + bh) * 16;
#pragma unroll
            for (int k = 0; k < 9; ++k) Ad[k] = od[k];
            Td[0] = od[9]; Td[1] = od[10]; Td[2] = od[11]; pass = a.pass_t[b];
        }
    }
    auto is_inlier = [&](size_t g) -> bool {
        if (dmode) return sqrt(trans_res2(Ad, Td, s[g], s[g + 1], s[g + 2], t[g], t[g + 1], t[g + 2])) < pass;
        return sim3_inlier(A, T, thr2, s[g], s[g + 1], s[g + 2], t[g], t[g + 1], t[g + 2]);
    };
    // pass 1: inlier bitmap, count, sums
    double ps[3] = {0.0, 0.0, 0.0}, pt[3] = {0.0, 0.0, 0.0}, pc = 0.0;
    for (int i0 = 0; i0 < nw * 32; i0 += S3_T) {
        const int i = i0 + tid;
        bool in = false;
        if (i < n && bh >= 0) {
            const size_t g = 3 * (size_t)i;
            in = is_inlier(g);
            if (in) {
#pragma unroll
                for (int c = 0; c < 3; ++c) { ps[c] = ps[c] + (double)s[g + c]; pt[c] = pt[c] + (double)t[g + c]; }
                pc = pc + 1.0;
            }
        }
        const uint32_t word = __ballot_sync(0xffffffffu, in);
        if (a.bitmap && (tid & 31) == 0 && (i >> 5) < nw) a.bitmap[(size_t)b * nw + (i >> 5)] = word;
    }
    const double cnt = block_sum_d(pc, red);
    double ms[3], mt[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) { ms[c] = block_sum_d(ps[c], red) / fmax(cnt, 1.0); mt[c] = block_sum_d(pt[c], red) / fmax(cnt, 1.0); }
    // pass 2: centred moments of the inliers
    double pC[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, pv = 0.0;
    for (int i = tid; i < n && bh >= 0; i += S3_T) {
        const size_t g = 3 * (size_t)i;
        if (is_inlier(g)) {
            double ds[3], dt[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) { ds[c] = (double)s[g + c] - ms[c]; dt[c] = (double)t[g + c] - mt[c]; }
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int c = 0; c < 3; ++c) pC[3 * r + c] = pC[3 * r + c] + dt[r] * ds[c];
            pv = pv + (ds[0] * ds[0] + ds[1] * ds[1] + ds[2] * ds[2]);
        }
    }
    double C[3][3];
#pragma unroll
    for (int k = 0; k < 9; ++k) C[k / 3][k % 3] = block_sum_d(pC[k], red) / fmax(cnt, 1.0);
    const double varp = block_sum_d(pv, red) / fmax(cnt, 1.0);
    if (tid == 0) {
        double scale = 0.0, R[9] = {1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0}, Tt[3] = {0.0, 0.0, 0.0};
        if (status == PB_STATUS_OK) {
            double Rr[9], Tr[3], sc;
            if (umeyama_core(ms, mt, C, varp, sc, Rr, Tr)) {
                scale = sc;
#pragma unroll
                for (int k = 0; k < 9; ++k) R[k] = Rr[k];
                Tt[0] = Tr[0]; Tt[1] = Tr[1]; Tt[2] = Tr[2];
            } else {
                status = PB_STATUS_NO_MODEL;
            }
        }
        a.scale[b] = (float)scale;
        if (dmode) {                               // trans.py:104 returns (U Vh)^T
            const double Rt[9] = {R[0], R[3], R[6], R[1], R[4], R[7], R[2], R[5], R[8]};
#pragma unroll
            for (int k = 0; k < 9; ++k) R[k] = Rt[k];
        }
#pragma unroll
        for (int k = 0; k < 9; ++k) a.R[9 * (size_t)b + k] = (float)R[k];
#pragma unroll
        for (int k = 0; k < 3; ++k) a.t[3 * (size_t)b + k] = (float)Tt[k];
        if (a.n_inl) a.n_inl[b] = (int)cnt;
        if (a.best_hyp) a.best_hyp[b] = bh;
        a.status[b] = status;
    }
}

// ---- batched rigid (Kabsch) fit: leastsquares.py:10-48 ------------------------------------------------
// One CTA per problem: B ~ R A + t with R the proper rotation closest to the cross-covariance (what
// Vt^T U^T with the reflection fix of :38-43 is) -- Horn's eigen-solve of umeyama_core, scale dropped.
// dst_idx != NULL: the i-th target point is dst[dst_idx[b,i]] (the KNN correspondences of the ICP loop).
// R_io / t_io != NULL: the fitted (opt_R, opt_t) is composed into the running pose
//   compose 0 (train_declarative.py:98-99 as written):  R <- R opt_R,      t <- t + opt_t
//   compose 1 (rigid-motion composition):               R <- opt_R R,      t <- opt_R t + opt_t
struct KabschArgs {
    const float* src; const float* dst; const int64_t* dst_idx; const int32_t* count;
    int cap, dst_cap, dst_shared, compose;
    float* R; float* t; float* cA; float* cB; int32_t* status; float* R_io; float* t_io;
};

__global__ void __launch_bounds__(S3_T) kabsch_kernel(KabschArgs a) {
    __shared__ double red[S3_T / 32];
    const int b = blockIdx.x, tid = threadIdx.x;
    const int n = a.count ? max(0, min(a.count[b], a.cap)) : a.cap;
    const float* s = a.src + (size_t)b * a.cap * 3;
    const float* d = a.dst + (a.dst_shared ? 0 : (size_t)b * a.dst_cap * 3);
    const int64_t* ix = a.dst_idx ? a.dst_idx + (size_t)b * a.cap : nullptr;
    auto dpt = [&](int i, int c) -> double {
        long long j = ix ? (long long)ix[i] : i;
        j = j < 0 ? 0 : (j >= a.dst_cap ? a.dst_cap - 1 : j);
        return (double)d[3 * (size_t)j + c];
    };
    double ps[3] = {0.0, 0.0, 0.0}, pt[3] = {0.0, 0.0, 0.0};
    for (int i = tid; i < n; i += S3_T)
#pragma unroll
        for (int c = 0; c < 3; ++c) { ps[c] = ps[c] + (double)s[3 * (size_t)i + c]; pt[c] = pt[c] + dpt(i, c); }
    double ms[3], mt[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) { ms[c] = block_sum_d(ps[c], red) / (double)max(n, 1); mt[c] = block_sum_d(pt[c], red) / (double)max(n, 1); }
    double pC[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, pv = 0.0;
    for (int i = tid; i < n; i += S3_T) {
        double ds[3], dt[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) { ds[c] = (double)s[3 * (size_t)i + c] - ms[c]; dt[c] = dpt(i, c) - mt[c]; }
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c) pC[3 * r + c] = pC[3 * r + c] + dt[r] * ds[c];
        pv = pv + (ds[0] * ds[0] + ds[1] * ds[1] + ds[2] * ds[2]);
    }
    double C[3][3];
#pragma unroll
    for (int k = 0; k < 9; ++k) C[k / 3][k % 3] = block_sum_d(pC[k], red) / (double)max(n, 1);
    const double varp = block_sum_d(pv, red) / (double)max(n, 1);
    if (tid == 0) {
        double R[9] = {1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0}, T[3], sc;
        int st = n < 1 ? PB_STATUS_TOO_FEW_POINTS : PB_STATUS_OK;
        if (st == PB_STATUS_OK) {
            double Rr[9], Tr[3];
            if (umeyama_core(ms, mt, C, varp, sc, Rr, Tr)) {
#pragma unroll
                for (int k = 0; k < 9; ++k) R[k] = Rr[k];
            } else {
                st = PB_STATUS_NO_MODEL;                    // zero spread: identity rotation, centroid shift
            }
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) T[i] = mt[i] - (R[3 * i] * ms[0] + R[3 * i + 1] * ms[1] + R[3 * i + 2] * ms[2]);   // :45
        if (n < 1) T[0] = T[1] = T[2] = 0.0;
        if (a.R)
#pragma unroll
            for (int k = 0; k < 9; ++k) a.R[9 * (size_t)b + k] = (float)R[k];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (a.t) a.t[3 * (size_t)b + k] = (float)T[k];
            if (a.cA) a.cA[3 * (size_t)b + k] = (float)ms[k];
            if (a.cB) a.cB[3 * (size_t)b + k] = (float)mt[k];
        }
        if (a.status) a.status[b] = st;
        if (a.R_io) {
            float* Rp = a.R_io + 9 * (size_t)b;
            float* tp = a.t_io + 3 * (size_t)b;
            double Ro[9], to[3];
#pragma unroll
            for (int k = 0; k < 9; ++k) Ro[k] = (double)Rp[k];
#pragma unroll
            for (int k = 0; k < 3; ++k) to[k] = (double)tp[k];
            double Rn[9], tn[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
#pragma unroll
                for (int j = 0; j < 3; ++j)
                    Rn[3 * i + j] = a.compose ? R[3 * i] * Ro[j] + R[3 * i + 1] * Ro[3 + j] + R[3 * i + 2] * Ro[6 + j]
                                              : Ro[3 * i] * R[j] + Ro[3 * i + 1] * R[3 + j] + Ro[3 * i + 2] * R[6 + j];
                tn[i] = a.compose ? R[3 * i] * to[0] + R[3 * i + 1] * to[1] + R[3 * i + 2] * to[2] + T[i] : to[i] + T[i];
            }
#pragma unroll
            for (int k = 0; k < 9; ++k) Rp[k] = (float)Rn[k];
#pragma unroll
            for (int k = 0; k < 3; ++k) tp[k] = (float)tn[k];
        }
    }
}

size_t sim3_ws_bytes(int B, int H, int rule) {
    size_t n = (size_t)B * sizeof(double) + (size_t)B * H * 16 * sizeof(float) + (size_t)B * H * sizeof(int32_t) + 256;
    if (rule == 1) n += (size_t)B * H * (16 + 1) * sizeof(double) + (size_t)B * sizeof(double) + (size_t)B * H * sizeof(int32_t) + 256;
    return n;
}

}  // namespace

extern "C" size_t pb_sim3_ransac_workspace_bytes(int B, const pb_sim3_params* prm) {
    if (B <= 0 || !prm || prm->n_iter <= 0) return 0;
    return sim3_ws_bytes(B, prm->n_iter, prm->rule);
}

extern "C" int pb_sim3_ransac(const float* src, const float* dst, const int32_t* count, int B, int cap, uint32_t prob0,
                              const pb_sim3_params* prm, const int32_t* sample_idx, float* scale, float* R, float* t,
                              int32_t* n_inl, int32_t* best_hyp, int32_t* status, int32_t* hyp_counts, uint32_t* inl_bitmap,
                              void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(B >= 0 && cap > 0 && prm, "pb_sim3_ransac: bad arguments");
    if (B == 0) return 0;
    const int H = prm->n_iter;
    PB_REQUIRE(H > 0 && H <= 4096, "pb_sim3_ransac: n_iter out of range (1..4096)");
    PB_REQUIRE(src && dst && count && scale && R && t && n_inl && best_hyp && status, "pb_sim3_ransac: null pointer argument");
    PB_REQUIRE(prm->rule == 0 || prm->rule == 1, "pb_sim3_ransac: rule must be 0 (umeyama.py) or 1 (trans.py)");
    PB_REQUIRE(ws && ws_bytes >= sim3_ws_bytes(B, H, prm->rule), "pb_sim3_ransac: workspace too small");
    PB_REQUIRE(reinterpret_cast<uintptr_t>(ws) % 16 == 0, "pb_sim3_ransac: workspace must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* base = static_cast<char*>(ws);
    Sim3Args a{};
    a.src = src; a.dst = dst; a.count = count; a.B = B; a.cap = cap; a.H = H;
    a.seed_lo = prm->seed_lo; a.seed_hi = prm->seed_hi; a.prob0 = prob0;
    a.confidence = prm->confidence; a.min_ratio = prm->min_inlier_ratio;
    a.sample_idx = sample_idx;
    a.hyp = reinterpret_cast<float*>(base);
    a.inlier_t = reinterpret_cast<double*>(base + (size_t)B * H * 16 * sizeof(float));
    int32_t* own_counts = reinterpret_cast<int32_t*>(base + (size_t)B * H * 16 * sizeof(float) + (size_t)B * sizeof(double));
    a.hyp_count = hyp_counts ? hyp_counts : own_counts;
    a.scale = scale; a.R = R; a.t = t; a.n_inl = n_inl; a.best_hyp = best_hyp; a.status = status; a.bitmap = inl_bitmap;
    a.rule = prm->rule;
    if (a.rule == 1) {
        size_t off = ((size_t)B * H * 16 * sizeof(float) + (size_t)B * sizeof(double) + (size_t)B * H * sizeof(int32_t) + 255) / 256 * 256;
        a.hypd = reinterpret_cast<double*>(base + off); off += (size_t)B * H * 16 * sizeof(double);
        a.residual = reinterpret_cast<double*>(base + off); off += (size_t)B * H * sizeof(double);
        a.pass_t = reinterpret_cast<double*>(base + off); off += (size_t)B * sizeof(double);
        a.hyp_inl = reinterpret_cast<int32_t*>(base + off);
    }
    sim3_prepare_kernel<<<B, S3_T, 0, st>>>(a);
    if (pb_check_launch("sim3_prepare_kernel")) return -2;
    const long long nh = (long long)B * H;
    sim3_hypothesis_kernel<<<(unsigned)((nh + 127) / 128), 128, 0, st>>>(a);
    if (pb_check_launch("sim3_hypothesis_kernel")) return -2;
    if (a.rule == 1) sim3_trans_score_kernel<<<B * ((H + S3_ST - 1) / S3_ST), S3_ST, 0, st>>>(a);
    else sim3_score_kernel<<<B * ((H + S3_ST - 1) / S3_ST), S3_ST, 0, st>>>(a);
    if (pb_check_launch("sim3_score_kernel")) return -2;
    sim3_finalize_kernel<<<B, S3_T, 0, st>>>(a);
    return pb_check_launch("sim3_finalize_kernel") ? -2 : 0;
}

// batched estimateSimilarityUmeyama (lib/transform/umeyama.py:65-97): every pair counts
extern "C" int pb_umeyama_fit(const float* src, const float* dst, const int32_t* count, int B, int cap, float* scale,
                              float* R, float* t, int32_t* status, void* stream) {
    PB_REQUIRE(B >= 0 && cap > 0, "pb_umeyama_fit: bad arguments");
    if (B == 0) return 0;
    PB_REQUIRE(src && dst && count && scale && R && t && status, "pb_umeyama_fit: null pointer argument");
    Sim3Args a{};
    a.src = src; a.dst = dst; a.count = count; a.B = B; a.cap = cap; a.H = 1; a.fit_all = 1;
    a.scale = scale; a.R = R; a.t = t; a.status = status;
    sim3_finalize_kernel<<<B, S3_T, 0, static_cast<cudaStream_t>(stream)>>>(a);
    return pb_check_launch("sim3_finalize_kernel") ? -2 : 0;
}

// batched rigid fit (version/transparent/lib/declarative/leastsquares.py:10-48, best_batch_fit_transform)
extern "C" int pb_kabsch_fit(const float* A, const float* Bm, const int32_t* count, int B, int cap, float* R, float* t,
                             float* centroid_A, float* centroid_B, int32_t* status, void* stream) {
    PB_REQUIRE(B >= 0 && cap > 0, "pb_kabsch_fit: bad arguments");
    if (B == 0) return 0;
    PB_REQUIRE(A && Bm && R && t, "pb_kabsch_fit: null pointer argument");
    KabschArgs a{};
    a.src = A; a.dst = Bm; a.count = count; a.cap = cap; a.dst_cap = cap;
    a.R = R; a.t = t; a.cA = centroid_A; a.cB = centroid_B; a.status = status;
    kabsch_kernel<<<B, S3_T, 0, static_cast<cudaStream_t>(stream)>>>(a);
    return pb_check_launch("kabsch_kernel") ? -2 : 0;
}

// the ICP-style loop of version/transparent/scripts/train_declarative.py:80-100
extern "C" size_t pb_icp_refine_workspace_bytes(int B, int N, int M) {
    if (B <= 0 || N <= 0 || M <= 0) return 0;
    return 256 + ((size_t)B * N * 3 * sizeof(float) + 255) / 256 * 256 + ((size_t)B * N * sizeof(int64_t) + 255) / 256 * 256 +
           pb_knn1_tc_workspace_bytes(B, N, M) + 256;
}

extern "C" int pb_icp_refine(const float* model, int model_shared, const float* target, int target_shared, float* R,
                             float* t, int B, int N, int M, int n_steps, int compose, int use_tc, int32_t* status,
                             void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(B >= 0 && N > 0 && M > 0 && n_steps >= 0, "pb_icp_refine: bad sizes");
    PB_REQUIRE(compose == 0 || compose == 1, "pb_icp_refine: compose must be 0 (reference) or 1 (rigid)");
    if (B == 0 || n_steps == 0) return 0;
    PB_REQUIRE(model && target && R && t, "pb_icp_refine: null pointer argument");
    PB_REQUIRE(ws && ws_bytes >= pb_icp_refine_workspace_bytes(B, N, M), "pb_icp_refine: workspace too small");
    PB_REQUIRE(reinterpret_cast<uintptr_t>(ws) % 128 == 0, "pb_icp_refine: workspace must be 128-byte aligned");
    char* base = static_cast<char*>(ws);
    float* pred = reinterpret_cast<float*>(base);
    size_t off = ((size_t)B * N * 3 * sizeof(float) + 255) / 256 * 256;
    int64_t* idx = reinterpret_cast<int64_t*>(base + off);
    off += ((size_t)B * N * sizeof(int64_t) + 255) / 256 * 256;
    off = (off + 127) / 128 * 128;
    void* kws = base + off;
    const size_t kws_bytes = ws_bytes - off;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    for (int step = 0; step < n_steps; ++step) {
        int rc = pb_transform_points(model, model_shared, R, t, B, N, pred, stream);                 // :81
        if (rc) return rc;
        if (use_tc) rc = pb_internal_knn1_tc("pb_icp_refine", pred, target, B, N, M, target_shared, idx, nullptr, kws, kws_bytes, nullptr, 0, st);
        else rc = pb_knn1(pred, target, B, N, M, target_shared, idx, nullptr, stream);               // :82
        if (rc) return rc;
        KabschArgs a{};                                                                               // :84-86, :98-99
        a.src = pred; a.dst = target; a.dst_idx = idx; a.cap = N; a.dst_cap = M; a.dst_shared = target_shared;
        a.compose = compose; a.status = status; a.R_io = R; a.t_io = t;
        kabsch_kernel<<<B, S3_T, 0, st>>>(a);
        if (pb_check_launch("kabsch_kernel")) return -2;
    }
    return 0;
}
