/* oracle/sim3_ref.c -- CPU oracle of the 3D-3D RANSAC + Umeyama path.
 *
 * TEST INFRASTRUCTURE ONLY (same rules as pose_ref.c).  Plain-C, bit-defined restatement of what
 * pose_estimation_b200/csrc/sim3.cu computes, which in turn follows the reference's
 * lib/transform/umeyama.py:8-62 (estimateSimilarityTransform) and :65-97 (estimateSimilarityUmeyama).
 * Parity status: PINNED against the reference itself -- tests/golden/sim3_ref.npz holds inputs, the index
 * sets np.random produced, per-iteration inlier counts and final transforms obtained by importing
 * /root/reference/lib/transform/umeyama.py (tests/golden/gen_golden_sim3.py); tests/test_oracle_golden.py
 * checks this file against them.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "pose_ref.h"

#define S3_T 256
#define S3_TAG 0x53494D33u

/* per-thread strided partials -> xor butterfly per warp -> the 8 warp sums in order (sim3.cu block_sum_d) */
static double tree_sum_d(const double part[S3_T]) {
    double tot = 0.0;
    for (int w = 0; w < S3_T / 32; ++w) {
        double v[32], nv[32];
        memcpy(v, part + 32 * w, sizeof(v));
        for (int off = 16; off >= 1; off >>= 1) {
            for (int l = 0; l < 32; ++l) nv[l] = v[l] + v[l ^ off];
            memcpy(v, nv, sizeof(v));
        }
        tot = (w == 0) ? v[0] : tot + v[0];
    }
    return tot;
}

/* Umeyama from moments; lib/transform/umeyama.py:72-95 with the SVD replaced by Horn's quaternion matrix */
int ref_umeyama_core(const double ms[3], const double mt[3], const double C[3][3], double varp, double* scale,
                     double R[9], double T[3]) {
    const double Sxx = C[0][0], Sxy = C[1][0], Sxz = C[2][0];
    const double Syx = C[0][1], Syy = C[1][1], Syz = C[2][1];
    const double Szx = C[0][2], Szy = C[1][2], Szz = C[2][2];
    double a[4][4], v[4][4];
    a[0][0] = Sxx + Syy + Szz; a[0][1] = Syz - Szy; a[0][2] = Szx - Sxz; a[0][3] = Sxy - Syx;
    a[1][1] = Sxx - Syy - Szz; a[1][2] = Sxy + Syx; a[1][3] = Szx + Sxz;
    a[2][2] = -Sxx + Syy - Szz; a[2][3] = Syz + Szy;
    a[3][3] = -Sxx - Syy + Szz;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            if (j < i) a[i][j] = a[j][i];
            v[i][j] = (i == j) ? 1.0 : 0.0;
        }
    for (int sweep = 0; sweep < 10; ++sweep)
        for (int p = 0; p < 3; ++p)
            for (int q = p + 1; q < 4; ++q) {
                const double apq = a[p][q];
                if (apq != 0.0) {
                    const double theta = (a[q][q] - a[p][p]) / (2.0 * apq);
                    const double tt = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                    const double c = 1.0 / sqrt(tt * tt + 1.0), s = tt * c;
                    for (int k = 0; k < 4; ++k) {
                        const double akp = a[k][p], akq = a[k][q];
                        a[k][p] = c * akp - s * akq;
                        a[k][q] = s * akp + c * akq;
                    }
                    for (int k = 0; k < 4; ++k) {
                        const double apk = a[p][k], aqk = a[q][k];
                        a[p][k] = c * apk - s * aqk;
                        a[q][k] = s * apk + c * aqk;
                    }
                    for (int k = 0; k < 4; ++k) {
                        const double vkp = v[k][p], vkq = v[k][q];
                        v[k][p] = c * vkp - s * vkq;
                        v[k][q] = s * vkp + c * vkq;
                    }
                }
            }
    int best = 0;
    double lam = a[0][0];
    for (int i = 1; i < 4; ++i)
        if (a[i][i] > lam) { lam = a[i][i]; best = i; }
    double q[4];
    for (int k = 0; k < 4; ++k) q[k] = v[k][best];
    const double nn = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    if (!(nn > 0.0) || !(varp > 0.0) || !isfinite(lam) || !isfinite(varp)) return 0;
    const double sg = q[0] < 0.0 ? -1.0 : 1.0;
    const double w = sg * q[0] / nn, x = sg * q[1] / nn, y = sg * q[2] / nn, z = sg * q[3] / nn;
    R[0] = 1.0 - 2.0 * (y * y + z * z); R[1] = 2.0 * (x * y - w * z);       R[2] = 2.0 * (x * z + w * y);
    R[3] = 2.0 * (x * y + w * z);       R[4] = 1.0 - 2.0 * (x * x + z * z); R[5] = 2.0 * (y * z - w * x);
    R[6] = 2.0 * (x * z - w * y);       R[7] = 2.0 * (y * z + w * x);       R[8] = 1.0 - 2.0 * (x * x + y * y);
    *scale = lam / varp;
    for (int i = 0; i < 3; ++i)
        T[i] = mt[i] - *scale * (R[3 * i] * ms[0] + R[3 * i + 1] * ms[1] + R[3 * i + 2] * ms[2]);
    return isfinite(*scale) && isfinite(T[0]) && isfinite(T[1]) && isfinite(T[2]);
}

static int sim3_inlier(const float A[9], const float T[3], float thr2, const float* s, const float* t) {
    const float dx = t[0] - fmaf(A[0], s[0], fmaf(A[1], s[1], fmaf(A[2], s[2], T[0])));
    const float dy = t[1] - fmaf(A[3], s[0], fmaf(A[4], s[1], fmaf(A[5], s[2], T[1])));
    const float dz = t[2] - fmaf(A[6], s[0], fmaf(A[7], s[1], fmaf(A[8], s[2], T[2])));
    return fmaf(dx, dx, fmaf(dy, dy, dz * dz)) < thr2;
}

/* InlierT = SourceDiameter / 10 (umeyama.py:16-20) */
double ref_sim3_inlier_t(const float* src, int n) {
    double m[3];
    for (int c = 0; c < 3; ++c) {
        double part[S3_T];
        for (int v = 0; v < S3_T; ++v) {
            double p = 0.0;
            for (int i = v; i < n; i += S3_T) p = p + (double)src[3 * (size_t)i + c];
            part[v] = p;
        }
        m[c] = tree_sum_d(part) / (double)(n > 1 ? n : 1);
    }
    double mx = 0.0;
    for (int i = 0; i < n; ++i) {
        const double x = (double)src[3 * (size_t)i] - m[0], y = (double)src[3 * (size_t)i + 1] - m[1],
                     z = (double)src[3 * (size_t)i + 2] - m[2];
        mx = fmax(mx, x * x + y * y + z * z);
    }
    return 2.0 * sqrt(mx) / 10.0;
}

/* one hypothesis from 5 index pairs: hyp[16] = sR (9), T (3), thr^2, valid */
int ref_sim3_hypothesis(const float* src, const float* dst, int n, const int32_t idx[5], double inlier_t, float hyp[16]) {
    int valid = n >= 1;
    for (int k = 0; k < 5; ++k) valid = valid && idx[k] >= 0 && idx[k] < n;
    double scale = 0.0, R[9], T[3];
    if (valid) {
        double ps[5][3], pt[5][3], ms[3] = {0.0, 0.0, 0.0}, mt[3] = {0.0, 0.0, 0.0};
        for (int k = 0; k < 5; ++k)
            for (int c = 0; c < 3; ++c) {
                ps[k][c] = (double)src[3 * (size_t)idx[k] + c];
                pt[k][c] = (double)dst[3 * (size_t)idx[k] + c];
                ms[c] = ms[c] + ps[k][c];
                mt[c] = mt[c] + pt[k][c];
            }
        for (int c = 0; c < 3; ++c) { ms[c] = ms[c] / 5.0; mt[c] = mt[c] / 5.0; }
        double C[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}}, varp = 0.0;
        for (int k = 0; k < 5; ++k) {
            double ds[3], dt[3];
            for (int c = 0; c < 3; ++c) { ds[c] = ps[k][c] - ms[c]; dt[c] = pt[k][c] - mt[c]; }
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) C[i][j] = C[i][j] + dt[i] * ds[j];
            varp = varp + (ds[0] * ds[0] + ds[1] * ds[1] + ds[2] * ds[2]);
        }
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) C[i][j] = C[i][j] / 5.0;
        varp = varp / 5.0;
        valid = ref_umeyama_core(ms, mt, C, varp, &scale, R, T);
    }
    memset(hyp, 0, 16 * sizeof(float));
    if (valid) {
        for (int k = 0; k < 9; ++k) hyp[k] = (float)(scale * R[k]);
        hyp[9] = (float)T[0]; hyp[10] = (float)T[1]; hyp[11] = (float)T[2];
        const double thr = scale * inlier_t;
        hyp[12] = thr > 0.0 ? (float)(thr * thr) : 0.0f;
        hyp[13] = 1.0f;
    }
    return valid;
}

/* 5 indices with replacement (np.random.randint(n, size=5), umeyama.py:32) from the counter RNG */
void ref_sim3_sample5(uint32_t seed_lo, uint32_t seed_hi, uint32_t prob, uint32_t hyp, int n, int32_t idx[5]) {
    const uint32_t key[2] = {seed_lo, seed_hi};
    uint32_t c0[4] = {hyp, prob, 0u, S3_TAG}, c1[4] = {hyp, prob, 1u, S3_TAG}, r0[4], r1[4];
    ref_philox4x32(c0, key, r0);
    ref_philox4x32(c1, key, r1);
    for (int k = 0; k < 4; ++k) idx[k] = (int32_t)(((uint64_t)r0[k] * (uint32_t)n) >> 32);
    idx[4] = (int32_t)(((uint64_t)r1[0] * (uint32_t)n) >> 32);
}

/* whole problem.  sample_idx [H,5] or NULL.  Returns status (0 ok, 1 too few points, 2 no model). */
int ref_sim3_ransac(const float* src, const float* dst, int n, int H, float confidence, float min_ratio, uint32_t seed_lo,
                    uint32_t seed_hi, uint32_t prob, const int32_t* sample_idx, float* scale_out, float R_out[9],
                    float t_out[3], int32_t* n_inl, int32_t* best_hyp, int32_t* hyp_counts /*[H] or NULL*/,
                    uint32_t* bitmap /*[ceil(n/32)] or NULL*/) {
    const int fit_all = H <= 0;                       /* ref_umeyama_fit: every pair is an inlier */
    const double inlier_t = ref_sim3_inlier_t(src, n);
    int best = -1;
    float bestA[16];
    memset(bestA, 0, sizeof(bestA));
    if (fit_all) { bestA[12] = INFINITY; best = n >= 1 ? 0 : -1; }
    double best_ratio = fit_all ? 1.0 : 0.0, base = 1.0, pw = 1.0;
    int stopped = 0;
    for (int i = 0; i < H; ++i) {
        int32_t idx[5];
        if (sample_idx) memcpy(idx, sample_idx + 5 * (size_t)i, sizeof(idx));
        else ref_sim3_sample5(seed_lo, seed_hi, prob, (uint32_t)i, n, idx);
        float hyp[16];
        ref_sim3_hypothesis(src, dst, n, idx, inlier_t, hyp);
        int c = 0;
        for (int k = 0; k < n; ++k) c += sim3_inlier(hyp, hyp + 9, hyp[12], src + 3 * (size_t)k, dst + 3 * (size_t)k);
        if (hyp_counts) hyp_counts[i] = c;
        if (stopped || n <= 0) continue;              /* the GPU scores every iteration; the choice stops here */
        const double ratio = (double)c / (double)n;
        if (ratio > best_ratio) {
            best_ratio = ratio; best = i;
            memcpy(bestA, hyp, sizeof(bestA));
            base = 1.0 - best_ratio * best_ratio * best_ratio * best_ratio * best_ratio;
            pw = 1.0;
            for (int k = 0; k < i; ++k) pw = pw * base;
        } else if (i > 0) {
            pw = pw * base;
        }
        if (1.0 - pw > (double)confidence) stopped = 1;
    }
    int status = n < 1 ? 1 : ((best >= 0 && !(best_ratio < (double)min_ratio)) ? 0 : 2);
    /* inliers of the chosen iteration, sums in the kernel's order */
    const int nw = (n + 31) / 32;
    if (bitmap) memset(bitmap, 0, (size_t)nw * sizeof(uint32_t));
    double pc[S3_T], ps[3][S3_T], pt[3][S3_T];
    memset(pc, 0, sizeof(pc)); memset(ps, 0, sizeof(ps)); memset(pt, 0, sizeof(pt));
    for (int i = 0; i < n && best >= 0; ++i)
        if (sim3_inlier(bestA, bestA + 9, bestA[12], src + 3 * (size_t)i, dst + 3 * (size_t)i)) {
            const int v = i % S3_T;
            if (bitmap) bitmap[i >> 5] |= 1u << (i & 31);
            for (int c = 0; c < 3; ++c) { ps[c][v] = ps[c][v] + (double)src[3 * (size_t)i + c]; pt[c][v] = pt[c][v] + (double)dst[3 * (size_t)i + c]; }
            pc[v] = pc[v] + 1.0;
        }
    const double cnt = tree_sum_d(pc);
    double ms[3], mt[3];
    for (int c = 0; c < 3; ++c) { ms[c] = tree_sum_d(ps[c]) / fmax(cnt, 1.0); mt[c] = tree_sum_d(pt[c]) / fmax(cnt, 1.0); }
    double pC[9][S3_T], pv[S3_T];
    memset(pC, 0, sizeof(pC)); memset(pv, 0, sizeof(pv));
    for (int i = 0; i < n && best >= 0; ++i)
        if (sim3_inlier(bestA, bestA + 9, bestA[12], src + 3 * (size_t)i, dst + 3 * (size_t)i)) {
            const int v = i % S3_T;
            double ds[3], dt[3];
            for (int c = 0; c < 3; ++c) { ds[c] = (double)src[3 * (size_t)i + c] - ms[c]; dt[c] = (double)dst[3 * (size_t)i + c] - mt[c]; }
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) pC[3 * r + c][v] = pC[3 * r + c][v] + dt[r] * ds[c];
            pv[v] = pv[v] + (ds[0] * ds[0] + ds[1] * ds[1] + ds[2] * ds[2]);
        }
    double C[3][3];
    for (int k = 0; k < 9; ++k) C[k / 3][k % 3] = tree_sum_d(pC[k]) / fmax(cnt, 1.0);
    const double varp = tree_sum_d(pv) / fmax(cnt, 1.0);
    double scale = 0.0, R[9] = {1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0}, T[3] = {0.0, 0.0, 0.0};
    if (status == 0) {
        double Rr[9], Tr[3], sc;
        if (ref_umeyama_core(ms, mt, C, varp, &sc, Rr, Tr)) {
            scale = sc; memcpy(R, Rr, sizeof(R)); memcpy(T, Tr, sizeof(T));
        } else {
            status = 2;
        }
    }
    *scale_out = (float)scale;
    for (int k = 0; k < 9; ++k) R_out[k] = (float)R[k];
    for (int k = 0; k < 3; ++k) t_out[k] = (float)T[k];
    *n_inl = (int32_t)cnt;
    *best_hyp = best;
    return status;
}

/* estimateSimilarityUmeyama over all n pairs (umeyama.py:65-97) with the kernel's summation order */
int ref_umeyama_fit(const float* src, const float* dst, int n, float* scale_out, float R_out[9], float t_out[3]) {
    int32_t n_inl, best;
    return ref_sim3_ransac(src, dst, n, 0, 0.99f, 0.0f, 0u, 0u, 0u, 0, scale_out, R_out, t_out, &n_inl, &best, 0, 0);
}

/* farthest point sampling, tools/script/sample_model.py:33-63: float32 sqrt((dx*dx+dy*dy)+dz*dz), start at
 * index 0, np.argmax (first maximum) */
void ref_fps(const float* pts, int n, int n_samples, int32_t* out) {
    if (n <= 0) { for (int i = 0; i < n_samples; ++i) out[i] = -1; return; }
    float* dist = (float*)malloc((size_t)n * sizeof(float));
    for (int i = 0; i < n; ++i) dist[i] = INFINITY;
    int cur = 0;
    for (int it = 0; it < n_samples; ++it) {
        out[it] = cur;
        const float cx = pts[3 * (size_t)cur], cy = pts[3 * (size_t)cur + 1], cz = pts[3 * (size_t)cur + 2];
        int best = 0;
        float bv = -1.0f;
        for (int i = 0; i < n; ++i) {
            const float dx = pts[3 * (size_t)i] - cx, dy = pts[3 * (size_t)i + 1] - cy, dz = pts[3 * (size_t)i + 2] - cz;
            const float a = dx * dx, b2 = dy * dy, c = dz * dz;
            const float d = sqrtf((a + b2) + c);
            const float m = dist[i] < d ? dist[i] : d;
            dist[i] = m;
            if (m > bv) { bv = m; best = i; }
        }
        cur = best;
    }
    free(dist);
}
