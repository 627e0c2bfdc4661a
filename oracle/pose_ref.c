/* oracle/pose_ref.c -- see pose_ref.h.  TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC  (see oracle/Makefile)
 * Every floating-point expression below is written as individual IEEE-754 binary32 operations
 * (fmaf where a fused multiply-add is meant, never implicitly); the CUDA kernels use the same
 * sequence with __fmaf_rn / -fmad=false, so integer results derived from them are bit-identical.
 */
#include "pose_ref.h"
#include <math.h>
#include <string.h>
#include <stdlib.h>

/* ------------------------------------------------------------------ Philox4x32-10 */
void ref_philox4x32(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static inline uint32_t mulhi32(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }

#define SAMPLE_TAG 0x50335031u /* "P3P1" */
#define SAMPLE_ATTEMPTS 16

int ref_sample4(uint32_t seed_lo, uint32_t seed_hi, uint32_t frame, uint32_t hyp, int P, int32_t idx[4]) {
    if (P < 4) return 0;
    uint32_t key[2] = {seed_lo, seed_hi};
    for (uint32_t a = 0; a < SAMPLE_ATTEMPTS; ++a) {
        uint32_t ctr[4] = {hyp, frame, a, SAMPLE_TAG}, r[4];
        ref_philox4x32(ctr, key, r);
        int32_t i0 = (int32_t)mulhi32(r[0], (uint32_t)P), i1 = (int32_t)mulhi32(r[1], (uint32_t)P);
        int32_t i2 = (int32_t)mulhi32(r[2], (uint32_t)P), i3 = (int32_t)mulhi32(r[3], (uint32_t)P);
        if (i0 != i1 && i0 != i2 && i0 != i3 && i1 != i2 && i1 != i3 && i2 != i3) {
            idx[0] = i0; idx[1] = i1; idx[2] = i2; idx[3] = i3;
            return 1;
        }
    }
    return 0;
}

/* ------------------------------------------------------------------ small vector helpers */
static inline float dot3(const float a[3], const float b[3]) {
    return fmaf(a[0], b[0], fmaf(a[1], b[1], a[2] * b[2]));
}
static inline void cross3(const float a[3], const float b[3], float c[3]) {
    c[0] = fmaf(a[1], b[2], -(a[2] * b[1]));
    c[1] = fmaf(a[2], b[0], -(a[0] * b[2]));
    c[2] = fmaf(a[0], b[1], -(a[1] * b[0]));
}
static inline int finite3(const float a[3]) { return isfinite(a[0]) && isfinite(a[1]) && isfinite(a[2]); }

/* symmetric 3x3 stored as s[6] = {00,01,02,11,12,22} */
static inline void sym_adj(const float s[6], float A[6]) {
    A[0] = fmaf(s[3], s[5], -(s[4] * s[4]));
    A[1] = fmaf(s[2], s[4], -(s[1] * s[5]));
    A[2] = fmaf(s[1], s[4], -(s[2] * s[3]));
    A[3] = fmaf(s[0], s[5], -(s[2] * s[2]));
    A[4] = fmaf(s[1], s[2], -(s[0] * s[4]));
    A[5] = fmaf(s[0], s[3], -(s[1] * s[1]));
}
static inline float sym_det(const float s[6], const float A[6]) {
    return fmaf(s[0], A[0], fmaf(s[1], A[1], s[2] * A[2]));
}
static inline float sym_tr_prod(const float A[6], const float t[6]) {
    float d = fmaf(A[0], t[0], fmaf(A[3], t[3], A[5] * t[5]));
    float o = fmaf(A[1], t[1], fmaf(A[2], t[2], A[4] * t[4]));
    return fmaf(2.0f, o, d);
}
/* v^T S w */
static inline float sym_quad(const float s[6], const float v[3], const float w[3]) {
    float r0 = fmaf(s[0], w[0], fmaf(s[1], w[1], s[2] * w[2]));
    float r1 = fmaf(s[1], w[0], fmaf(s[3], w[1], s[4] * w[2]));
    float r2 = fmaf(s[2], w[0], fmaf(s[4], w[1], s[5] * w[2]));
    return fmaf(v[0], r0, fmaf(v[1], r1, v[2] * r2));
}

/* deterministic cube-root upper bound (bit hack + 2 Newton steps, inflated 2%) */
static inline float cbrt_up(float m) {
    if (!(m > 0.0f)) return 0.0f;
    union { float f; int32_t i; } u;
    u.f = m;
    u.i = u.i / 3 + 709921077;
    float y = u.f;
    y = y - (y * y * y - m) / (3.0f * y * y);
    y = y - (y * y * y - m) / (3.0f * y * y);
    return y * 1.02f;
}

static inline float cubic_eval(float a, float b, float c, float x) {
    return fmaf(fmaf(x + a, x, b), x, c);
}

/* one real root of x^3 + a x^2 + b x + c, Newton from a start with guaranteed monotone approach */
static float cubic_real_root(float a, float b, float c) {
    float disc = fmaf(a, a, -(3.0f * b));
    float x;
    if (disc > 0.0f) {
        float s = sqrtf(disc);
        float xmax = (-a - s) / 3.0f;
        float fmax = cubic_eval(a, b, c, xmax);
        if (fmax >= 0.0f) {
            x = xmax - sqrtf(fmax / s);
        } else {
            float xmin = (-a + s) / 3.0f;
            float fmin = cubic_eval(a, b, c, xmin);
            x = xmin + sqrtf(-fmin / s);
        }
    } else {
        float xi = -a / 3.0f;
        float fi = cubic_eval(a, b, c, xi);
        float d0 = cbrt_up(fabsf(fi));
        x = xi - copysignf(d0, fi);
    }
    for (int it = 0; it < 24; ++it) {
        float fx = cubic_eval(a, b, c, x);
        if (fx == 0.0f) break;
        float fp = fmaf(fmaf(3.0f, x, 2.0f * a), x, b);
        if (fp == 0.0f) break;
        float xn = x - fx / fp;
        if (xn == x) break;
        x = xn;
    }
    return x;
}

/* unit null vector of (E - sigma I), E symmetric; returns 0 on failure */
static int sym_eigvec(const float e[6], float sigma, float v[3]) {
    float r0[3] = {e[0] - sigma, e[1], e[2]};
    float r1[3] = {e[1], e[3] - sigma, e[4]};
    float r2[3] = {e[2], e[4], e[5] - sigma};
    float c01[3], c02[3], c12[3];
    cross3(r0, r1, c01); cross3(r0, r2, c02); cross3(r1, r2, c12);
    float n01 = dot3(c01, c01), n02 = dot3(c02, c02), n12 = dot3(c12, c12);
    const float* best = c01; float nb = n01;
    if (n02 > nb) { best = c02; nb = n02; }
    if (n12 > nb) { best = c12; nb = n12; }
    if (!(nb > 0.0f)) return 0;
    float inv = 1.0f / sqrtf(nb);
    v[0] = best[0] * inv; v[1] = best[1] * inv; v[2] = best[2] * inv;
    return 1;
}

/* ------------------------------------------------------------------ P3P */
int ref_p3p(const float x[3][3], const float y[3][3], float Rs[4][9], float ts[4][3]) {
    float d12[3] = {x[0][0] - x[1][0], x[0][1] - x[1][1], x[0][2] - x[1][2]};
    float d13[3] = {x[0][0] - x[2][0], x[0][1] - x[2][1], x[0][2] - x[2][2]};
    float d23[3] = {x[1][0] - x[2][0], x[1][1] - x[2][1], x[1][2] - x[2][2]};
    float a12 = dot3(d12, d12), a13 = dot3(d13, d13), a23 = dot3(d23, d23);
    if (!(a12 > 1e-12f) || !(a13 > 1e-12f) || !(a23 > 1e-12f)) return 0;
    float b12 = dot3(y[0], y[1]), b13 = dot3(y[0], y[2]), b23 = dot3(y[1], y[2]);

    /* object-side frame, shared by all solutions: Xinv = [d12 d13 d12xd13]^-1 */
    float xc[3];
    cross3(d12, d13, xc);
    float detX = dot3(xc, xc);
    if (!(detX > 1e-24f)) return 0;
    float Xinv[9];
    {
        float r0[3], r1[3];
        cross3(d13, xc, r0);   /* rows of the inverse (before 1/det) */
        cross3(xc, d12, r1);
        float id = 1.0f / detX;
        Xinv[0] = r0[0] * id; Xinv[1] = r0[1] * id; Xinv[2] = r0[2] * id;
        Xinv[3] = r1[0] * id; Xinv[4] = r1[1] * id; Xinv[5] = r1[2] * id;
        Xinv[6] = xc[0] * id; Xinv[7] = xc[1] * id; Xinv[8] = xc[2] * id;
    }

    float D1[6] = {a23, -(a23 * b12), 0.0f, a23 - a12, a12 * b23, -a12};
    float D2[6] = {a23, 0.0f, -(a23 * b13), -a13, a13 * b23, a23 - a13};
    float A1[6], A2[6];
    sym_adj(D1, A1); sym_adj(D2, A2);
    float c0 = sym_det(D1, A1), c3 = sym_det(D2, A2);
    float c1 = sym_tr_prod(A1, D2), c2 = sym_tr_prod(A2, D1);

    int swapped = fabsf(c0) > fabsf(c3);
    float lead = swapped ? c0 : c3;
    if (!(fabsf(lead) > 1e-36f)) return 0;
    float ca, cb, cc;
    if (!swapped) { ca = c2 / lead; cb = c1 / lead; cc = c0 / lead; }
    else          { ca = c1 / lead; cb = c2 / lead; cc = c3 / lead; }
    float g = cubic_real_root(ca, cb, cc);
    if (!isfinite(g)) return 0;

    float E[6];
    const float* Q;
    if (!swapped) { for (int i = 0; i < 6; ++i) E[i] = fmaf(g, D2[i], D1[i]); Q = D2; }
    else          { for (int i = 0; i < 6; ++i) E[i] = fmaf(g, D1[i], D2[i]); Q = D1; }

    float tr = E[0] + E[3] + E[5];
    float mm = fmaf(E[0], E[3], -(E[1] * E[1])) + fmaf(E[0], E[5], -(E[2] * E[2])) +
               fmaf(E[3], E[5], -(E[4] * E[4]));
    if (!(mm < 0.0f)) return 0;
    float sq = sqrtf(fmaf(tr, tr, -(4.0f * mm)));
    float s1, s2;
    if (tr >= 0.0f) { s1 = 0.5f * (tr + sq); s2 = mm / s1; }
    else            { s2 = 0.5f * (tr - sq); s1 = mm / s2; }
    float e1[3], e2[3];
    if (!sym_eigvec(E, s1, e1) || !sym_eigvec(E, s2, e2)) return 0;
    float sr = sqrtf(-s2 / s1);
    float asum = a12 + a13 + a23;

    int ns = 0;
    for (int sgi = 0; sgi < 2; ++sgi) {
        float sg = sgi == 0 ? sr : -sr;
        float n[3] = {fmaf(sg, e2[0], e1[0]), fmaf(sg, e2[1], e1[1]), fmaf(sg, e2[2], e1[2])};
        float nn = dot3(n, n);
        if (!(nn > 0.0f)) continue;
        float inn = 1.0f / sqrtf(nn);
        n[0] *= inn; n[1] *= inn; n[2] *= inn;
        /* branch-free orthonormal basis of the plane n.L = 0 */
        float sz = copysignf(1.0f, n[2]);
        float aa = -1.0f / (sz + n[2]);
        float bb = n[0] * n[1] * aa;
        float p[3] = {fmaf(sz * n[0], n[0] * aa, 1.0f), sz * bb, -(sz * n[0])};
        float q[3] = {bb, fmaf(n[1], n[1] * aa, sz), -n[1]};
        float qa = sym_quad(Q, p, p), qb = sym_quad(Q, p, q), qc = sym_quad(Q, q, q);
        float dsc = fmaf(qb, qb, -(qa * qc));
        if (!(dsc >= 0.0f)) continue;
        float sd = sqrtf(dsc);
        int use_a = fabsf(qa) >= fabsf(qc);
        float den = use_a ? qa : qc;
        if (den == 0.0f) continue;
        for (int ri = 0; ri < 2; ++ri) {
            float r = ((ri == 0 ? sd : -sd) - qb) / den;
            float L[3];
            if (use_a) { L[0] = fmaf(r, p[0], q[0]); L[1] = fmaf(r, p[1], q[1]); L[2] = fmaf(r, p[2], q[2]); }
            else       { L[0] = fmaf(r, q[0], p[0]); L[1] = fmaf(r, q[1], p[1]); L[2] = fmaf(r, q[2], p[2]); }
            int pos = (L[0] > 0.0f) && (L[1] > 0.0f) && (L[2] > 0.0f);
            int neg = (L[0] < 0.0f) && (L[1] < 0.0f) && (L[2] < 0.0f);
            if (!pos && !neg) continue;
            float sq_sum = fmaf(L[0], L[0], fmaf(L[1], L[1], L[2] * L[2]));
            float cr_sum = fmaf(b12 * L[0], L[1], fmaf(b13 * L[0], L[2], (b23 * L[1]) * L[2]));
            float ms = 2.0f * (sq_sum - cr_sum);
            if (!(ms > 0.0f)) continue;
            float k = sqrtf(asum / ms);
            if (neg) k = -k;
            L[0] *= k; L[1] *= k; L[2] *= k;
            /* Gauss-Newton polish of the three depth constraints */
            for (int it = 0; it < 3; ++it) {
                float f0 = fmaf(L[0], L[0], fmaf(L[1], L[1], -(2.0f * b12) * (L[0] * L[1]))) - a12;
                float f1 = fmaf(L[0], L[0], fmaf(L[2], L[2], -(2.0f * b13) * (L[0] * L[2]))) - a13;
                float f2 = fmaf(L[1], L[1], fmaf(L[2], L[2], -(2.0f * b23) * (L[1] * L[2]))) - a23;
                float j00 = 2.0f * fmaf(-b12, L[1], L[0]), j01 = 2.0f * fmaf(-b12, L[0], L[1]);
                float j10 = 2.0f * fmaf(-b13, L[2], L[0]), j12 = 2.0f * fmaf(-b13, L[0], L[2]);
                float j21 = 2.0f * fmaf(-b23, L[2], L[1]), j22 = 2.0f * fmaf(-b23, L[1], L[2]);
                /* J = [j00 j01 0; j10 0 j12; 0 j21 j22] */
                float det = -(fmaf(j00 * j12, j21, (j01 * j10) * j22));
                if (!(fabsf(det) > 1e-30f)) break;
                float idet = 1.0f / det;
                /* cofactor inverse applied to f */
                float g0 = fmaf(-(j12 * j21), f0, fmaf(-(j01 * j22), f1, (j01 * j12) * f2));
                float g1 = fmaf(-(j10 * j22), f0, fmaf(j00 * j22, f1, -(j00 * j12) * f2));
                float g2 = fmaf(j10 * j21, f0, fmaf(-(j00 * j21), f1, -(j01 * j10) * f2));
                L[0] = fmaf(-g0, idet, L[0]);
                L[1] = fmaf(-g1, idet, L[1]);
                L[2] = fmaf(-g2, idet, L[2]);
            }
            if (!((L[0] > 0.0f) && (L[1] > 0.0f) && (L[2] > 0.0f))) continue;
            float p0[3] = {L[0] * y[0][0], L[0] * y[0][1], L[0] * y[0][2]};
            float p1[3] = {L[1] * y[1][0], L[1] * y[1][1], L[1] * y[1][2]};
            float p2[3] = {L[2] * y[2][0], L[2] * y[2][1], L[2] * y[2][2]};
            float y12[3] = {p0[0] - p1[0], p0[1] - p1[1], p0[2] - p1[2]};
            float y13[3] = {p0[0] - p2[0], p0[1] - p2[1], p0[2] - p2[2]};
            float yc[3];
            cross3(y12, y13, yc);
            float* R = Rs[ns];
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j)
                    R[3 * i + j] = fmaf(y12[i], Xinv[j], fmaf(y13[i], Xinv[3 + j], yc[i] * Xinv[6 + j]));
            float* t = ts[ns];
            for (int i = 0; i < 3; ++i)
                t[i] = p0[i] - fmaf(R[3 * i], x[0][0], fmaf(R[3 * i + 1], x[0][1], R[3 * i + 2] * x[0][2]));
            int ok = finite3(t);
            for (int i = 0; i < 9; ++i) ok = ok && isfinite(R[i]);
            if (ok) ++ns;
        }
    }
    return ns;
}

/* ------------------------------------------------------------------ hypothesis / scoring */
static inline void make_proj(const float K4[4], const float R[9], const float t[3], float M[12]) {
    float fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3];
    M[0] = fmaf(fx, R[0], cx * R[6]); M[1] = fmaf(fx, R[1], cx * R[7]);
    M[2] = fmaf(fx, R[2], cx * R[8]); M[3] = fmaf(fx, t[0], cx * t[2]);
    M[4] = fmaf(fy, R[3], cy * R[6]); M[5] = fmaf(fy, R[4], cy * R[7]);
    M[6] = fmaf(fy, R[5], cy * R[8]); M[7] = fmaf(fy, t[1], cy * t[2]);
    M[8] = R[6]; M[9] = R[7]; M[10] = R[8]; M[11] = t[2];
}

/* Inlier test in threshold units (division-free; cv2 tests |proj - uv|^2 <= thr^2):
 *   Ms = K[R|t] with rows 0,1 multiplied by inv_thr = 1/thr; (us,vs) = (u,v)*inv_thr;
 *   s  = (x - us*w)^2 + (y - vs*w)^2 - w*|w|      inlier  <=>  s < 0
 * (w <= 0 makes -w*|w| >= 0, so points behind the camera never pass; NaN never passes).
 * The operation order below is the contract shared with csrc/common.cuh::inlier_s. */
static inline void scale_proj(const float M[12], float inv_thr, float Ms[12]) {
    for (int k = 0; k < 8; ++k) Ms[k] = M[k] * inv_thr;
    for (int k = 8; k < 12; ++k) Ms[k] = M[k];
}
static inline int inlier_test(const float Ms[12], float X, float Y, float Z, float us, float vs) {
    float x = fmaf(Ms[0], X, fmaf(Ms[1], Y, fmaf(Ms[2], Z, Ms[3])));
    float y = fmaf(Ms[4], X, fmaf(Ms[5], Y, fmaf(Ms[6], Z, Ms[7])));
    float w = fmaf(Ms[8], X, fmaf(Ms[9], Y, fmaf(Ms[10], Z, Ms[11])));
    float dx = fmaf(-us, w, x);
    float dy = fmaf(-vs, w, y);
    float q = w * fabsf(w);
    float t = fmaf(dy, dy, -q);
    float s = fmaf(dx, dx, t);
    return s < 0.0f;
}

int ref_hypothesis(const float* X, const float* Y, const float* Z, const float* U, const float* V,
                   const float K4[4], const int32_t idx[4], float M[12], float R[9], float t[3]) {
    float fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3];
    float x[3][3], y[3][3];
    for (int k = 0; k < 3; ++k) {
        int i = idx[k];
        x[k][0] = X[i]; x[k][1] = Y[i]; x[k][2] = Z[i];
        float xn = (U[i] - cx) / fx, yn = (V[i] - cy) / fy;
        float inv = 1.0f / sqrtf(fmaf(xn, xn, fmaf(yn, yn, 1.0f)));
        y[k][0] = xn * inv; y[k][1] = yn * inv; y[k][2] = inv;
    }
    float Rs[4][9], ts[4][3];
    int ns = ref_p3p(x, y, Rs, ts);
    int i4 = idx[3];
    float X4 = X[i4], Y4 = Y[i4], Z4 = Z[i4], u4 = U[i4], v4 = V[i4];
    int best = -1; float beste = 0.0f;
    for (int s = 0; s < ns; ++s) {
        const float* Rr = Rs[s]; const float* tt = ts[s];
        float xc = fmaf(Rr[0], X4, fmaf(Rr[1], Y4, fmaf(Rr[2], Z4, tt[0])));
        float yc = fmaf(Rr[3], X4, fmaf(Rr[4], Y4, fmaf(Rr[5], Z4, tt[1])));
        float zc = fmaf(Rr[6], X4, fmaf(Rr[7], Y4, fmaf(Rr[8], Z4, tt[2])));
        if (!(zc > 0.0f)) continue;
        float du = fmaf(fx, xc / zc, cx) - u4;
        float dv = fmaf(fy, yc / zc, cy) - v4;
        float e = fmaf(du, du, dv * dv);
        if (!isfinite(e)) continue;
        if (best < 0 || e < beste) { best = s; beste = e; }
    }
    if (best < 0) return 0;
    memcpy(R, Rs[best], 9 * sizeof(float));
    memcpy(t, ts[best], 3 * sizeof(float));
    make_proj(K4, R, t, M);
    return 1;
}

int ref_score(const float M[12], const float* X, const float* Y, const float* Z, const float* U,
              const float* V, int P, float inv_thr, uint32_t* bitmap) {
    int n = 0;
    float Ms[12];
    scale_proj(M, inv_thr, Ms);
    if (bitmap) memset(bitmap, 0, sizeof(uint32_t) * (size_t)((P + 31) / 32));
    for (int i = 0; i < P; ++i) {
        int in = inlier_test(Ms, X[i], Y[i], Z[i], U[i] * inv_thr, V[i] * inv_thr);
        n += in;
        if (bitmap && in) bitmap[i >> 5] |= 1u << (i & 31);
    }
    return n;
}

/* ------------------------------------------------------------------ refine (Gauss-Newton) */
#define GN_EPS 1e-6f /* Gauss-Newton stops early once a step is this small (rad / m) */
#define RT 256 /* threads of the finalize CTA whose reduction tree is emulated */

/* emulate: per-thread strided partials -> xor-butterfly per warp -> sequential sum over warps */
static float tree_sum(float part[RT]) {
    float tot = 0.0f;
    for (int w = 0; w < RT / 32; ++w) {
        float v[32], nv[32];
        memcpy(v, part + 32 * w, sizeof(v));
        for (int off = 16; off >= 1; off >>= 1) {
            for (int l = 0; l < 32; ++l) nv[l] = v[l] + v[l ^ off];
            memcpy(v, nv, sizeof(v));
        }
        tot = (w == 0) ? v[0] : tot + v[0];
    }
    return tot;
}

void ref_rot_to_quat(const float R[9], float q[4]) {
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

static void quat_to_rot(const float q[4], float R[9]) {
    float w = q[0], x = q[1], y = q[2], z = q[3];
    float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z;
    float wx = w * x, wy = w * y, wz = w * z;
    R[0] = 1.0f - 2.0f * (yy + zz); R[1] = 2.0f * (xy - wz);        R[2] = 2.0f * (xz + wy);
    R[3] = 2.0f * (xy + wz);        R[4] = 1.0f - 2.0f * (xx + zz); R[5] = 2.0f * (yz - wx);
    R[6] = 2.0f * (xz - wy);        R[7] = 2.0f * (yz + wx);        R[8] = 1.0f - 2.0f * (xx + yy);
}

/* 6x6 SPD solve in double (Cholesky); A upper triangle packed row-major 21, returns 0 if not PD */
static int chol6_solve(const double A21[21], const double g[6], double d[6]) {
    double L[6][6];
    memset(L, 0, sizeof(L));
    double A[6][6];
    int k = 0;
    for (int i = 0; i < 6; ++i)
        for (int j = i; j < 6; ++j) { A[i][j] = A21[k]; A[j][i] = A21[k]; ++k; }
    for (int j = 0; j < 6; ++j) {
        double s = A[j][j];
        for (int m = 0; m < j; ++m) s = s - L[j][m] * L[j][m];
        if (!(s > 0.0)) return 0;
        double ljj = sqrt(s);
        L[j][j] = 1.0 / ljj;    /* the diagonal holds the RECIPROCAL (one division per column) */
        for (int i = j + 1; i < 6; ++i) {
            double v = A[i][j];
            for (int m = 0; m < j; ++m) v = v - L[i][m] * L[j][m];
            L[i][j] = v * L[j][j];
        }
    }
    double z[6];
    for (int i = 0; i < 6; ++i) {
        double v = g[i];
        for (int m = 0; m < i; ++m) v = v - L[i][m] * z[m];
        z[i] = v * L[i][i];
    }
    for (int i = 5; i >= 0; --i) {
        double v = z[i];
        for (int m = i + 1; m < 6; ++m) v = v - L[m][i] * d[m];
        d[i] = v * L[i][i];
    }
    return 1;
}

/* refit pose on the inlier bitmap; q,t in/out. returns 0 if skipped */
static int refine_pose(const float* X, const float* Y, const float* Z, const float* U, const float* V,
                       int P, const uint32_t* bitmap, const float K4[4], int gn_iters, float q[4], float t[3]) {
    float fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3];
    static float part[28][RT];
    /* centroid of inlier object points */
    int cnt = 0;
    for (int th = 0; th < RT; ++th) {
        float sx = 0.0f, sy = 0.0f, sz = 0.0f;
        for (int i = th; i < P; i += RT)
            if (bitmap[i >> 5] >> (i & 31) & 1u) { sx += X[i]; sy += Y[i]; sz += Z[i]; ++cnt; }
        part[0][th] = sx; part[1][th] = sy; part[2][th] = sz;
    }
    if (cnt < 4) return 0;
    float fn = (float)cnt;
    float cb[3] = {tree_sum(part[0]) / fn, tree_sum(part[1]) / fn, tree_sum(part[2]) / fn};
    float R[9];
    quat_to_rot(q, R);
    float tc[3];
    for (int i = 0; i < 3; ++i)
        tc[i] = fmaf(R[3 * i], cb[0], fmaf(R[3 * i + 1], cb[1], fmaf(R[3 * i + 2], cb[2], t[i])));
    for (int it = 0; it < gn_iters; ++it) {
        quat_to_rot(q, R);
        for (int th = 0; th < RT; ++th) {
            float acc[27];
            for (int k = 0; k < 27; ++k) acc[k] = 0.0f;
            for (int i = th; i < P; i += RT) {
                if (!(bitmap[i >> 5] >> (i & 31) & 1u)) continue;
                float xo = X[i] - cb[0], yo = Y[i] - cb[1], zo = Z[i] - cb[2];
                float xr = fmaf(R[0], xo, fmaf(R[1], yo, R[2] * zo));
                float yr = fmaf(R[3], xo, fmaf(R[4], yo, R[5] * zo));
                float zr = fmaf(R[6], xo, fmaf(R[7], yo, R[8] * zo));
                float xc = xr + tc[0], yc = yr + tc[1], zc = zr + tc[2];
                float iz = 1.0f / zc;
                float un = xc * iz, vn = yc * iz;
                float ru = fmaf(fx, un, cx) - U[i];
                float rv = fmaf(fy, vn, cy) - V[i];
                float a = fx * iz, b = fy * iz;
                float c = -(a * un), d = -(b * vn);
                /* Ju = (a,0,c), Jv = (0,b,d) wrt camera point; d/dw = Xr x J */
                float ju[6] = {yr * c, fmaf(zr, a, -(xr * c)), -(yr * a), a, 0.0f, c};
                float jv[6] = {fmaf(yr, d, -(zr * b)), -(xr * d), xr * b, 0.0f, b, d};
                int k = 0;
                for (int r = 0; r < 6; ++r)
                    for (int s = r; s < 6; ++s) { acc[k] = fmaf(ju[r], ju[s], fmaf(jv[r], jv[s], acc[k])); ++k; }
                for (int r = 0; r < 6; ++r) acc[21 + r] = fmaf(ju[r], ru, fmaf(jv[r], rv, acc[21 + r]));
            }
            for (int k = 0; k < 27; ++k) part[k][th] = acc[k];
        }
        double A21[21], g[6], d[6];
        for (int k = 0; k < 21; ++k) A21[k] = (double)tree_sum(part[k]);
        for (int k = 0; k < 6; ++k) g[k] = -(double)tree_sum(part[21 + k]);
        if (!chol6_solve(A21, g, d)) break;
        float dw[3] = {(float)d[0], (float)d[1], (float)d[2]};
        float dt[3] = {(float)d[3], (float)d[4], (float)d[5]};
        if (!finite3(dw) || !finite3(dt)) break;
        /* q <- normalize((1, dw/2) * q) */
        float hx = 0.5f * dw[0], hy = 0.5f * dw[1], hz = 0.5f * dw[2];
        float w0 = q[0], x0 = q[1], y0 = q[2], z0 = q[3];
        float nw = w0 - fmaf(hx, x0, fmaf(hy, y0, hz * z0));
        float nx = x0 + fmaf(hx, w0, fmaf(hy, z0, -(hz * y0)));
        float ny = y0 + fmaf(hy, w0, fmaf(hz, x0, -(hx * z0)));
        float nz = z0 + fmaf(hz, w0, fmaf(hx, y0, -(hy * x0)));
        float in = 1.0f / sqrtf(fmaf(nw, nw, fmaf(nx, nx, fmaf(ny, ny, nz * nz))));
        q[0] = nw * in; q[1] = nx * in; q[2] = ny * in; q[3] = nz * in;
        tc[0] += dt[0]; tc[1] += dt[1]; tc[2] += dt[2];
        /* converged: the step just applied is below 1e-6 rad and 1e-6 m in every component */
        if (fmaxf(fmaxf(fabsf(dw[0]), fabsf(dw[1])), fabsf(dw[2])) < GN_EPS &&
            fmaxf(fmaxf(fabsf(dt[0]), fabsf(dt[1])), fabsf(dt[2])) < GN_EPS) break;
    }
    quat_to_rot(q, R);
    for (int i = 0; i < 3; ++i)
        t[i] = tc[i] - fmaf(R[3 * i], cb[0], fmaf(R[3 * i + 1], cb[1], R[3 * i + 2] * cb[2]));
    return 1;
}

int ref_ransac_frame(const float* X, const float* Y, const float* Z, const float* U, const float* V,
                     int P, const float K4[4], const ref_ransac_params* prm, uint32_t frame,
                     const int32_t* sample_idx, float R[9], float t[3], int32_t* n_inl,
                     int32_t* best_hyp, int32_t* n_inl_refined, int32_t* hyp_counts,
                     uint32_t* bitmap, float R0[9], float t0[3]) {
    static const float I9[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    memcpy(R, I9, sizeof(I9)); t[0] = t[1] = t[2] = 0.0f;
    if (R0) memcpy(R0, I9, sizeof(I9));
    if (t0) t0[0] = t0[1] = t0[2] = 0.0f;
    *n_inl = 0; *best_hyp = -1; *n_inl_refined = 0;
    int words = (P + 31) / 32;
    if (bitmap) memset(bitmap, 0, sizeof(uint32_t) * (size_t)words);
    int H = prm->n_hyp;
    if (hyp_counts) for (int h = 0; h < H; ++h) hyp_counts[h] = -1;
    if (P < 4) return 1;
    float inv_thr = 1.0f / prm->thresh_px;
    int bc = -1, bh = -1;
    float bR[9], bt[3], bM[12];
    for (int h = 0; h < H; ++h) {
        int32_t idx[4];
        if (sample_idx) {
            memcpy(idx, sample_idx + 4 * h, sizeof(idx));
            int ok = 1;
            for (int k = 0; k < 4; ++k) ok = ok && idx[k] >= 0 && idx[k] < P;
            if (!ok) continue;
        } else if (!ref_sample4(prm->seed_lo, prm->seed_hi, frame, (uint32_t)h, P, idx)) {
            continue;
        }
        float M[12], Rh[9], th[3];
        if (!ref_hypothesis(X, Y, Z, U, V, K4, idx, M, Rh, th)) continue;
        int c = ref_score(M, X, Y, Z, U, V, P, inv_thr, NULL);
        if (hyp_counts) hyp_counts[h] = c;
        if (c > bc) { bc = c; bh = h; memcpy(bR, Rh, sizeof(bR)); memcpy(bt, th, sizeof(bt)); memcpy(bM, M, sizeof(bM)); }
    }
    if (bh < 0 || bc < prm->min_inliers) return 2;
    *n_inl = bc; *best_hyp = bh;
    if (R0) memcpy(R0, bR, sizeof(bR));
    if (t0) memcpy(t0, bt, sizeof(bt));
    uint32_t* bm = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)words);
    ref_score(bM, X, Y, Z, U, V, P, inv_thr, bm);
    if (bitmap) memcpy(bitmap, bm, sizeof(uint32_t) * (size_t)words);
    float q[4], tt[3] = {bt[0], bt[1], bt[2]};
    ref_rot_to_quat(bR, q);
    int nref = bc;
    for (int round = 0; round <= prm->lo_rounds; ++round) {
        if (!refine_pose(X, Y, Z, U, V, P, bm, K4, prm->gn_iters, q, tt)) break;
        float Rq[9], Mq[12];
        quat_to_rot(q, Rq);
        make_proj(K4, Rq, tt, Mq);
        nref = ref_score(Mq, X, Y, Z, U, V, P, inv_thr, bm);
    }
    /* a refit that lost its consensus (fewer than half of the minimal hypothesis' inliers, or fewer than
       min_inliers, under the refitted pose) is discarded: the minimal hypothesis is returned instead */
    int need = (bc + 1) / 2 > prm->min_inliers ? (bc + 1) / 2 : prm->min_inliers;
    if (nref < need) {
        memcpy(R, bR, sizeof(bR)); t[0] = bt[0]; t[1] = bt[1]; t[2] = bt[2];
        nref = bc;
    } else {
        quat_to_rot(q, R);
        t[0] = tt[0]; t[1] = tt[1]; t[2] = tt[2];
    }
    *n_inl_refined = nref;
    free(bm);
    return 0;
}

/* ------------------------------------------------------------------ map compaction */
int ref_compact_map(const float* xyz, const uint8_t* mask, int Hh, int Ww, const float extent[3],
                    const float lfborder[3], int rmin, int cmin, int cap,
                    float* X, float* Y, float* Z, float* U, float* V, int32_t* total) {
    size_t plane = (size_t)Hh * Ww;
    int n = 0, tot = 0;
    for (int r = 0; r < Hh; ++r)
        for (int c = 0; c < Ww; ++c) {
            size_t o = (size_t)r * Ww + c;
            float a = xyz[o], b = xyz[plane + o], d = xyz[2 * plane + o];
            int m = mask ? (mask[o] != 0) : (a != 0.0f || b != 0.0f || d != 0.0f);
            if (!m) continue;
            if (n < cap) {
                X[n] = fmaf(a, extent[0], lfborder[0]);
                Y[n] = fmaf(b, extent[1], lfborder[1]);
                Z[n] = fmaf(d, extent[2], lfborder[2]);
                U[n] = (float)(c + cmin);
                V[n] = (float)(r + rmin);
                ++n;
            }
            ++tot;
        }
    if (total) *total = tot;
    return n;
}

/* ------------------------------------------------------------------ KNN / ADD / metrics */
static inline float sqdist3(const float* a, const float* b) {
    float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    return fmaf(dz, dz, fmaf(dy, dy, dx * dx));
}

void ref_knn1(const float* x, const float* y, int N, int M, int64_t* idx, float* d2) {
    for (int i = 0; i < N; ++i) {
        float best = INFINITY; int64_t bj = 0;
        for (int j = 0; j < M; ++j) {
            float d = sqdist3(x + 3 * i, y + 3 * j);
            if (d < best) { best = d; bj = j; }
        }
        idx[i] = bj;
        if (d2) d2[i] = best;
    }
}

void ref_add_adds(const float* pred, const float* tgt, int N, int M, float* add, float* adds) {
    if (add) {
        double s = 0.0;
        int n = N < M ? N : M;
        for (int i = 0; i < n; ++i) s += (double)sqrtf(sqdist3(pred + 3 * i, tgt + 3 * i));
        *add = (float)(s / (double)n);
    }
    if (adds) {
        double s = 0.0;
        for (int i = 0; i < M; ++i) { /* for each target point: nearest predicted point */
            float best = INFINITY;
            for (int j = 0; j < N; ++j) {
                float d = sqdist3(tgt + 3 * i, pred + 3 * j);
                if (d < best) best = d;
            }
            s += (double)sqrtf(best);
        }
        *adds = (float)(s / (double)M);
    }
}

void ref_transform(const float* model, int N, const float R[9], const float t[3], float* out) {
    for (int i = 0; i < N; ++i) {
        float x = model[3 * i], y = model[3 * i + 1], z = model[3 * i + 2];
        out[3 * i]     = fmaf(R[0], x, fmaf(R[1], y, fmaf(R[2], z, t[0])));
        out[3 * i + 1] = fmaf(R[3], x, fmaf(R[4], y, fmaf(R[5], z, t[1])));
        out[3 * i + 2] = fmaf(R[6], x, fmaf(R[7], y, fmaf(R[8], z, t[2])));
    }
}

float ref_angular_deg(const float R1[9], const float R2[9]) {
    float q1[4], q2[4];
    ref_rot_to_quat(R1, q1); ref_rot_to_quat(R2, q2);
    float d = fabsf(fmaf(q1[0], q2[0], fmaf(q1[1], q2[1], fmaf(q1[2], q2[2], q1[3] * q2[3]))));
    const float eps = 1e-7f;
    if (d > 1.0f - eps) d = 1.0f - eps;
    if (d < -1.0f + eps) d = -1.0f + eps;
    return 2.0f * acosf(d) / 3.14159265358979323846f * 180.0f;
}

void ref_vote_translation(const float* cloud, const float* t_res, int NP, float t[3]) {
    for (int c = 0; c < 3; ++c) {
        double s = 0.0;
        for (int i = 0; i < NP; ++i) s += (double)(cloud[3 * i + c] + t_res[(size_t)c * NP + i]);
        t[c] = (float)(s / (double)NP);
    }
}

#define RESAMPLE_TAG 0x52534d50u

void ref_resample_inliers(const uint32_t* bitmap, int P, int n, uint32_t seed_lo, uint32_t seed_hi,
                          uint32_t frame, int32_t* out) {
    int m = 0;
    for (int i = 0; i < P; ++i) m += (bitmap[i >> 5] >> (i & 31)) & 1u;
    if (m == 0) { for (int k = 0; k < n; ++k) out[k] = -1; return; }
    uint32_t key[2] = {seed_lo, seed_hi};
    if (m > n) {
        int sel = 0, j = 0;
        uint32_t r[4] = {0, 0, 0, 0};
        for (int i = 0; i < P && sel < n; ++i) {
            if (!((bitmap[i >> 5] >> (i & 31)) & 1u)) continue;
            if ((j & 3) == 0) {
                uint32_t ctr[4] = {(uint32_t)(j >> 2), frame, 0u, RESAMPLE_TAG};
                ref_philox4x32(ctr, key, r);
            }
            uint32_t pick = mulhi32(r[j & 3], (uint32_t)(m - j));
            if (pick < (uint32_t)(n - sel)) out[sel++] = i;
            ++j;
        }
    } else {
        int k = 0;
        for (int i = 0; i < P; ++i)
            if ((bitmap[i >> 5] >> (i & 31)) & 1u) out[k++] = i;
        for (; k < n; ++k) out[k] = out[k - m];
    }
}
