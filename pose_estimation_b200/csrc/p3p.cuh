// p3p.cuh -- closed-form, trig-free P3P minimal solver (one hypothesis per thread).
//
// Replaces the 5-point EPnP minimal solve inside cv2.solvePnPRansac (tools/trainer.py:423-427).
// Formulation: the three depth constraints  l_i^2 + l_j^2 - 2 b_ij l_i l_j = a_ij  are reduced to
// two homogeneous quadrics L^T D1 L = L^T D2 L = 0; a real root g of det(D1 + g D2) = 0 (cubic,
// Newton with a monotone start, no acos/cos/cbrt) makes the pencil member degenerate, i.e. a pair
// of planes obtained from its 2x2 eigen-structure; each plane cut with the remaining quadric is a
// quadratic -> up to four depth triples, polished by three Gauss-Newton steps, then R,t from the
// aligned difference frames.  The fourth sampled point picks the solution (smallest squared pixel
// error, positive depth).  Only + - * / sqrt fma are used, in a written order, so the CPU oracle
// reproduces every bit.
#pragma once
#include "common.cuh"

namespace p3p {

// symmetric 3x3 as s[6] = {00,01,02,11,12,22}
__device__ __forceinline__ void sym_adj(const float s[6], float A[6]) {
    A[0] = fmaf(s[3], s[5], -(s[4] * s[4]));
    A[1] = fmaf(s[2], s[4], -(s[1] * s[5]));
    A[2] = fmaf(s[1], s[4], -(s[2] * s[3]));
    A[3] = fmaf(s[0], s[5], -(s[2] * s[2]));
    A[4] = fmaf(s[1], s[2], -(s[0] * s[4]));
    A[5] = fmaf(s[0], s[3], -(s[1] * s[1]));
}
__device__ __forceinline__ float sym_det(const float s[6], const float A[6]) {
    return fmaf(s[0], A[0], fmaf(s[1], A[1], s[2] * A[2]));
}
__device__ __forceinline__ float sym_tr_prod(const float A[6], const float t[6]) {
    float d = fmaf(A[0], t[0], fmaf(A[3], t[3], A[5] * t[5]));
    float o = fmaf(A[1], t[1], fmaf(A[2], t[2], A[4] * t[4]));
    return fmaf(2.0f, o, d);
}
__device__ __forceinline__ float sym_quad(const float s[6], const float v[3], const float w[3]) {
    float r0 = fmaf(s[0], w[0], fmaf(s[1], w[1], s[2] * w[2]));
    float r1 = fmaf(s[1], w[0], fmaf(s[3], w[1], s[4] * w[2]));
    float r2 = fmaf(s[2], w[0], fmaf(s[4], w[1], s[5] * w[2]));
    return fmaf(v[0], r0, fmaf(v[1], r1, v[2] * r2));
}

__device__ __forceinline__ float cbrt_up(float m) {
    if (!(m > 0.0f)) return 0.0f;
    int i = __float_as_int(m);
    i = i / 3 + 709921077;
    float y = __int_as_float(i);
    y = y - (y * y * y - m) / (3.0f * y * y);
    y = y - (y * y * y - m) / (3.0f * y * y);
    return y * 1.02f;
}
__device__ __forceinline__ float cubic_eval(float a, float b, float c, float x) {
    return fmaf(fmaf(x + a, x, b), x, c);
}
__device__ __noinline__ float cubic_real_root(float a, float b, float c) {
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

__device__ __forceinline__ bool sym_eigvec(const float e[6], float sigma, float v[3]) {
    float r0[3] = {e[0] - sigma, e[1], e[2]};
    float r1[3] = {e[1], e[3] - sigma, e[4]};
    float r2[3] = {e[2], e[4], e[5] - sigma};
    float c01[3], c02[3], c12[3];
    cross3(r0, r1, c01); cross3(r0, r2, c02); cross3(r1, r2, c12);
    float n01 = dot3(c01, c01), n02 = dot3(c02, c02), n12 = dot3(c12, c12);
    float b0 = c01[0], b1 = c01[1], b2 = c01[2], nb = n01;
    if (n02 > nb) { b0 = c02[0]; b1 = c02[1]; b2 = c02[2]; nb = n02; }
    if (n12 > nb) { b0 = c12[0]; b1 = c12[1]; b2 = c12[2]; nb = n12; }
    if (!(nb > 0.0f)) return false;
    float inv = 1.0f / sqrtf(nb);
    v[0] = b0 * inv; v[1] = b1 * inv; v[2] = b2 * inv;
    return true;
}

__device__ __forceinline__ bool fin(float a) { return isfinite(a); }

// One hypothesis from 4 correspondences (x = object points, (u,v) pixels).  The first three feed
// P3P, the fourth disambiguates.  Writes the chosen R[9], t[3]; returns false if no model.
__device__ __noinline__ bool solve(const float x[4][3], const float uv[4][2], float fx, float fy, float cx,
                                   float cy, float Rout[9], float tout[3]) {
    float y[3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float xn = (uv[k][0] - cx) / fx, yn = (uv[k][1] - cy) / fy;
        float inv = 1.0f / sqrtf(fmaf(xn, xn, fmaf(yn, yn, 1.0f)));
        y[k][0] = xn * inv; y[k][1] = yn * inv; y[k][2] = inv;
    }
    float d12[3] = {x[0][0] - x[1][0], x[0][1] - x[1][1], x[0][2] - x[1][2]};
    float d13[3] = {x[0][0] - x[2][0], x[0][1] - x[2][1], x[0][2] - x[2][2]};
    float d23[3] = {x[1][0] - x[2][0], x[1][1] - x[2][1], x[1][2] - x[2][2]};
    float a12 = dot3(d12, d12), a13 = dot3(d13, d13), a23 = dot3(d23, d23);
    if (!(a12 > 1e-12f) || !(a13 > 1e-12f) || !(a23 > 1e-12f)) return false;
    float b12 = dot3(y[0], y[1]), b13 = dot3(y[0], y[2]), b23 = dot3(y[1], y[2]);

    float xc[3];
    cross3(d12, d13, xc);
    float detX = dot3(xc, xc);
    if (!(detX > 1e-24f)) return false;
    float Xinv[9];
    {
        float r0[3], r1[3];
        cross3(d13, xc, r0);
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

    bool swapped = fabsf(c0) > fabsf(c3);
    float lead = swapped ? c0 : c3;
    if (!(fabsf(lead) > 1e-36f)) return false;
    float ca, cb, cc;
    if (!swapped) { ca = c2 / lead; cb = c1 / lead; cc = c0 / lead; }
    else          { ca = c1 / lead; cb = c2 / lead; cc = c3 / lead; }
    float g = cubic_real_root(ca, cb, cc);
    if (!fin(g)) return false;

    float E[6], Q[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        E[i] = swapped ? fmaf(g, D1[i], D2[i]) : fmaf(g, D2[i], D1[i]);
        Q[i] = swapped ? D1[i] : D2[i];
    }
    float tr = E[0] + E[3] + E[5];
    float mm = fmaf(E[0], E[3], -(E[1] * E[1])) + fmaf(E[0], E[5], -(E[2] * E[2])) +
               fmaf(E[3], E[5], -(E[4] * E[4]));
    if (!(mm < 0.0f)) return false;
    float sq = sqrtf(fmaf(tr, tr, -(4.0f * mm)));
    float s1, s2;
    if (tr >= 0.0f) { s1 = 0.5f * (tr + sq); s2 = mm / s1; }
    else            { s2 = 0.5f * (tr - sq); s1 = mm / s2; }
    float e1[3], e2[3];
    if (!sym_eigvec(E, s1, e1) || !sym_eigvec(E, s2, e2)) return false;
    float sr = sqrtf(-s2 / s1);
    float asum = a12 + a13 + a23;

    const float X4 = x[3][0], Y4 = x[3][1], Z4 = x[3][2], u4 = uv[3][0], v4 = uv[3][1];
    bool found = false;
    float beste = 0.0f;

#pragma unroll 1
    for (int sgi = 0; sgi < 2; ++sgi) {
        float sg = sgi == 0 ? sr : -sr;
        float n[3] = {fmaf(sg, e2[0], e1[0]), fmaf(sg, e2[1], e1[1]), fmaf(sg, e2[2], e1[2])};
        float nn = dot3(n, n);
        if (!(nn > 0.0f)) continue;
        float inn = 1.0f / sqrtf(nn);
        n[0] *= inn; n[1] *= inn; n[2] *= inn;
        float sz = copysignf(1.0f, n[2]);
        float aa = -1.0f / (sz + n[2]);
        float bb = n[0] * n[1] * aa;
        float p[3] = {fmaf(sz * n[0], n[0] * aa, 1.0f), sz * bb, -(sz * n[0])};
        float q[3] = {bb, fmaf(n[1], n[1] * aa, sz), -n[1]};
        float qa = sym_quad(Q, p, p), qb = sym_quad(Q, p, q), qc = sym_quad(Q, q, q);
        float dsc = fmaf(qb, qb, -(qa * qc));
        if (!(dsc >= 0.0f)) continue;
        float sd = sqrtf(dsc);
        bool use_a = fabsf(qa) >= fabsf(qc);
        float den = use_a ? qa : qc;
        if (den == 0.0f) continue;
#pragma unroll 1
        for (int ri = 0; ri < 2; ++ri) {
            float r = ((ri == 0 ? sd : -sd) - qb) / den;
            float L[3];
            if (use_a) { L[0] = fmaf(r, p[0], q[0]); L[1] = fmaf(r, p[1], q[1]); L[2] = fmaf(r, p[2], q[2]); }
            else       { L[0] = fmaf(r, q[0], p[0]); L[1] = fmaf(r, q[1], p[1]); L[2] = fmaf(r, q[2], p[2]); }
            bool pos = (L[0] > 0.0f) && (L[1] > 0.0f) && (L[2] > 0.0f);
            bool neg = (L[0] < 0.0f) && (L[1] < 0.0f) && (L[2] < 0.0f);
            if (!pos && !neg) continue;
            float sq_sum = fmaf(L[0], L[0], fmaf(L[1], L[1], L[2] * L[2]));
            float cr_sum = fmaf(b12 * L[0], L[1], fmaf(b13 * L[0], L[2], (b23 * L[1]) * L[2]));
            float ms = 2.0f * (sq_sum - cr_sum);
            if (!(ms > 0.0f)) continue;
            float k = sqrtf(asum / ms);
            if (neg) k = -k;
            L[0] *= k; L[1] *= k; L[2] *= k;
#pragma unroll 1
            for (int it = 0; it < 3; ++it) {
                float f0 = fmaf(L[0], L[0], fmaf(L[1], L[1], -(2.0f * b12) * (L[0] * L[1]))) - a12;
                float f1 = fmaf(L[0], L[0], fmaf(L[2], L[2], -(2.0f * b13) * (L[0] * L[2]))) - a13;
                float f2 = fmaf(L[1], L[1], fmaf(L[2], L[2], -(2.0f * b23) * (L[1] * L[2]))) - a23;
                float j00 = 2.0f * fmaf(-b12, L[1], L[0]), j01 = 2.0f * fmaf(-b12, L[0], L[1]);
                float j10 = 2.0f * fmaf(-b13, L[2], L[0]), j12 = 2.0f * fmaf(-b13, L[0], L[2]);
                float j21 = 2.0f * fmaf(-b23, L[2], L[1]), j22 = 2.0f * fmaf(-b23, L[1], L[2]);
                float det = -(fmaf(j00 * j12, j21, (j01 * j10) * j22));
                if (!(fabsf(det) > 1e-30f)) break;
                float idet = 1.0f / det;
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
            float R[9], t[3];
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j)
                    R[3 * i + j] = fmaf(y12[i], Xinv[j], fmaf(y13[i], Xinv[3 + j], yc[i] * Xinv[6 + j]));
#pragma unroll
            for (int i = 0; i < 3; ++i)
                t[i] = p0[i] - fmaf(R[3 * i], x[0][0], fmaf(R[3 * i + 1], x[0][1], R[3 * i + 2] * x[0][2]));
            bool ok = fin(t[0]) && fin(t[1]) && fin(t[2]);
#pragma unroll
            for (int i = 0; i < 9; ++i) ok = ok && fin(R[i]);
            if (!ok) continue;
            // fourth-point test
            float xcam = fmaf(R[0], X4, fmaf(R[1], Y4, fmaf(R[2], Z4, t[0])));
            float ycam = fmaf(R[3], X4, fmaf(R[4], Y4, fmaf(R[5], Z4, t[1])));
            float zcam = fmaf(R[6], X4, fmaf(R[7], Y4, fmaf(R[8], Z4, t[2])));
            if (!(zcam > 0.0f)) continue;
            float du = fmaf(fx, xcam / zcam, cx) - u4;
            float dv = fmaf(fy, ycam / zcam, cy) - v4;
            float e = fmaf(du, du, dv * dv);
            if (!fin(e)) continue;
            if (!found || e < beste) {
                found = true; beste = e;
#pragma unroll
                for (int i = 0; i < 9; ++i) Rout[i] = R[i];
                tout[0] = t[0]; tout[1] = t[1]; tout[2] = t[2];
            }
        }
    }
    return found;
}

}  // namespace p3p
