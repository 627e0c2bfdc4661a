// bpnp.cu -- back-propagatable PnP (reference: lib/network/dnn/BPnP.py).
//
//   forward  (BPnP.py:24-51)   pose refinement on ALL correspondences from an initial pose: Levenberg-Marquardt
//                              on (angle-axis, t) minimising the squared reprojection error, FP64.  The initial pose
//                              is the caller's ini_pose or the RANSAC pose of pb_pnp_ransac_pts (host glue).
//   backward (BPnP.py:54-117)  implicit-function gradient.  With  p = K (R(w) z + t),  pi = p[0:2]/p[2],  S = p[2]
//                              the reference builds  f_j = sum_i -2 S_i (x_i - pi_i) . d pi_i/d y_j  (j = 0..5),
//                              takes the Jacobians of f wrt y, x, z, K by autograd (a double backward through the
//                              projection) and returns  g . (-J_fy^-1 J_f{x,z,K}).  Here the same quantities come in
//                              closed form from the first and second derivatives of the rotation formula (small
//                              forward-mode jets over the 3 angle-axis components, evaluated once per sample) and the
//                              quotient rule per point; no autograd, no per-sample Python loop.
//
// One CTA per batch sample; threads stride the sample's points; all sums in FP64 in a fixed order (lane-strided
// partial sums, xor butterfly, warps in index order), so results do not depend on anything but the inputs.
#include "common.cuh"

namespace {

constexpr int BP_T = 128;

// ---- second-order jet over the three angle-axis components ------------------------------------------------
struct J3 {
    double v, g[3], h[6];   // h: (0,0) (0,1) (0,2) (1,1) (1,2) (2,2)
};
__device__ __forceinline__ int hidx(int a, int b) {   // a <= b
    return a == 0 ? b : (a == 1 ? 2 + b : 5);
}
__device__ __forceinline__ J3 jconst(double c) {
    J3 r; r.v = c;
    for (int i = 0; i < 3; ++i) r.g[i] = 0.0;
    for (int i = 0; i < 6; ++i) r.h[i] = 0.0;
    return r;
}
__device__ __forceinline__ J3 jvar(double c, int k) {
    J3 r = jconst(c); r.g[k] = 1.0;
    return r;
}
__device__ __forceinline__ J3 operator+(const J3& a, const J3& b) {
    J3 r; r.v = a.v + b.v;
    for (int i = 0; i < 3; ++i) r.g[i] = a.g[i] + b.g[i];
    for (int i = 0; i < 6; ++i) r.h[i] = a.h[i] + b.h[i];
    return r;
}
__device__ __forceinline__ J3 operator-(const J3& a, const J3& b) {
    J3 r; r.v = a.v - b.v;
    for (int i = 0; i < 3; ++i) r.g[i] = a.g[i] - b.g[i];
    for (int i = 0; i < 6; ++i) r.h[i] = a.h[i] - b.h[i];
    return r;
}
__device__ __forceinline__ J3 operator-(const J3& a) {
    J3 r; r.v = -a.v;
    for (int i = 0; i < 3; ++i) r.g[i] = -a.g[i];
    for (int i = 0; i < 6; ++i) r.h[i] = -a.h[i];
    return r;
}
__device__ __forceinline__ J3 operator*(const J3& a, const J3& b) {
    J3 r; r.v = a.v * b.v;
    for (int i = 0; i < 3; ++i) r.g[i] = a.g[i] * b.v + a.v * b.g[i];
    for (int i = 0; i < 3; ++i)
        for (int k = i; k < 3; ++k)
            r.h[hidx(i, k)] = a.h[hidx(i, k)] * b.v + a.g[i] * b.g[k] + a.g[k] * b.g[i] + a.v * b.h[hidx(i, k)];
    return r;
}
// f(a) for a scalar function with value f0 and derivatives f1, f2 at a.v
__device__ __forceinline__ J3 jchain(const J3& a, double f0, double f1, double f2) {
    J3 r; r.v = f0;
    for (int i = 0; i < 3; ++i) r.g[i] = f1 * a.g[i];
    for (int i = 0; i < 3; ++i)
        for (int k = i; k < 3; ++k) r.h[hidx(i, k)] = f1 * a.h[hidx(i, k)] + f2 * a.g[i] * a.g[k];
    return r;
}
__device__ __forceinline__ J3 jrecip(const J3& a) {
    const double w = 1.0 / a.v;
    return jchain(a, w, -w * w, 2.0 * w * w * w);
}
__device__ __forceinline__ J3 jsqrt(const J3& a) {
    const double s = sqrt(a.v);
    return jchain(a, s, 0.5 / s, -0.25 / (s * a.v));
}

// Rotation matrix and its first / second derivatives wrt the angle-axis vector.
//   eps = 1e-6: the formula the reference differentiates (kornia angle_axis_to_rotation_matrix: axis = aa/(theta+eps),
//               first-order branch I + [aa]x when theta^2 <= eps);
//   eps = 0   : exact Rodrigues (what cv2 optimises over), with the series branch only below theta^2 = 1e-16.
// out[0..8] = R (row major), out[9+9k+e] = dR_e/dw_k, out[36+9*hidx(j,k)+e] = d2R_e/dw_j dw_k       (90 doubles)
__device__ void rotation_jets(const double w[3], double eps, double* out) {
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    const bool taylor = eps > 0.0 ? !(th2 > eps) : !(th2 > 1e-16);
    if (taylor) {
        for (int i = 0; i < 90; ++i) out[i] = 0.0;
        out[0] = out[4] = out[8] = 1.0;
        out[1] = -w[2]; out[2] = w[1]; out[3] = w[2]; out[5] = -w[0]; out[6] = -w[1]; out[7] = w[0];
        out[9 + 5] = -1.0; out[9 + 7] = 1.0;          // d/dw0
        out[18 + 2] = 1.0; out[18 + 6] = -1.0;        // d/dw1
        out[27 + 1] = -1.0; out[27 + 3] = 1.0;        // d/dw2
        return;
    }
    J3 a[3] = {jvar(w[0], 0), jvar(w[1], 1), jvar(w[2], 2)};
    J3 t2 = a[0] * a[0] + a[1] * a[1] + a[2] * a[2];
    J3 th = jsqrt(t2);
    J3 inv = jrecip(th + jconst(eps));
    J3 u[3] = {a[0] * inv, a[1] * inv, a[2] * inv};
    const double c0 = cos(th.v), s0 = sin(th.v);
    J3 c = jchain(th, c0, -s0, -c0);
    J3 s = jchain(th, s0, c0, -s0);
    J3 omc = jconst(1.0) - c;
    J3 R[9];
    R[0] = c + u[0] * u[0] * omc;
    R[3] = u[2] * s + u[0] * u[1] * omc;
    R[6] = -(u[1] * s) + u[0] * u[2] * omc;
    R[1] = u[0] * u[1] * omc - u[2] * s;
    R[4] = c + u[1] * u[1] * omc;
    R[7] = u[0] * s + u[1] * u[2] * omc;
    R[2] = u[1] * s + u[0] * u[2] * omc;
    R[5] = -(u[0] * s) + u[1] * u[2] * omc;
    R[8] = c + u[2] * u[2] * omc;
    for (int e = 0; e < 9; ++e) {
        out[e] = R[e].v;
        for (int k = 0; k < 3; ++k) out[9 + 9 * k + e] = R[e].g[k];
        for (int k = 0; k < 6; ++k) out[36 + 9 * k + e] = R[e].h[k];
    }
}

// sum of v[0..N) over the CTA, result in every thread; fixed order (butterfly inside a warp, warps in index order)
template <int N>
__device__ __forceinline__ void cta_allsum(double* v, double* red /* [BP_T/32][N] */) {
#pragma unroll
    for (int k = 0; k < N; ++k) {
        double x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x = x + __shfl_xor_sync(0xffffffffu, x, o);
        v[k] = x;
    }
    const int nw = blockDim.x >> 5;
    if (nw == 1) return;
    __syncthreads();
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < N; ++k) red[(threadIdx.x >> 5) * N + k] = v[k];
    __syncthreads();
    for (int k = 0; k < N; ++k) {
        double s = red[k];
        for (int w = 1; w < nw; ++w) s = s + red[w * N + k];
        v[k] = s;
    }
}

__device__ __forceinline__ void mat3vec(const double* M, const double v[3], double o[3]) {
    o[0] = M[0] * v[0] + M[1] * v[1] + M[2] * v[2];
    o[1] = M[3] * v[0] + M[4] * v[1] + M[5] * v[2];
    o[2] = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
}

// in-place solve of A x = b (6x6, row major), partial pivoting; false if singular
__device__ bool solve6(double A[36], double b[6]) {
    for (int c = 0; c < 6; ++c) {
        int piv = c; double best = fabs(A[c * 6 + c]);
        for (int r = c + 1; r < 6; ++r)
            if (fabs(A[r * 6 + c]) > best) { best = fabs(A[r * 6 + c]); piv = r; }
        if (!(best > 0.0) || !isfinite(best)) return false;
        if (piv != c) {
            for (int k = 0; k < 6; ++k) { double t = A[c * 6 + k]; A[c * 6 + k] = A[piv * 6 + k]; A[piv * 6 + k] = t; }
            double t = b[c]; b[c] = b[piv]; b[piv] = t;
        }
        const double inv = 1.0 / A[c * 6 + c];
        for (int r = c + 1; r < 6; ++r) {
            const double f = A[r * 6 + c] * inv;
            for (int k = c; k < 6; ++k) A[r * 6 + k] -= f * A[c * 6 + k];
            b[r] -= f * b[c];
        }
    }
    for (int c = 5; c >= 0; --c) {
        double s = b[c];
        for (int k = c + 1; k < 6; ++k) s -= A[c * 6 + k] * b[k];
        b[c] = s / A[c * 6 + c];
    }
    return true;
}

struct PointGeom {
    double pi[2], e[2], w, S;
    double dpi[6][2];      // d pi_a / d y_j
    double dp2[6];         // d p[2] / d y_j
};

// KJ (shared): [0..8] K R, [9+9k..] K dR_k, [36+9h..] K ddR_h; Kt = K t; Kd = K (row major)
__device__ __forceinline__ void point_first_order(const double* KJ, const double* Kd, const double Kt[3], const double z[3],
                                                  const double x[2], PointGeom& g, double dp[6][3]) {
    double p[3];
    mat3vec(KJ, z, p);
    p[0] += Kt[0]; p[1] += Kt[1]; p[2] += Kt[2];
    for (int j = 0; j < 3; ++j) mat3vec(KJ + 9 + 9 * j, z, dp[j]);
    for (int j = 0; j < 3; ++j) { dp[3 + j][0] = Kd[j]; dp[3 + j][1] = Kd[3 + j]; dp[3 + j][2] = Kd[6 + j]; }
    g.w = 1.0 / p[2];
    g.S = p[2];
    g.pi[0] = p[0] * g.w; g.pi[1] = p[1] * g.w;
    g.e[0] = x[0] - g.pi[0]; g.e[1] = x[1] - g.pi[1];
    for (int j = 0; j < 6; ++j) {
        g.dp2[j] = dp[j][2];
        g.dpi[j][0] = (dp[j][0] - g.pi[0] * dp[j][2]) * g.w;
        g.dpi[j][1] = (dp[j][1] - g.pi[1] * dp[j][2]) * g.w;
    }
}

struct BpnpBwdArgs {
    const float* x;       // [B,n,2]
    const float* z;       // [n,3] or [B,n,3]
    const float* K;       // [9]
    const float* pose;    // [B,6]
    const float* gout;    // [B,6]
    int n, z_batched;
    float* grad_x;        // [B,n,2]
    float* grad_z_b;      // [B,n,3] (z_batched) or null
    double* gz_part;      // [B,n,3] (shared z) or null
    double* gK_part;      // [B,9]
    int32_t* status;      // [B]
};

__global__ void __launch_bounds__(BP_T) bpnp_backward_kernel(BpnpBwdArgs a) {
    __shared__ double RJ[90], KJ[90], Kd[9], y[6], Kt[3], red[(BP_T / 32) * 36], vsh[6], KRv[9], Rv[9];
    __shared__ int ok_sh;
    const int b = blockIdx.x, n = a.n, tid = threadIdx.x;
    if (tid < 9) Kd[tid] = (double)a.K[tid];
    if (tid < 6) y[tid] = (double)a.pose[(size_t)b * 6 + tid];
    __syncthreads();
    if (tid == 0) rotation_jets(y, 1e-6, RJ);
    __syncthreads();
    for (int m = tid; m < 10 * 9; m += BP_T) {           // K times each of the 10 matrices
        const int mat = m / 9, e = m % 9, r = e / 3, c = e % 3;
        const double* M = RJ + 9 * mat;
        KJ[m] = Kd[r * 3] * M[c] + Kd[r * 3 + 1] * M[3 + c] + Kd[r * 3 + 2] * M[6 + c];
    }
    if (tid < 3) Kt[tid] = Kd[tid * 3] * y[3] + Kd[tid * 3 + 1] * y[4] + Kd[tid * 3 + 2] * y[5];
    __syncthreads();
    const float* xb = a.x + (size_t)b * n * 2;
    const float* zb = a.z + (a.z_batched ? (size_t)b * n * 3 : 0);

    // ---- pass 1: J_fy ------------------------------------------------------------------------------------
    double J[36];
    for (int k = 0; k < 36; ++k) J[k] = 0.0;
    for (int i = tid; i < n; i += BP_T) {
        const double z[3] = {(double)zb[3 * i], (double)zb[3 * i + 1], (double)zb[3 * i + 2]};
        const double x[2] = {(double)xb[2 * i], (double)xb[2 * i + 1]};
        PointGeom g; double dp[6][3];
        point_first_order(KJ, Kd, Kt, z, x, g, dp);
#pragma unroll
        for (int j = 0; j < 6; ++j)
#pragma unroll
            for (int k = 0; k < 6; ++k) {
                double ddp[3] = {0.0, 0.0, 0.0};
                if (j < 3 && k < 3) mat3vec(KJ + 36 + 9 * (j <= k ? hidx(j, k) : hidx(k, j)), z, ddp);
                double acc = 0.0;
                for (int c = 0; c < 2; ++c) {
                    const double d2 = (ddp[c] - g.dpi[k][c] * g.dp2[j] - g.pi[c] * ddp[2]) * g.w - g.dpi[j][c] * g.dp2[k] * g.w;
                    acc += g.dp2[k] * g.e[c] * g.dpi[j][c] - g.S * g.dpi[k][c] * g.dpi[j][c] + g.S * g.e[c] * d2;
                }
                J[j * 6 + k] += -2.0 * acc;
            }
    }
    cta_allsum<36>(J, red);
    if (tid == 0) {
        double A[36], rhs[6];
        for (int r = 0; r < 6; ++r) {
            for (int c = 0; c < 6; ++c) A[r * 6 + c] = J[c * 6 + r];      // J_fy^T v = -g
            rhs[r] = -(double)a.gout[(size_t)b * 6 + r];
        }
        const bool ok = solve6(A, rhs);
        ok_sh = ok ? 1 : 0;
        for (int r = 0; r < 6; ++r) vsh[r] = ok ? rhs[r] : 0.0;
        a.status[b] = ok ? 0 : 1;
    }
    __syncthreads();
    // direction v in pose space: Rv = sum_j v_j dR_j, K Rv
    if (tid < 9) {
        Rv[tid] = vsh[0] * RJ[9 + tid] + vsh[1] * RJ[18 + tid] + vsh[2] * RJ[27 + tid];
        KRv[tid] = vsh[0] * KJ[9 + tid] + vsh[1] * KJ[18 + tid] + vsh[2] * KJ[27 + tid];
    }
    __syncthreads();

    // ---- pass 2: v . J_fx, v . J_fz, v . J_fK -------------------------------------------------------------
    double gK[9];
    for (int k = 0; k < 9; ++k) gK[k] = 0.0;
    for (int i = tid; i < n; i += BP_T) {
        const double z[3] = {(double)zb[3 * i], (double)zb[3 * i + 1], (double)zb[3 * i + 2]};
        const double x[2] = {(double)xb[2 * i], (double)xb[2 * i + 1]};
        PointGeom g; double dp[6][3];
        point_first_order(KJ, Kd, Kt, z, x, g, dp);
        double q[3], dqv[3], dpv[3];
        mat3vec(RJ, z, q);
        q[0] += y[3]; q[1] += y[4]; q[2] += y[5];
        mat3vec(Rv, z, dqv);
        dqv[0] += vsh[3]; dqv[1] += vsh[4]; dqv[2] += vsh[5];
        mat3vec(Kd, dqv, dpv);
        const double piv[2] = {(dpv[0] - g.pi[0] * dpv[2]) * g.w, (dpv[1] - g.pi[1] * dpv[2]) * g.w};
        a.grad_x[((size_t)b * n + i) * 2] = (float)(-2.0 * g.S * piv[0]);
        a.grad_x[((size_t)b * n + i) * 2 + 1] = (float)(-2.0 * g.S * piv[1]);
        // derivative of F = sum_j v_j f_j in a direction m with first-order dp_m and mixed second-order ddp_vm
        auto dirF = [&](const double dpm[3], const double ddpvm[3]) {
            double acc = 0.0;
            for (int c = 0; c < 2; ++c) {
                const double pim = (dpm[c] - g.pi[c] * dpm[2]) * g.w;
                const double pivm = (ddpvm[c] - pim * dpv[2] - g.pi[c] * ddpvm[2]) * g.w - piv[c] * dpm[2] * g.w;
                acc += dpm[2] * g.e[c] * piv[c] - g.S * pim * piv[c] + g.S * g.e[c] * pivm;
            }
            return -2.0 * acc;
        };
        double gz[3];
        for (int c = 0; c < 3; ++c) {
            const double dpm[3] = {KJ[c], KJ[3 + c], KJ[6 + c]};
            const double dd[3] = {KRv[c], KRv[3 + c], KRv[6 + c]};
            gz[c] = dirF(dpm, dd);
        }
        if (a.z_batched)
            for (int c = 0; c < 3; ++c) a.grad_z_b[((size_t)b * n + i) * 3 + c] = (float)gz[c];
        else
            for (int c = 0; c < 3; ++c) a.gz_part[((size_t)b * n + i) * 3 + c] = gz[c];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) {
                double dpm[3] = {0.0, 0.0, 0.0}, dd[3] = {0.0, 0.0, 0.0};
                dpm[r] = q[c]; dd[r] = dqv[c];
                gK[r * 3 + c] += dirF(dpm, dd);
            }
    }
    cta_allsum<9>(gK, red);
    if (tid < 9) a.gK_part[(size_t)b * 9 + tid] = ok_sh ? gK[tid] : 0.0;
}

// sums the per-sample partial gradients over the batch in sample order
__global__ void bpnp_reduce_kernel(const double* __restrict__ gz_part, const double* __restrict__ gK_part, int B, int n3,
                                   float* __restrict__ grad_z, float* __restrict__ grad_K) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (gz_part && i < n3) {
        double s = 0.0;
        for (int b = 0; b < B; ++b) s = s + gz_part[(size_t)b * n3 + i];
        grad_z[i] = (float)s;
    }
    if (i < 9) {
        double s = 0.0;
        for (int b = 0; b < B; ++b) s = s + gK_part[(size_t)b * 9 + i];
        grad_K[i] = (float)s;
    }
}

// rotation matrix (row major) -> angle-axis vector, FP64
__device__ void rotation_log(const double R[9], double w[3]) {
    const double ax = 0.5 * (R[7] - R[5]), ay = 0.5 * (R[2] - R[6]), az = 0.5 * (R[3] - R[1]);
    const double s = sqrt(ax * ax + ay * ay + az * az);
    double c = 0.5 * (R[0] + R[4] + R[8] - 1.0);
    c = fmin(1.0, fmax(-1.0, c));
    const double th = atan2(s, c);
    if (s > 1e-7 && c > -0.99) {
        const double f = th / s;
        w[0] = ax * f; w[1] = ay * f; w[2] = az * f;
    } else if (c > 0.0) {                       // theta ~ 0: sin(theta)/theta -> 1
        w[0] = ax; w[1] = ay; w[2] = az;
    } else {                                    // theta near pi: axis from the symmetric part
        const double d0 = fmax(0.0, (R[0] - c) / (1.0 - c)), d1 = fmax(0.0, (R[4] - c) / (1.0 - c)),
                     d2 = fmax(0.0, (R[8] - c) / (1.0 - c));
        double u[3] = {sqrt(d0), sqrt(d1), sqrt(d2)};
        const int m = (u[0] >= u[1] && u[0] >= u[2]) ? 0 : (u[1] >= u[2] ? 1 : 2);
        const double s01 = R[1] + R[3], s02 = R[2] + R[6], s12 = R[5] + R[7];
        if (m == 0) { if (s01 < 0) u[1] = -u[1]; if (s02 < 0) u[2] = -u[2]; }
        else if (m == 1) { if (s01 < 0) u[0] = -u[0]; if (s12 < 0) u[2] = -u[2]; }
        else { if (s02 < 0) u[0] = -u[0]; if (s12 < 0) u[1] = -u[1]; }
        const double a3[3] = {ax, ay, az};
        if (a3[0] * u[0] + a3[1] * u[1] + a3[2] * u[2] < 0) { u[0] = -u[0]; u[1] = -u[1]; u[2] = -u[2]; }
        const double nn = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
        for (int k = 0; k < 3; ++k) w[k] = nn > 0 ? u[k] / nn * th : 0.0;
    }
}

struct RefineArgs {
    const float* x;        // [B,n,2]
    const float* z;        // [n,3] or [B,n,3]
    const float* K;        // [9]
    const float* pose0;    // [B,6] or null
    const float* R0;       // [B,9] (used when pose0 is null)
    const float* t0;       // [B,3]
    int n, z_batched, max_iter;
    float* pose;           // [B,6]
    float* rmse;           // [B] or null
    int32_t* status;       // [B]
};

// normal equations of the reprojection error at pose y: acc[0..20] upper triangle of J^T J, acc[21..26] J^T e, acc[27] cost
__device__ __forceinline__ void refine_accumulate(const float* xb, const float* zb, int n, const double* KJ, const double* Kd,
                                                  const double Kt[3], double acc[28]) {
    for (int k = 0; k < 28; ++k) acc[k] = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const double z[3] = {(double)zb[3 * i], (double)zb[3 * i + 1], (double)zb[3 * i + 2]};
        const double x[2] = {(double)xb[2 * i], (double)xb[2 * i + 1]};
        PointGeom g; double dp[6][3];
        point_first_order(KJ, Kd, Kt, z, x, g, dp);
        int q = 0;
#pragma unroll
        for (int j = 0; j < 6; ++j) {
#pragma unroll
            for (int k = j; k < 6; ++k) acc[q++] += g.dpi[j][0] * g.dpi[k][0] + g.dpi[j][1] * g.dpi[k][1];
            acc[21 + j] += g.dpi[j][0] * g.e[0] + g.dpi[j][1] * g.e[1];
        }
        acc[27] += g.e[0] * g.e[0] + g.e[1] * g.e[1];
    }
}

__global__ void __launch_bounds__(BP_T) bpnp_refine_kernel(RefineArgs a) {
    __shared__ double RJ[90], KJ[36], Kd[9], Kt[3], red[(BP_T / 32) * 28], ysh[6];
    const int b = blockIdx.x, n = a.n, tid = threadIdx.x;
    if (tid < 9) Kd[tid] = (double)a.K[tid];
    if (tid == 0) {
        if (a.pose0) {
            for (int k = 0; k < 6; ++k) ysh[k] = (double)a.pose0[(size_t)b * 6 + k];
        } else {
            double R[9];
            for (int k = 0; k < 9; ++k) R[k] = (double)a.R0[(size_t)b * 9 + k];
            rotation_log(R, ysh);
            for (int k = 0; k < 3; ++k) ysh[3 + k] = (double)a.t0[(size_t)b * 3 + k];
        }
    }
    __syncthreads();
    const float* xb = a.x + (size_t)b * n * 2;
    const float* zb = a.z + (a.z_batched ? (size_t)b * n * 3 : 0);
    double y[6];
    for (int k = 0; k < 6; ++k) y[k] = ysh[k];

    auto evaluate = [&](const double yy[6], double acc[28]) {
        __syncthreads();
        if (tid == 0) rotation_jets(yy, 0.0, RJ);
        __syncthreads();
        for (int m = tid; m < 36; m += blockDim.x) {
            const int mat = m / 9, e = m % 9, r = e / 3, c = e % 3;
            const double* M = RJ + 9 * mat;
            KJ[m] = Kd[r * 3] * M[c] + Kd[r * 3 + 1] * M[3 + c] + Kd[r * 3 + 2] * M[6 + c];
        }
        if (tid < 3) Kt[tid] = Kd[tid * 3] * yy[3] + Kd[tid * 3 + 1] * yy[4] + Kd[tid * 3 + 2] * yy[5];
        __syncthreads();
        refine_accumulate(xb, zb, n, KJ, Kd, Kt, acc);
        cta_allsum<28>(acc, red);
    };

    double cur[28];
    evaluate(y, cur);
    int st = 1;                                  // 1 = iteration limit reached
    if (!isfinite(cur[27])) st = 2;
    double lambda = 1e-3;
    for (int it = 0; it < a.max_iter && st == 1; ++it) {
        double A[36], d[6];
        int q = 0;
        for (int j = 0; j < 6; ++j)
            for (int k = j; k < 6; ++k) { A[j * 6 + k] = cur[q]; A[k * 6 + j] = cur[q]; ++q; }
        for (int j = 0; j < 6; ++j) { A[j * 6 + j] += lambda * A[j * 6 + j] + 1e-300; d[j] = cur[21 + j]; }
        if (!solve6(A, d)) { st = 2; break; }
        double yn[6], step = 0.0;
        for (int k = 0; k < 6; ++k) { yn[k] = y[k] + d[k]; step = fmax(step, fabs(d[k])); }
        double nxt[28];
        evaluate(yn, nxt);
        if (isfinite(nxt[27]) && nxt[27] <= cur[27]) {
            const double drop = cur[27] - nxt[27];
            for (int k = 0; k < 6; ++k) y[k] = yn[k];
            for (int k = 0; k < 28; ++k) cur[k] = nxt[k];
            lambda = fmax(lambda * 0.1, 1e-15);
            if (step < 1e-12 || drop <= 1e-16 * cur[27]) st = 0;
        } else {
            lambda *= 10.0;
            if (lambda > 1e12) st = 0;           // no descent direction left: at the minimum to working precision
        }
    }
    if (tid == 0) {
        for (int k = 0; k < 6; ++k) a.pose[(size_t)b * 6 + k] = (float)y[k];
        if (a.rmse) a.rmse[b] = (float)sqrt(cur[27] / (double)n);
        a.status[b] = st;
    }
}

}  // namespace

extern "C" size_t pb_bpnp_backward_workspace_bytes(int B, int n) {
    if (B < 1 || n < 1) return 0;
    return ((size_t)B * n * 3 + (size_t)B * 9) * sizeof(double) + 256;
}

extern "C" int pb_bpnp_backward(const float* pts2d, const float* pts3d, int pts3d_batched, const float* K, const float* pose6,
                                const float* grad_out, int B, int n, float* grad_pts2d, float* grad_pts3d, float* grad_K,
                                int32_t* status, void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(pts2d && pts3d && K && pose6 && grad_out && grad_pts2d && grad_pts3d && grad_K && status,
               "pb_bpnp_backward: null pointer");
    PB_REQUIRE(B >= 1 && n >= 3, "pb_bpnp_backward: need B >= 1 and n >= 3 (B=%d n=%d)", B, n);
    PB_REQUIRE(ws && ws_bytes >= pb_bpnp_backward_workspace_bytes(B, n) && ((uintptr_t)ws & 7) == 0,
               "pb_bpnp_backward: workspace too small or misaligned (%zu bytes)", ws_bytes);
    cudaStream_t st = (cudaStream_t)stream;
    BpnpBwdArgs a;
    a.x = pts2d; a.z = pts3d; a.K = K; a.pose = pose6; a.gout = grad_out; a.n = n; a.z_batched = pts3d_batched ? 1 : 0;
    a.grad_x = grad_pts2d;
    a.gK_part = (double*)ws;
    a.gz_part = pts3d_batched ? nullptr : a.gK_part + (size_t)B * 9;
    a.grad_z_b = pts3d_batched ? grad_pts3d : nullptr;
    a.status = status;
    bpnp_backward_kernel<<<B, BP_T, 0, st>>>(a);
    const int n3 = n * 3, work = n3 > 9 ? n3 : 9;
    bpnp_reduce_kernel<<<(work + 255) / 256, 256, 0, st>>>(a.gz_part, a.gK_part, B, n3, grad_pts3d, grad_K);
    return pb_check_launch("pb_bpnp_backward");
}

extern "C" int pb_pnp_refine(const float* pts2d, const float* pts3d, int pts3d_batched, const float* K, const float* pose6_init,
                             const float* R_init, const float* t_init, int B, int n, int max_iter, float* pose6, float* rmse,
                             int32_t* status, void* stream) {
    PB_REQUIRE(pts2d && pts3d && K && pose6 && status, "pb_pnp_refine: null pointer");
    PB_REQUIRE(pose6_init || (R_init && t_init), "pb_pnp_refine: need pose6_init or R_init + t_init");
    PB_REQUIRE(B >= 1 && n >= 3 && max_iter >= 0, "pb_pnp_refine: need B >= 1, n >= 3, max_iter >= 0 (B=%d n=%d)", B, n);
    RefineArgs a;
    a.x = pts2d; a.z = pts3d; a.K = K; a.pose0 = pose6_init; a.R0 = R_init; a.t0 = t_init; a.n = n;
    a.z_batched = pts3d_batched ? 1 : 0; a.max_iter = max_iter; a.pose = pose6; a.rmse = rmse; a.status = status;
    const int threads = n <= 64 ? 32 : BP_T;
    bpnp_refine_kernel<<<B, threads, 0, (cudaStream_t)stream>>>(a);
    return pb_check_launch("pb_pnp_refine");
}
