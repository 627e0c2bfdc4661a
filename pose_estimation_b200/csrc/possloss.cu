// possloss.cu -- per-hypothesis pose loss of the transparent-object branch (reference:
// version/transparent/lib/networks/loss.py:197-267 PossLoss, lib/transform/allocentric.py:15-79).
//
// H = bs * num_p pose hypotheses (quaternion, translation, confidence) against one model cloud per batch element:
//   R_h   = allo_to_ego(t_h) * R(q_h / |q_h|)                                   (loss.py:209-212)
//   dis_h = mean_i | R_h m_i + t_h - target_{nn(i)} |      nn = identity, or for symmetric classes (not refine) the
//           nearest of target[0]'s points: the 250 000 x 500 KeOps argmin of loss.py:230, here fused with the
//           transform and the distance so the [H*N,3] cloud and the index list never exist in HBM
//   loss  = mean_h (dis_h c_h - w log(c_h + 1e-8)),  loss_rot likewise over the per-axis cosine term  (:239-247)
// forward  = one CTA per hypothesis (target tile in shared memory, four query points per thread in flight) + a
// one-CTA finalize; backward = one CTA per hypothesis: closed-form d dis / d (R, t) summed over the points, then
// the chain through the rotation construction by first-order forward-mode duals over (q, t) in FP64.
#include "common.cuh"

namespace {

constexpr int PL_T = 128;
constexpr int PL_Q = 4;             // query points per thread processed together
constexpr int PL_TILE = 2048;       // target points per shared-memory tile

// ---- first-order dual over the 7 inputs (q0..q3, t0..t2) ---------------------------------------------------
struct D7 {
    double v, g[7];
};
__device__ __forceinline__ D7 dmake(double v) { D7 r; r.v = v; for (int i = 0; i < 7; ++i) r.g[i] = 0.0; return r; }
__device__ __forceinline__ D7 dvar(double v, int k) { D7 r = dmake(v); r.g[k] = 1.0; return r; }
__device__ __forceinline__ D7 operator+(const D7& a, const D7& b) { D7 r; r.v = a.v + b.v; for (int i = 0; i < 7; ++i) r.g[i] = a.g[i] + b.g[i]; return r; }
__device__ __forceinline__ D7 operator-(const D7& a, const D7& b) { D7 r; r.v = a.v - b.v; for (int i = 0; i < 7; ++i) r.g[i] = a.g[i] - b.g[i]; return r; }
__device__ __forceinline__ D7 operator-(const D7& a) { D7 r; r.v = -a.v; for (int i = 0; i < 7; ++i) r.g[i] = -a.g[i]; return r; }
__device__ __forceinline__ D7 operator*(const D7& a, const D7& b) { D7 r; r.v = a.v * b.v; for (int i = 0; i < 7; ++i) r.g[i] = a.g[i] * b.v + a.v * b.g[i]; return r; }
__device__ __forceinline__ D7 operator*(double s, const D7& a) { D7 r; r.v = s * a.v; for (int i = 0; i < 7; ++i) r.g[i] = s * a.g[i]; return r; }
__device__ __forceinline__ D7 operator+(const D7& a, double s) { D7 r = a; r.v += s; return r; }
__device__ __forceinline__ D7 operator-(double s, const D7& a) { D7 r = -a; r.v += s; return r; }
__device__ __forceinline__ D7 dfun(const D7& a, double f0, double f1) { D7 r; r.v = f0; for (int i = 0; i < 7; ++i) r.g[i] = f1 * a.g[i]; return r; }
__device__ __forceinline__ D7 operator/(const D7& a, const D7& b) { const double w = 1.0 / b.v; return a * dfun(b, w, -w * w); }
__device__ __forceinline__ D7 xsqrt(const D7& a) { const double s = sqrt(a.v); return dfun(a, s, s > 0.0 ? 0.5 / s : 0.0); }
__device__ __forceinline__ D7 xacos(const D7& a) { return dfun(a, acos(a.v), -1.0 / sqrt(fmax(1.0 - a.v * a.v, 1e-300))); }
__device__ __forceinline__ D7 xsin(const D7& a) { return dfun(a, sin(a.v), cos(a.v)); }
__device__ __forceinline__ D7 xcos(const D7& a) { return dfun(a, cos(a.v), -sin(a.v)); }
__device__ __forceinline__ D7 xmaxc(const D7& a, double c) { return a.v >= c ? a : dmake(c); }
__device__ __forceinline__ double xsqrt(double a) { return sqrt(a); }
__device__ __forceinline__ double xacos(double a) { return acos(a); }
__device__ __forceinline__ double xsin(double a) { return sin(a); }
__device__ __forceinline__ double xcos(double a) { return cos(a); }
__device__ __forceinline__ double xmaxc(double a, double c) { return a >= c ? a : c; }

// quaternion (w,x,y,z), already normalised -> matrix, the reference's two spellings
template <typename T>
__device__ void quat_to_mat_pred(const T r[4], T M[9]) {          // loss.py:255-265
    M[0] = 1.0 - 2.0 * (r[2] * r[2] + r[3] * r[3]);
    M[1] = 2.0 * (r[1] * r[2]) - 2.0 * (r[0] * r[3]);
    M[2] = 2.0 * (r[0] * r[2]) + 2.0 * (r[1] * r[3]);
    M[3] = 2.0 * (r[1] * r[2]) + 2.0 * (r[3] * r[0]);
    M[4] = 1.0 - 2.0 * (r[1] * r[1] + r[3] * r[3]);
    M[5] = 2.0 * (r[2] * r[3]) - 2.0 * (r[0] * r[1]);
    M[6] = 2.0 * (r[1] * r[3]) - 2.0 * (r[0] * r[2]);
    M[7] = 2.0 * (r[0] * r[1]) + 2.0 * (r[2] * r[3]);
    M[8] = 1.0 - 2.0 * (r[1] * r[1] + r[2] * r[2]);
}

// R_ego = allo_to_ego(t) * R(q/|q|)   (loss.py:209-212, allocentric.py:15-45); la = sum_k axis_k (1 - cos(col_k R, col_k Rt))
template <typename T>
__device__ void ego_rotation(const T q[4], const T t[3], const double* target_r, const double* axis, T R[9], T& la) {
    const double eps = 1e-4;
    T qn = xsqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    T r[4] = {q[0] / qn, q[1] / qn, q[2] / qn, q[3] / qn};
    T Ra[9];
    quat_to_mat_pred(r, Ra);
    T tn = xsqrt(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]) + eps;
    T o[3] = {t[0] / tn, t[1] / tn, t[2] / tn};
    T ang = xacos(o[2]);
    T ax[3] = {-o[1], o[0], 0.0 * o[0]};                           // (0,0,1) x obj_ray
    T an = xsqrt(ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2]) + eps;
    T half = 0.5 * ang;
    T sh = xsin(half);
    T qa[4] = {xcos(half), ax[0] / an * sh, ax[1] / an * sh, ax[2] / an * sh};
    T qan = xsqrt(qa[0] * qa[0] + qa[1] * qa[1] + qa[2] * qa[2] + qa[3] * qa[3]);   // quat2mat_torch normalises again
    T w = qa[0] / qan, x = qa[1] / qan, y = qa[2] / qan, z = qa[3] / qan;
    T X = 2.0 * x, Y = 2.0 * y, Z = 2.0 * z;
    T A[9] = {1.0 - (y * Y + z * Z), x * Y - w * Z, x * Z + w * Y,
              x * Y + w * Z, 1.0 - (x * X + z * Z), y * Z - w * X,
              x * Z - w * Y, y * Z + w * X, 1.0 - (x * X + y * Y)};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R[i * 3 + j] = A[i * 3] * Ra[j] + A[i * 3 + 1] * Ra[3 + j] + A[i * 3 + 2] * Ra[6 + j];
    la = 0.0 * R[0];
    for (int k = 0; k < 3; ++k) {
        const double b0 = target_r[k], b1 = target_r[3 + k], b2 = target_r[6 + k];
        const double bn = fmax(sqrt(b0 * b0 + b1 * b1 + b2 * b2), 1e-6);
        T dot = b0 * R[k] + b1 * R[3 + k] + b2 * R[6 + k];
        T nn = xmaxc(xsqrt(R[k] * R[k] + R[3 + k] * R[3 + k] + R[6 + k] * R[6 + k]), 1e-6);
        la = la + axis[k] * (1.0 - dot / (bn * nn));
    }
}

struct PossArgs {
    const float* pred_r;   // [H,4]
    const float* pred_t;   // [H,3]
    const float* pred_c;   // [H]
    const float* model;    // [bs,N,3]
    const float* target;   // [bs,N,3]
    int bs, num_p, N, nn_mode;
    float w;
    const float* axis;     // [3]  (device)
    const float* target_r; // [9]  (device)
    float* dis;            // [H]       (workspace)
    float* la;             // [H]
    int32_t* which;        // [1]
    int32_t* nn;           // [H,N]     (nn_mode only)
};

__device__ __forceinline__ void load_pose(const PossArgs& a, int h, double q[4], double t[3], double tr[9], double ax[3]) {
    for (int k = 0; k < 4; ++k) q[k] = (double)a.pred_r[(size_t)h * 4 + k];
    for (int k = 0; k < 3; ++k) t[k] = (double)a.pred_t[(size_t)h * 3 + k];
    for (int k = 0; k < 9; ++k) tr[k] = (double)a.target_r[k];
    for (int k = 0; k < 3; ++k) ax[k] = (double)a.axis[k];
}

__global__ void __launch_bounds__(PL_T) poss_fwd_kernel(PossArgs a) {
    extern __shared__ float4 tile[];
    __shared__ float Rs[12];
    __shared__ double red[PL_T / 32];
    const int h = blockIdx.x, tid = threadIdx.x, N = a.N, b = h / a.num_p;
    if (tid == 0) {
        double q[4], t[3], tr[9], ax[3], R[9], la;
        load_pose(a, h, q, t, tr, ax);
        ego_rotation<double>(q, t, tr, ax, R, la);
        for (int k = 0; k < 9; ++k) Rs[k] = (float)R[k];
        for (int k = 0; k < 3; ++k) Rs[9 + k] = a.pred_t[(size_t)h * 3 + k];
        a.la[h] = (float)la;
    }
    __syncthreads();
    const float* mb = a.model + (size_t)b * N * 3;
    float sum = 0.f;
    if (!a.nn_mode) {
        const float* tb = a.target + (size_t)b * N * 3;
        for (int i = tid; i < N; i += PL_T) {
            const float mx = mb[3 * i], my = mb[3 * i + 1], mz = mb[3 * i + 2];
            const float dx = fmaf(Rs[0], mx, fmaf(Rs[1], my, fmaf(Rs[2], mz, Rs[9]))) - tb[3 * i];
            const float dy = fmaf(Rs[3], mx, fmaf(Rs[4], my, fmaf(Rs[5], mz, Rs[10]))) - tb[3 * i + 1];
            const float dz = fmaf(Rs[6], mx, fmaf(Rs[7], my, fmaf(Rs[8], mz, Rs[11]))) - tb[3 * i + 2];
            sum += sqrtf(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
        }
    } else {
        // groups of PL_Q * PL_T query points; every group sweeps the target tiles
        for (int g0 = 0; g0 < N; g0 += PL_Q * PL_T) {
            float px[PL_Q], py[PL_Q], pz[PL_Q], best[PL_Q];
            int bi[PL_Q];
#pragma unroll
            for (int k = 0; k < PL_Q; ++k) {
                const int i = g0 + k * PL_T + tid;
                const int ii = i < N ? i : N - 1;
                const float mx = mb[3 * ii], my = mb[3 * ii + 1], mz = mb[3 * ii + 2];
                px[k] = fmaf(Rs[0], mx, fmaf(Rs[1], my, fmaf(Rs[2], mz, Rs[9])));
                py[k] = fmaf(Rs[3], mx, fmaf(Rs[4], my, fmaf(Rs[5], mz, Rs[10])));
                pz[k] = fmaf(Rs[6], mx, fmaf(Rs[7], my, fmaf(Rs[8], mz, Rs[11])));
                best[k] = INFINITY; bi[k] = 0;
            }
            for (int j0 = 0; j0 < N; j0 += PL_TILE) {
                const int nt = min(PL_TILE, N - j0);
                __syncthreads();
                for (int j = tid; j < nt; j += PL_T)
                    tile[j] = make_float4(a.target[3 * (j0 + j)], a.target[3 * (j0 + j) + 1], a.target[3 * (j0 + j) + 2], 0.f);
                __syncthreads();
                for (int j = 0; j < nt; ++j) {
                    const float4 q = tile[j];
#pragma unroll
                    for (int k = 0; k < PL_Q; ++k) {
                        const float dx = px[k] - q.x, dy = py[k] - q.y, dz = pz[k] - q.z;
                        const float d = (dx * dx + dy * dy) + dz * dz;          // SqDist, direct differences
                        if (d < best[k]) { best[k] = d; bi[k] = j0 + j; }
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < PL_Q; ++k) {
                const int i = g0 + k * PL_T + tid;
                if (i < N) {
                    sum += sqrtf(best[k]);
                    a.nn[(size_t)h * N + i] = bi[k];
                }
            }
        }
    }
    double s = (double)sum;
    for (int o = 16; o > 0; o >>= 1) s = s + __shfl_xor_sync(0xffffffffu, s, o);
    if ((tid & 31) == 0) red[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
        double tot = red[0];
        for (int w = 1; w < PL_T / 32; ++w) tot = tot + red[w];
        a.dis[h] = (float)(tot / (double)N);
    }
}

// loss = mean_h(dis c - w log(c + 1e-8)), loss_rot = mean_h(c la - w log(c + 1e-8)), dis[0][argmax c[0]]  (loss.py:237-251)
__global__ void __launch_bounds__(256) poss_finalize_kernel(PossArgs a, float* __restrict__ vals) {
    __shared__ double r0[8], r1[8];
    __shared__ unsigned long long rk[8];
    const int H = a.bs * a.num_p, tid = threadIdx.x;
    double s0 = 0.0, s1 = 0.0;
    unsigned long long key = 0;                                   // (confidence bits, ~index): first maximum wins
    for (int h = tid; h < H; h += 256) {
        const float c = a.pred_c[h];
        const double lg = (double)a.w * (double)logf(c + 1e-8f);
        s0 += (double)a.dis[h] * (double)c - lg;
        s1 += (double)c * (double)a.la[h] - lg;
        if (h < a.num_p) {
            uint32_t bits = __float_as_uint(c);
            bits = (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
            const unsigned long long k = ((unsigned long long)bits << 32) | (uint32_t)(~(uint32_t)h);
            key = k > key ? k : key;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        s0 = s0 + __shfl_xor_sync(0xffffffffu, s0, o);
        s1 = s1 + __shfl_xor_sync(0xffffffffu, s1, o);
        const unsigned long long ok = __shfl_xor_sync(0xffffffffu, key, o);
        key = ok > key ? ok : key;
    }
    if ((tid & 31) == 0) { r0[tid >> 5] = s0; r1[tid >> 5] = s1; rk[tid >> 5] = key; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < 8; ++w) { s0 = s0 + r0[w]; s1 = s1 + r1[w]; key = rk[w] > key ? rk[w] : key; }
        const int which = (int)(~(uint32_t)(key & 0xffffffffu));
        a.which[0] = which;
        vals[0] = (float)(s0 / (double)H);
        vals[1] = a.dis[which];
        vals[2] = (float)(s1 / (double)H);
    }
}

__global__ void __launch_bounds__(PL_T) poss_bwd_kernel(PossArgs a, const float* __restrict__ gout, float* __restrict__ grad_r,
                                                        float* __restrict__ grad_t, float* __restrict__ grad_c) {
    __shared__ float Rs[12];
    __shared__ double red[(PL_T / 32) * 12];
    const int h = blockIdx.x, tid = threadIdx.x, N = a.N, b = h / a.num_p, H = a.bs * a.num_p;
    double q[4], t[3], tr[9], ax[3];
    if (tid == 0) {
        double R[9], la;
        load_pose(a, h, q, t, tr, ax);
        ego_rotation<double>(q, t, tr, ax, R, la);
        for (int k = 0; k < 9; ++k) Rs[k] = (float)R[k];
        for (int k = 0; k < 3; ++k) Rs[9 + k] = a.pred_t[(size_t)h * 3 + k];
    }
    __syncthreads();
    const float* mb = a.model + (size_t)b * N * 3;
    const float* tb = a.target + (a.nn_mode ? 0 : (size_t)b * N * 3);
    double acc[12];
    for (int k = 0; k < 12; ++k) acc[k] = 0.0;
    for (int i = tid; i < N; i += PL_T) {
        const float mx = mb[3 * i], my = mb[3 * i + 1], mz = mb[3 * i + 2];
        const int j = a.nn_mode ? a.nn[(size_t)h * N + i] : i;
        const float dx = fmaf(Rs[0], mx, fmaf(Rs[1], my, fmaf(Rs[2], mz, Rs[9]))) - tb[3 * j];
        const float dy = fmaf(Rs[3], mx, fmaf(Rs[4], my, fmaf(Rs[5], mz, Rs[10]))) - tb[3 * j + 1];
        const float dz = fmaf(Rs[6], mx, fmaf(Rs[7], my, fmaf(Rs[8], mz, Rs[11]))) - tb[3 * j + 2];
        const float nrm = sqrtf(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
        const double inv = nrm > 0.f ? 1.0 / (double)nrm : 0.0;   // the norm's subgradient at 0 is 0, as in torch
        const double gx = dx * inv, gy = dy * inv, gz = dz * inv;
        acc[0] += gx * mx; acc[1] += gx * my; acc[2] += gx * mz;
        acc[3] += gy * mx; acc[4] += gy * my; acc[5] += gy * mz;
        acc[6] += gz * mx; acc[7] += gz * my; acc[8] += gz * mz;
        acc[9] += gx; acc[10] += gy; acc[11] += gz;
    }
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        double x = acc[k];
        for (int o = 16; o > 0; o >>= 1) x = x + __shfl_xor_sync(0xffffffffu, x, o);
        acc[k] = x;
    }
    if ((tid & 31) == 0)
        for (int k = 0; k < 12; ++k) red[(tid >> 5) * 12 + k] = acc[k];
    __syncthreads();
    if (tid == 0) {
        for (int k = 0; k < 12; ++k) {
            double s = red[k];
            for (int w = 1; w < PL_T / 32; ++w) s = s + red[w * 12 + k];
            acc[k] = s;
        }
        const double g0 = (double)gout[0], g1 = (double)gout[1], g2 = (double)gout[2];
        const double c = (double)a.pred_c[h];
        const double cd = (g0 * c / (double)H + (h == a.which[0] ? g1 : 0.0)) / (double)N;    // weight of sum_i |.|
        const double cl = g2 * c / (double)H;                                                   // weight of la_h
        const double dlog = (double)a.w / ((double)(a.pred_c[h] + 1e-8f));
        grad_c[h] = (float)((g0 * ((double)a.dis[h] - dlog) + g2 * ((double)a.la[h] - dlog)) / (double)H);
        D7 qd[4] = {dvar(q[0], 0), dvar(q[1], 1), dvar(q[2], 2), dvar(q[3], 3)};
        D7 td[3] = {dvar(t[0], 4), dvar(t[1], 5), dvar(t[2], 6)};
        D7 R[9], la;
        ego_rotation<D7>(qd, td, tr, ax, R, la);
        double g[7];
        for (int k = 0; k < 7; ++k) {
            double s = cl * la.g[k];
            for (int e = 0; e < 9; ++e) s += cd * acc[e] * R[e].g[k];
            g[k] = s;
        }
        for (int k = 0; k < 4; ++k) grad_r[(size_t)h * 4 + k] = (float)g[k];
        for (int k = 0; k < 3; ++k) grad_t[(size_t)h * 3 + k] = (float)(g[4 + k] + cd * acc[9 + k]);
    }
}

size_t ws_layout(int H, int N, size_t* off_la, size_t* off_which, size_t* off_nn) {
    size_t o = 0;
    o += (size_t)H * sizeof(float); *off_la = o;
    o += (size_t)H * sizeof(float); *off_which = o;
    o += 16; *off_nn = o;
    o += (size_t)H * N * sizeof(int32_t);
    return o + 64;
}

int fill_args(const char* who, PossArgs& a, const float* pred_r, const float* pred_t, const float* pred_c, const float* model,
              const float* target, int bs, int num_p, int N, int nn_mode, float w, const float* axis3, const float* target_r9,
              void* ws, size_t ws_bytes) {
    PB_REQUIRE(pred_r && pred_t && pred_c && model && target && axis3 && target_r9, "%s: null pointer", who);
    PB_REQUIRE(bs >= 1 && num_p >= 1 && N >= 1, "%s: need bs, num_p, N >= 1", who);
    PB_REQUIRE((long long)bs * num_p <= 0x7fffffffLL / (N > 4 ? N : 4), "%s: bs * num_p * N exceeds the int32 range", who);
    size_t ol, ow, on;
    const size_t need = ws_layout(bs * num_p, N, &ol, &ow, &on);
    PB_REQUIRE(ws && ws_bytes >= need && ((uintptr_t)ws & 15) == 0, "%s: workspace too small or misaligned (%zu < %zu)", who,
               ws_bytes, need);
    a.pred_r = pred_r; a.pred_t = pred_t; a.pred_c = pred_c; a.model = model; a.target = target;
    a.bs = bs; a.num_p = num_p; a.N = N; a.nn_mode = nn_mode ? 1 : 0; a.w = w;
    a.axis = axis3; a.target_r = target_r9;
    char* base = (char*)ws;
    a.dis = (float*)base; a.la = (float*)(base + ol); a.which = (int32_t*)(base + ow); a.nn = (int32_t*)(base + on);
    return 0;
}

}  // namespace

extern "C" size_t pb_poss_loss_workspace_bytes(int bs, int num_p, int N) {
    if (bs < 1 || num_p < 1 || N < 1) return 0;
    size_t a, b, c;
    return ws_layout(bs * num_p, N, &a, &b, &c);
}

extern "C" int pb_poss_loss_fwd(const float* pred_r, const float* pred_t, const float* pred_c, const float* model,
                                const float* target, int bs, int num_p, int N, int nn_mode, float w, const float* axis3,
                                const float* target_r9, float* vals3, void* ws, size_t ws_bytes, void* stream) {
    PossArgs a;
    if (fill_args("pb_poss_loss_fwd", a, pred_r, pred_t, pred_c, model, target, bs, num_p, N, nn_mode, w, axis3, target_r9, ws,
                  ws_bytes)) return -1;
    PB_REQUIRE(vals3, "pb_poss_loss_fwd: null output");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = a.nn_mode ? (size_t)(N < PL_TILE ? N : PL_TILE) * sizeof(float4) : 0;
    poss_fwd_kernel<<<bs * num_p, PL_T, smem, st>>>(a);
    poss_finalize_kernel<<<1, 256, 0, st>>>(a, vals3);
    return pb_check_launch("pb_poss_loss_fwd");
}

extern "C" int pb_poss_loss_bwd(const float* pred_r, const float* pred_t, const float* pred_c, const float* model,
                                const float* target, int bs, int num_p, int N, int nn_mode, float w, const float* axis3,
                                const float* target_r9, const float* grad_vals3, const void* ws, size_t ws_bytes, float* grad_r,
                                float* grad_t, float* grad_c, void* stream) {
    PossArgs a;
    if (fill_args("pb_poss_loss_bwd", a, pred_r, pred_t, pred_c, model, target, bs, num_p, N, nn_mode, w, axis3, target_r9,
                  (void*)ws, ws_bytes)) return -1;
    PB_REQUIRE(grad_vals3 && grad_r && grad_t && grad_c, "pb_poss_loss_bwd: null pointer");
    poss_bwd_kernel<<<bs * num_p, PL_T, 0, (cudaStream_t)stream>>>(a, grad_vals3, grad_r, grad_t, grad_c);
    return pb_check_launch("pb_poss_loss_bwd");
}
