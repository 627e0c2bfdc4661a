// metrics.cu -- K6 per-frame pose metrics + accuracy table, symmetric-KNN pose loss, K7 T-vote.
//
//   pose_metrics_kernel   Trainer.cal_dis (tools/trainer.py:370-381): model*R^T+t fused with
//                         ADD / ADD-S (lib/utils/metric.py:17-36), angular_distance (:67-80) and
//                         translation_distance (:82-84); nothing is materialised.
//   accumulate_kernel     success tests + per-object sums (tools/trainer.py:156-157,174-221).
//   pose_loss_*           PoseLoss.forward (lib/network/loss.py:25-41) and its closed-form gradient.
//   vote_kernel           pred_t = (p_emb + t_res^T).mean(1)  (lib/network/krrn.py:153).
#include "common.cuh"
#include "nn.cuh"

namespace {

constexpr int PM_T = 512;            // threads per frame
constexpr int PM_PARTS = PM_T / KT;  // database parts scanned side by side by groups of KT threads

// One CTA of 512 threads per frame.  ADD frames: threads stride the model points.  ADD-S frames
// (N x N pairs): the CTA is KT query lanes (QPT queries each) x PM_PARTS database parts, so the
// serial scan per thread is N / PM_PARTS long; partial minima meet in shared memory (min is
// order-independent), sums use a fixed tree.
__global__ void __launch_bounds__(PM_T) pose_metrics_kernel(const float* __restrict__ R, const float* __restrict__ t,
                                                            const float* __restrict__ R_gt, const float* __restrict__ t_gt,
                                                            const float* __restrict__ model, const float* __restrict__ target,
                                                            const int32_t* __restrict__ cls, const uint8_t* __restrict__ is_sym,
                                                            int N, int C, float* __restrict__ add, float* __restrict__ r_deg,
                                                            float* __restrict__ t_m) {
    __shared__ float4 tile[TY];
    __shared__ float smin[PM_PARTS][KT * QPT];
    __shared__ float red[PM_T / 32];
    const int f = blockIdx.x, tid = threadIdx.x;
    int c = cls[f];
    c = c < 0 ? 0 : (c >= C ? C - 1 : c);
    const float* m = model + (size_t)c * N * 3;
    LoadTransformed Lp, Lg;
    Lp.m = m; Lg.m = m;
#pragma unroll
    for (int k = 0; k < 9; ++k) { Lp.R[k] = R[9 * (size_t)f + k]; Lg.R[k] = R_gt[9 * (size_t)f + k]; }
#pragma unroll
    for (int k = 0; k < 3; ++k) { Lp.t[k] = t[3 * (size_t)f + k]; Lg.t[k] = t_gt[3 * (size_t)f + k]; }
    const float* tg = target ? target + (size_t)f * N * 3 : nullptr;
    const bool sym = is_sym && is_sym[c] != 0;
    float s = 0.0f;
    if (!sym) {
        for (int i = tid; i < N; i += PM_T) {
            const float4 p = Lp(i);
            const float4 g = tg ? make_float4(tg[3 * (size_t)i], tg[3 * (size_t)i + 1], tg[3 * (size_t)i + 2], 0.0f) : Lg(i);
            s += sqrtf(sqdist3(p.x, p.y, p.z, g.x, g.y, g.z));
        }
    } else {
        // ADD-S: for every target point the closest predicted point (metric.py:27-31)
        const int ql = tid % KT, part = tid / KT;   // part is warp-uniform
        for (int base = 0; base < N; base += KT * QPT) {
            float qx[QPT], qy[QPT], qz[QPT], best[QPT];
#pragma unroll
            for (int q = 0; q < QPT; ++q) {
                const int i = base + q * KT + ql;
                float4 g = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                if (i < N) g = tg ? make_float4(tg[3 * (size_t)i], tg[3 * (size_t)i + 1], tg[3 * (size_t)i + 2], 0.0f) : Lg(i);
                qx[q] = g.x; qy[q] = g.y; qz[q] = g.z;
                best[q] = INFINITY;
            }
            for (int j0 = 0; j0 < N; j0 += TY) {
                const int n = min(TY, N - j0);
                __syncthreads();
                for (int j = tid; j < n; j += PM_T) tile[j] = Lp(j0 + j);
                __syncthreads();
                const int per = (n + PM_PARTS - 1) / PM_PARTS;
                const int j1 = min(n, (part + 1) * per);
#pragma unroll 4
                for (int j = part * per; j < j1; ++j) {
                    const float4 p = tile[j];
#pragma unroll
                    for (int q = 0; q < QPT; ++q) best[q] = fminf(best[q], sqdist3(qx[q], qy[q], qz[q], p.x, p.y, p.z));
                }
            }
#pragma unroll
            for (int q = 0; q < QPT; ++q) smin[part][q * KT + ql] = best[q];
            __syncthreads();
            if (part == 0) {
#pragma unroll
                for (int q = 0; q < QPT; ++q) {
                    float b = smin[0][q * KT + ql];
#pragma unroll
                    for (int pp = 1; pp < PM_PARTS; ++pp) b = fminf(b, smin[pp][q * KT + ql]);
                    if (base + q * KT + ql < N) s += sqrtf(b);
                }
            }
        }
    }
    // deterministic block sum: butterfly per warp, then warps in order
    s = warp_sum_bfly(s);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
        float tot = red[0];
        for (int w = 1; w < PM_T / 32; ++w) tot = tot + red[w];
        add[f] = tot / (float)N;
        float q1[4], q2[4];
        rot_to_quat(Lp.R, q1);
        rot_to_quat(Lg.R, q2);
        float d = fabsf(fmaf(q1[0], q2[0], fmaf(q1[1], q2[1], fmaf(q1[2], q2[2], q1[3] * q2[3]))));
        const float eps = 1e-7f;
        d = fminf(fmaxf(d, -1.0f + eps), 1.0f - eps);
        r_deg[f] = 2.0f * acosf(d) / 3.14159265358979323846f * 180.0f;
        const float dx = Lp.t[0] - Lg.t[0], dy = Lp.t[1] - Lg.t[1], dz = Lp.t[2] - Lg.t[2];
        t_m[f] = sqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
    }
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v = v + __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

// one CTA per class: deterministic sums over the frames of that class
__global__ void __launch_bounds__(256) accumulate_kernel(const float* __restrict__ add, const float* __restrict__ r_deg,
                                                         const float* __restrict__ t_m, const int32_t* __restrict__ cls,
                                                         const int32_t* __restrict__ status, const float* __restrict__ diameter,
                                                         int F, int C, int variant, float add_frac, float r_thr, float t_thr,
                                                         int count_frames, double* __restrict__ table) {
    __shared__ double red[8][8];
    const int c = blockIdx.x;
    const float thr_add = add_frac * diameter[c];
    double v[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // n, succ_rt, dis_rt, succ_r, dis_r, succ_t, dis_t, failed
    for (int f = threadIdx.x; f < F; f += 256) {
        // an out-of-range class index is clamped exactly as pose_metrics_kernel clamps it (the frame was measured
        // against that class' model) and the frame is counted in the `failed` row: every frame lands in one column
        const int cf = cls[f];
        const int cc = cf < 0 ? 0 : (cf >= C ? C - 1 : cf);
        if (cc != c) continue;
        const float a = add[f], r = r_deg[f], tt = t_m[f];
        v[0] += 1.0;
        v[1] += (a < thr_add) ? 1.0 : 0.0;
        v[2] += (double)a;
        v[3] += (r < r_thr) ? 1.0 : 0.0;
        v[4] += (double)r;
        v[5] += (tt < t_thr) ? 1.0 : 0.0;
        v[6] += (double)tt;
        v[7] += ((status && status[f] != 0) || cf != cc) ? 1.0 : 0.0;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = warp_sum_d(v[k]);
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 8; ++k) red[threadIdx.x >> 5][k] = v[k];
    __syncthreads();
    if (threadIdx.x < 8) {
        double s = red[0][threadIdx.x];
        for (int w = 1; w < 8; ++w) s = s + red[w][threadIdx.x];
        const int k = threadIdx.x;
        if (k == 0) {
            if (count_frames) { table[PB_ROW_ALL_NUM * C + c] += s; table[PB_ROW_OBJ_NUM * C + c] += s; }
        } else if (k == 7) {
            if (count_frames) table[PB_ROW_FAILED * C + c] += s;
        } else {
            table[(PB_ROW_VARIANT0 + 6 * variant + (k - 1)) * C + c] += s;
        }
    }
}

// nearest target index per predicted point for symmetric samples, identity otherwise
__global__ void __launch_bounds__(KT) pose_loss_nn_kernel(const float* __restrict__ pred, const float* __restrict__ target,
                                                          const uint8_t* __restrict__ sym, int N, int64_t* __restrict__ nn_idx) {
    __shared__ float4 tile[TY];
    const int b = blockIdx.y;
    const int i0 = blockIdx.x * (KT * QPT) + threadIdx.x;
    if (!sym[b]) {
#pragma unroll
        for (int q = 0; q < QPT; ++q) {
            const int i = i0 + q * KT;
            if (i < N) nn_idx[(size_t)b * N + i] = i;
        }
        return;
    }
    const float* pb = pred + (size_t)b * N * 3;
    float qx[QPT], qy[QPT], qz[QPT], best[QPT];
    int bidx[QPT];
#pragma unroll
    for (int q = 0; q < QPT; ++q) {
        const int i = i0 + q * KT;
        const bool ok = i < N;
        qx[q] = ok ? pb[3 * (size_t)i] : 0.0f;
        qy[q] = ok ? pb[3 * (size_t)i + 1] : 0.0f;
        qz[q] = ok ? pb[3 * (size_t)i + 2] : 0.0f;
        best[q] = INFINITY; bidx[q] = 0;
    }
    nn_scan<true>(LoadXYZ{target + (size_t)b * N * 3}, N, tile, qx, qy, qz, best, bidx);
#pragma unroll
    for (int q = 0; q < QPT; ++q) {
        const int i = i0 + q * KT;
        if (i < N) nn_idx[(size_t)b * N + i] = bidx[q];
    }
}

__global__ void __launch_bounds__(KT) gather_mean_l2_kernel(const float* __restrict__ pred, const float* __restrict__ target,
                                                            const int64_t* __restrict__ nn_idx, int N,
                                                            float* __restrict__ loss_b) {
    __shared__ float red[KT / 32];
    const int b = blockIdx.x;
    const float* pb = pred + (size_t)b * N * 3;
    const float* tb = target + (size_t)b * N * 3;
    float s = 0.0f;
    for (int i = threadIdx.x; i < N; i += KT) {
        const size_t j = (size_t)nn_idx[(size_t)b * N + i];
        s += sqrtf(sqdist3(pb[3 * (size_t)i], pb[3 * (size_t)i + 1], pb[3 * (size_t)i + 2], tb[3 * j], tb[3 * j + 1],
                           tb[3 * j + 2]));
    }
    const float v = block_sum_kt(s, red) / (float)N;
    if (threadIdx.x == 0) loss_b[b] = v;
}

// d loss / d R[b], d loss / d t[b] for loss = mean_b mean_i |pred_bi - tgt_bi|, pred = R m + t
__global__ void __launch_bounds__(KT) pose_loss_bwd_kernel(const float* __restrict__ pred, const float* __restrict__ model,
                                                           const float* __restrict__ target, const int64_t* __restrict__ nn_idx,
                                                           int N, float scale, float* __restrict__ grad_R,
                                                           float* __restrict__ grad_t) {
    __shared__ float red[KT / 32];
    const int b = blockIdx.x;
    const float* pb = pred + (size_t)b * N * 3;
    const float* mb = model + (size_t)b * N * 3;
    const float* tb = target + (size_t)b * N * 3;
    float acc[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) acc[k] = 0.0f;
    for (int i = threadIdx.x; i < N; i += KT) {
        const size_t j = (size_t)nn_idx[(size_t)b * N + i];
        const float dx = pb[3 * (size_t)i] - tb[3 * j], dy = pb[3 * (size_t)i + 1] - tb[3 * j + 1],
                    dz = pb[3 * (size_t)i + 2] - tb[3 * j + 2];
        const float n = sqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
        if (n > 0.0f) {
            const float inv = scale / n;
            const float g[3] = {dx * inv, dy * inv, dz * inv};
            const float mx = mb[3 * (size_t)i], my = mb[3 * (size_t)i + 1], mz = mb[3 * (size_t)i + 2];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                acc[3 * a] = fmaf(g[a], mx, acc[3 * a]);
                acc[3 * a + 1] = fmaf(g[a], my, acc[3 * a + 1]);
                acc[3 * a + 2] = fmaf(g[a], mz, acc[3 * a + 2]);
                acc[9 + a] += g[a];
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        const float v = block_sum_kt(acc[k], red);
        if (threadIdx.x == 0) {
            if (k < 9) grad_R[9 * (size_t)b + k] = v;
            else grad_t[3 * (size_t)b + (k - 9)] = v;
        }
    }
}

__global__ void __launch_bounds__(256) vote_kernel(const float* __restrict__ cloud, const float* __restrict__ t_res, int NP,
                                                   float* __restrict__ t) {
    __shared__ float red[8][3];
    const int b = blockIdx.x;
    const float* cb = cloud + (size_t)b * NP * 3;
    const float* rb = t_res + (size_t)b * 3 * NP;
    float s[3] = {0.0f, 0.0f, 0.0f};
    for (int i = threadIdx.x; i < NP; i += 256) {
        s[0] += cb[3 * (size_t)i] + rb[i];
        s[1] += cb[3 * (size_t)i + 1] + rb[(size_t)NP + i];
        s[2] += cb[3 * (size_t)i + 2] + rb[2 * (size_t)NP + i];
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) s[k] = warp_sum_bfly(s[k]);
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 3; ++k) red[threadIdx.x >> 5][k] = s[k];
    __syncthreads();
    if (threadIdx.x < 3) {
        float v = red[0][threadIdx.x];
        for (int w = 1; w < 8; ++w) v = v + red[w][threadIdx.x];
        t[3 * (size_t)b + threadIdx.x] = v / (float)NP;
    }
}

}  // namespace

extern "C" int pb_pose_metrics(const float* R, const float* t, const float* R_gt, const float* t_gt, const float* model,
                               const float* target, const int32_t* cls, const uint8_t* is_sym, int F, int N, int C,
                               float* add, float* r_deg, float* t_m, void* stream) {
    PB_REQUIRE(F >= 0 && N > 0 && C > 0, "pb_pose_metrics: bad sizes F=%d N=%d C=%d", F, N, C);
    if (F == 0) return 0;
    PB_REQUIRE(R && t && R_gt && t_gt && model && cls && add && r_deg && t_m, "pb_pose_metrics: null pointer argument");
    pose_metrics_kernel<<<F, PM_T, 0, static_cast<cudaStream_t>(stream)>>>(R, t, R_gt, t_gt, model, target, cls, is_sym, N, C,
                                                                       add, r_deg, t_m);
    return pb_check_launch("pose_metrics_kernel") ? -2 : 0;
}

extern "C" int pb_accumulate_results(const float* add, const float* r_deg, const float* t_m, const int32_t* cls,
                                     const int32_t* status, const float* diameter, int F, int C, int variant,
                                     float add_frac, float r_thr_deg, float t_thr_m, int count_frames, double* table,
                                     void* stream) {
    PB_REQUIRE(F >= 0 && C > 0 && variant >= 0 && variant < 3, "pb_accumulate_results: bad arguments");
    if (F == 0) return 0;
    PB_REQUIRE(add && r_deg && t_m && cls && diameter && table, "pb_accumulate_results: null pointer argument");
    accumulate_kernel<<<C, 256, 0, static_cast<cudaStream_t>(stream)>>>(add, r_deg, t_m, cls, status, diameter, F, C, variant,
                                                                      add_frac, r_thr_deg, t_thr_m, count_frames, table);
    return pb_check_launch("accumulate_kernel") ? -2 : 0;
}

extern "C" int pb_pose_loss_fwd(const float* R, const float* t, const float* model, const float* target,
                                const uint8_t* sym, int B, int N, float* pred, int64_t* nn_idx, float* loss_b,
                                void* stream) {
    PB_REQUIRE(B >= 0 && N > 0, "pb_pose_loss_fwd: bad sizes");
    PB_REQUIRE(B <= 65535, "pb_pose_loss_fwd: B too large (<= 65535)");
    if (B == 0) return 0;
    PB_REQUIRE(R && t && model && target && sym && pred && nn_idx && loss_b, "pb_pose_loss_fwd: null pointer argument");
    int rc = pb_transform_points(model, 0, R, t, B, N, pred, stream);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    dim3 grid((N + KT * QPT - 1) / (KT * QPT), B);
    pose_loss_nn_kernel<<<grid, KT, 0, st>>>(pred, target, sym, N, nn_idx);
    if (pb_check_launch("pose_loss_nn_kernel")) return -2;
    gather_mean_l2_kernel<<<B, KT, 0, st>>>(pred, target, nn_idx, N, loss_b);
    return pb_check_launch("gather_mean_l2_kernel") ? -2 : 0;
}

__global__ void identity_index_kernel(int64_t* __restrict__ nn_idx, int N) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) nn_idx[(size_t)blockIdx.y * N + i] = i;
}

// pb_pose_loss_fwd with the nearest-neighbour search of the symmetric samples on the tensor cores; sym_list is the
// device list of the n_sym batch elements with sym != 0 (sym NULL = every element is symmetric).  Same bits out.
extern "C" int pb_pose_loss_fwd_tc(const float* R, const float* t, const float* model, const float* target,
                                   const uint8_t* sym, const int* sym_list, int n_sym, int B, int N, float* pred,
                                   int64_t* nn_idx, float* loss_b, void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(B >= 0 && N > 0, "pb_pose_loss_fwd_tc: bad sizes");
    PB_REQUIRE(B <= 65535, "pb_pose_loss_fwd_tc: B too large (<= 65535)");
    if (B == 0) return 0;
    PB_REQUIRE(R && t && model && target && pred && nn_idx && loss_b, "pb_pose_loss_fwd_tc: null pointer argument");
    PB_REQUIRE(!sym || (n_sym >= 0 && n_sym <= B && (sym_list || n_sym == 0)), "pb_pose_loss_fwd_tc: sym needs its list of set elements");
    int rc = pb_transform_points(model, 0, R, t, B, N, pred, stream);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (sym) {
        identity_index_kernel<<<dim3((N + 255) / 256, B), 256, 0, st>>>(nn_idx, N);
        if (pb_check_launch("identity_index_kernel")) return -2;
    }
    if (!sym || n_sym > 0) {
        rc = pb_internal_knn1_tc("pb_pose_loss_fwd_tc", pred, target, B, N, N, 0, nn_idx, nullptr, ws, ws_bytes,
                                 sym ? sym_list : nullptr, n_sym, st);
        if (rc) return rc;
    }
    gather_mean_l2_kernel<<<B, KT, 0, st>>>(pred, target, nn_idx, N, loss_b);
    return pb_check_launch("gather_mean_l2_kernel") ? -2 : 0;
}

extern "C" int pb_pose_loss_bwd(const float* pred, const float* model, const float* target, const int64_t* nn_idx, int B,
                                int N, float gscale, float* grad_R, float* grad_t, void* stream) {
    PB_REQUIRE(B >= 0 && N > 0, "pb_pose_loss_bwd: bad sizes");
    if (B == 0) return 0;
    PB_REQUIRE(pred && model && target && nn_idx && grad_R && grad_t, "pb_pose_loss_bwd: null pointer argument");
    const float scale = gscale / ((float)B * (float)N);
    pose_loss_bwd_kernel<<<B, KT, 0, static_cast<cudaStream_t>(stream)>>>(pred, model, target, nn_idx, N, scale, grad_R, grad_t);
    return pb_check_launch("pose_loss_bwd_kernel") ? -2 : 0;
}

extern "C" int pb_vote_translation(const float* cloud, const float* t_res, int B, int NP, float* t, void* stream) {
    PB_REQUIRE(B >= 0 && NP > 0, "pb_vote_translation: bad sizes");
    if (B == 0) return 0;
    PB_REQUIRE(cloud && t_res && t, "pb_vote_translation: null pointer argument");
    vote_kernel<<<B, 256, 0, static_cast<cudaStream_t>(stream)>>>(cloud, t_res, NP, t);
    return pb_check_launch("vote_kernel") ? -2 : 0;
}
