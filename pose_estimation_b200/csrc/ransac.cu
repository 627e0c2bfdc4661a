// ransac.cu -- K1..K4: PnP-RANSAC on compacted correspondences (sm_100a).
//
// Replaces cv2.solvePnPRansac at tools/trainer.py:423-427 / lib/utils/utlis.py:59-63.
//   ransac_hypothesis_kernel / ransac_score_kernel   K1-K3, see ransac_score.cu.
//   K4 argmax over slices (lowest hypothesis index on ties), inlier bitmap by warp ballot,
//      Gauss-Newton refit on the inliers, optional re-score -> refit rounds.  The refit's sums have
//      ONE canonical order (256 virtual threads striding the points, xor-butterfly over each group
//      of 32, groups added in order) that the CPU oracle reproduces bit for bit;
//      ransac_finalize_kernel<G> realises it with G = 1, 2, 4 or 8 warps per frame (one CTA per
//      frame), G chosen from the frame count so that enough warps are in flight.
#include "common.cuh"
#include "ransac_plan.h"

namespace {

// ------------------------------------------------------------------------------------------
struct FinalArgs {
    const float* pts;
    const int32_t* count;
    const float* K4;
    const float* hypRt;
    const unsigned long long* slice_best;
    const int32_t* cnt_all;          // balanced scoring: [F,H] counts, the arg-max is taken here
    float* R; float* t; int32_t* n_inl; int32_t* best_hyp; int32_t* n_inl_refined; int32_t* status;
    uint32_t* inl_bitmap; float* R0; float* t0;
    int cap, H, HS, gn_iters, lo_rounds, min_inliers;
    float inv_thr;
};

__device__ __forceinline__ bool chol6_solve(const double A21[21], const double g[6], double d[6]) {
    // every loop is fully unrolled (constant indices) so A/L stay in registers, no local memory
    double A[6][6], L[6][6];
    int k = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = i; j < 6; ++j) { A[i][j] = A21[k]; A[j][i] = A21[k]; ++k; }
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 6; ++j) L[i][j] = 0.0;
    bool spd = true;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        double s = A[j][j];
#pragma unroll
        for (int m = 0; m < j; ++m) s = s - L[j][m] * L[j][m];
        spd = spd && (s > 0.0);
        double ljj = sqrt(s);
        L[j][j] = 1.0 / ljj;    // the diagonal holds the RECIPROCAL (one division per column)
#pragma unroll
        for (int i = j + 1; i < 6; ++i) {
            double v = A[i][j];
#pragma unroll
            for (int m = 0; m < j; ++m) v = v - L[i][m] * L[j][m];
            L[i][j] = v * L[j][j];
        }
    }
    if (!spd) return false;
    double z[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        double v = g[i];
#pragma unroll
        for (int m = 0; m < i; ++m) v = v - L[i][m] * z[m];
        z[i] = v * L[i][i];
    }
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        double v = z[i];
#pragma unroll
        for (int m = i + 1; m < 6; ++m) v = v - L[m][i] * d[m];
        d[i] = v * L[i][i];
    }
    return true;
}

// one inlier's contribution to the normal equations (21 upper-triangle terms of J^T J, 6 of J^T r)
__device__ __forceinline__ void gn_accumulate(float* acc, const float R[9], const float cb[3], const float tc[3],
                                              float fx, float fy, float cx, float cy, float Xi, float Yi, float Zi,
                                              float Ui, float Vi) {
    float xo = Xi - cb[0], yo = Yi - cb[1], zo = Zi - cb[2];
    float xr = fmaf(R[0], xo, fmaf(R[1], yo, R[2] * zo));
    float yr = fmaf(R[3], xo, fmaf(R[4], yo, R[5] * zo));
    float zr = fmaf(R[6], xo, fmaf(R[7], yo, R[8] * zo));
    float xc = xr + tc[0], yc = yr + tc[1], zc = zr + tc[2];
    float iz = 1.0f / zc;
    float un = xc * iz, vn = yc * iz;
    float ru = fmaf(fx, un, cx) - Ui;
    float rv = fmaf(fy, vn, cy) - Vi;
    float ja = fx * iz, jb = fy * iz;
    float jc = -(ja * un), jd = -(jb * vn);
    float ju[6] = {yr * jc, fmaf(zr, ja, -(xr * jc)), -(yr * ja), ja, 0.0f, jc};
    float jv[6] = {fmaf(yr, jd, -(zr * jb)), -(xr * jd), xr * jb, 0.0f, jb, jd};
    int k = 0;
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int s2 = r; s2 < 6; ++s2) { acc[k] = fmaf(ju[r], ju[s2], fmaf(jv[r], jv[s2], acc[k])); ++k; }
#pragma unroll
    for (int r = 0; r < 6; ++r) acc[21 + r] = fmaf(ju[r], ru, fmaf(jv[r], rv, acc[21 + r]));
}

// solve the 6x6 normal equations (FP64 Cholesky) -> FP32 step; false if not SPD / not finite
__device__ __forceinline__ bool gn_solve(const float* tot, float dl[6]) {
    double A21[21], g[6], d[6];
#pragma unroll
    for (int k = 0; k < 21; ++k) A21[k] = (double)tot[k];
#pragma unroll
    for (int k = 0; k < 6; ++k) g[k] = -(double)tot[21 + k];
    bool ok = chol6_solve(A21, g, d);
    if (ok) {
#pragma unroll
        for (int k = 0; k < 6; ++k) { dl[k] = (float)d[k]; ok = ok && isfinite(dl[k]); }
    }
    return ok;
}

// q <- normalise(q + 0.5 * (0, d_rot) * q), tc += d_trans
__device__ __forceinline__ void gn_apply(float q[4], float tc[3], const float dl[6]) {
    const float hx = 0.5f * dl[0], hy = 0.5f * dl[1], hz = 0.5f * dl[2];
    const float w0 = q[0], x0 = q[1], y0 = q[2], z0 = q[3];
    float nw = w0 - fmaf(hx, x0, fmaf(hy, y0, hz * z0));
    float nx = x0 + fmaf(hx, w0, fmaf(hy, z0, -(hz * y0)));
    float ny = y0 + fmaf(hy, w0, fmaf(hz, x0, -(hx * z0)));
    float nz = z0 + fmaf(hz, w0, fmaf(hx, y0, -(hy * x0)));
    float in = 1.0f / sqrtf(fmaf(nw, nw, fmaf(nx, nx, fmaf(ny, ny, nz * nz))));
    q[0] = nw * in; q[1] = nx * in; q[2] = ny * in; q[3] = nz * in;
    tc[0] += dl[3]; tc[1] += dl[4]; tc[2] += dl[5];
}

// 32-value transpose-reduce: lane k ends with the xor-butterfly sum (offsets 16,8,4,2,1) of value
// k over the warp.  16+8+4+2+1 shuffles instead of 27 separate butterflies; float add is
// commutative, so every partial sum has the butterfly's bits.
__device__ __forceinline__ float warp_transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
        const bool up = (lane & half) != 0;
#pragma unroll
        for (int k = 0; k < half; ++k) {
            const float send = up ? v[k] : v[k + half];
            const float keep = up ? v[k + half] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, half);
        }
    }
    return v[0];
}

constexpr int NVG = PB_THREADS / 32;  // virtual groups of the canonical summation order

// K4, one CTA of G warps per frame (G = 1, 2, 4, 8).  The refit's sums have ONE canonical order --
// 256 virtual threads stride the points, each group of 32 is xor-butterfly reduced, the 8 group
// sums are added in order -- whatever G is: warp w plays groups w, w+G, ...; the per-group sums
// meet in shared memory and every thread adds them in group order, then solves the 6x6 system
// redundantly (no broadcast).  So results do not depend on G, and the CPU oracle reproduces them.
template <int G>
__global__ void __launch_bounds__(32 * G) ransac_finalize_kernel(FinalArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* bm = reinterpret_cast<uint32_t*>(smem_raw);  // [ceil(cap/256)*8] inlier bitmap
    __shared__ float red[2][NVG][32];
    __shared__ int ired[NVG];
    constexpr int NT = 32 * G;

    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int P = min(a.count[f], a.cap);
    const int words = (a.cap + 31) / 32;
    const float* X = a.pts + (size_t)f * 5 * a.cap;
    const float* Y = X + a.cap;
    const float* Z = Y + a.cap;
    const float* U = Z + a.cap;
    const float* V = U + a.cap;
    const float fx = a.K4[4 * f], fy = a.K4[4 * f + 1], cx = a.K4[4 * f + 2], cy = a.K4[4 * f + 3];

    // argmax over slices (or, after balanced scoring, over the hypotheses' counts: every warp redundantly,
    // max count, lowest hypothesis index on ties -- the same packed key the scoring kernel builds)
    unsigned long long key = 0ull;
    if (a.cnt_all) {
        for (int h = lane; h < a.H; h += 32) {
            const int cv = a.cnt_all[(size_t)f * a.H + h];
            const unsigned long long k2 = ((unsigned long long)(uint32_t)(cv + 1) << 32) | (0xFFFFFFFFu - (uint32_t)h);
            key = k2 > key ? k2 : key;
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            const unsigned long long o = __shfl_xor_sync(0xffffffffu, key, off);
            key = o > key ? o : key;
        }
    } else {
        key = a.slice_best[(size_t)f * a.HS];
        for (int s = 1; s < a.HS; ++s) {
            unsigned long long k = a.slice_best[(size_t)f * a.HS + s];
            key = k > key ? k : key;
        }
    }
    const int bc = (int)(uint32_t)(key >> 32) - 1;
    const int bh = (int)(0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFull));
    const bool fail_pts = P < 4;
    const bool fail_model = !fail_pts && (bc < 0 || bc < a.min_inliers);
    if (fail_pts || fail_model) {
        for (int k = tid; k < 9; k += NT) {
            float v = (k % 4 == 0) ? 1.0f : 0.0f;
            a.R[9 * f + k] = v;
            if (a.R0) a.R0[9 * f + k] = v;
        }
        if (tid < 3) { a.t[3 * f + tid] = 0.0f; if (a.t0) a.t0[3 * f + tid] = 0.0f; }
        if (tid == 0) {
            a.n_inl[f] = 0; a.best_hyp[f] = -1; a.n_inl_refined[f] = 0;
            a.status[f] = fail_pts ? PB_STATUS_TOO_FEW_POINTS : PB_STATUS_NO_MODEL;
        }
        if (a.inl_bitmap)
            for (int w = tid; w < words; w += NT) a.inl_bitmap[(size_t)f * words + w] = 0u;
        return;
    }

    float Rb[9], tb[3];
    {
        const float* o = a.hypRt + ((size_t)f * a.H + bh) * 12;
#pragma unroll
        for (int k = 0; k < 9; ++k) Rb[k] = o[k];
        tb[0] = o[9]; tb[1] = o[10]; tb[2] = o[11];
    }
    if (tid < 9 && a.R0) a.R0[9 * f + tid] = Rb[tid];
    if (tid < 3 && a.t0) a.t0[3 * f + tid] = tb[tid];

    float q[4], tt[3] = {tb[0], tb[1], tb[2]};
    rot_to_quat(Rb, q);
    float M[12];
    make_proj_s(fx, fy, cx, cy, Rb, tb, a.inv_thr, M);

    int nref = bc;
    // round -1: bitmap of the best minimal hypothesis; rounds 0..lo_rounds: refit then re-score
    for (int round = -1; round <= a.lo_rounds; ++round) {
        if (round >= 0) {
            // ---- centroid of the inlier object points
            int cn = 0;
            for (int vw = wid; vw < NVG; vw += G) {
                float cs[3] = {0.0f, 0.0f, 0.0f};
                for (int i = vw * 32 + lane; i < P; i += PB_THREADS)
                    if ((bm[i >> 5] >> (i & 31)) & 1u) { cs[0] += X[i]; cs[1] += Y[i]; cs[2] += Z[i]; ++cn; }
#pragma unroll
                for (int k = 0; k < 3; ++k) cs[k] = warp_sum_bfly(cs[k]);
                if (lane < 3) red[0][vw][lane] = lane == 0 ? cs[0] : (lane == 1 ? cs[1] : cs[2]);
            }
            cn = __reduce_add_sync(0xffffffffu, cn);
            if (lane == 0) ired[wid] = cn;
            __syncthreads();
            float cb[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                float sacc = red[0][0][k];
#pragma unroll
                for (int vw = 1; vw < NVG; ++vw) sacc = sacc + red[0][vw][k];
                cb[k] = sacc;
            }
            int cnt_all = 0;
#pragma unroll
            for (int w = 0; w < G; ++w) cnt_all += ired[w];
            __syncthreads();         // red / ired may be rewritten
            if (cnt_all < 4) break;  // uniform
            const float fn = (float)cnt_all;
            cb[0] = cb[0] / fn; cb[1] = cb[1] / fn; cb[2] = cb[2] / fn;
            float R[9];
            quat_to_rot(q, R);
            float tc[3];
#pragma unroll
            for (int i = 0; i < 3; ++i)
                tc[i] = fmaf(R[3 * i], cb[0], fmaf(R[3 * i + 1], cb[1], fmaf(R[3 * i + 2], cb[2], tt[i])));
            // ---- Gauss-Newton on the reprojection error of the inliers
            for (int it = 0; it < a.gn_iters; ++it) {
                quat_to_rot(q, R);
                const int buf = it & 1;
#pragma unroll 1
                for (int vw = wid; vw < NVG; vw += G) {
                    float acc[32];
#pragma unroll
                    for (int k = 0; k < 32; ++k) acc[k] = 0.0f;
                    for (int i = vw * 32 + lane; i < P; i += PB_THREADS) {
                        if (!((bm[i >> 5] >> (i & 31)) & 1u)) continue;
                        gn_accumulate(acc, R, cb, tc, fx, fy, cx, cy, X[i], Y[i], Z[i], U[i], V[i]);
                    }
                    red[buf][vw][lane] = warp_transpose_reduce32(acc, lane);
                }
                __syncthreads();
                float totl = red[buf][0][lane];      // lane k: normal-equation term k
#pragma unroll
                for (int vw = 1; vw < NVG; ++vw) totl = totl + red[buf][vw][lane];
                float tot[27];
#pragma unroll
                for (int k = 0; k < 27; ++k) tot[k] = __shfl_sync(0xffffffffu, totl, k);
                float dl[6];
                if (!gn_solve(tot, dl)) break;       // every thread holds the same tot: uniform
                gn_apply(q, tc, dl);
                if (fmaxf(fmaxf(fabsf(dl[0]), fabsf(dl[1])), fabsf(dl[2])) < PB_GN_EPS &&
                    fmaxf(fmaxf(fabsf(dl[3]), fabsf(dl[4])), fabsf(dl[5])) < PB_GN_EPS) break;   // converged (uniform)
            }
            quat_to_rot(q, R);
#pragma unroll
            for (int i = 0; i < 3; ++i)
                tt[i] = tc[i] - fmaf(R[3 * i], cb[0], fmaf(R[3 * i + 1], cb[1], R[3 * i + 2] * cb[2]));
            make_proj_s(fx, fy, cx, cy, R, tt, a.inv_thr, M);
            __syncthreads();         // every warp is done reading bm and red
        }
        // ---- (re-)score: inlier bitmap by warp ballot
        int cn = 0;
        for (int base = 0; base < P; base += NT) {
            const int i = base + tid;
            bool in = i < P && inlier_test(M, X[i], Y[i], Z[i], U[i] * a.inv_thr, V[i] * a.inv_thr);
            unsigned b = __ballot_sync(0xffffffffu, in);
            if (lane == 0) { bm[(base >> 5) + wid] = b; cn += __popc(b); }
        }
        if (lane == 0) ired[wid] = cn;
        __syncthreads();
        int tot_in = 0;
#pragma unroll
        for (int w = 0; w < G; ++w) tot_in += ired[w];
        if (round < 0) {
            if (a.inl_bitmap) {
                const int pw = (P + 31) / 32;
                for (int w = tid; w < words; w += NT) a.inl_bitmap[(size_t)f * words + w] = w < pw ? bm[w] : 0u;
            }
        } else {
            nref = tot_in;
        }
        __syncthreads();
    }
    // a refit that lost its consensus (fewer than half of the minimal hypothesis' inliers, or fewer than
    // min_inliers, under the refitted pose) is discarded: the minimal hypothesis is returned instead (uniform)
    float R[9];
    if (nref < max((bc + 1) >> 1, a.min_inliers)) {
#pragma unroll
        for (int k = 0; k < 9; ++k) R[k] = Rb[k];
        tt[0] = tb[0]; tt[1] = tb[1]; tt[2] = tb[2];
        nref = bc;
    } else {
        quat_to_rot(q, R);
    }
    for (int k = tid; k < 9; k += NT) a.R[9 * f + k] = R[k];
    if (tid < 3) a.t[3 * f + tid] = tt[tid];
    if (tid == 0) { a.n_inl[f] = bc; a.best_hyp[f] = bh; a.n_inl_refined[f] = nref; a.status[f] = PB_STATUS_OK; }
}

template <int G>
int launch_finalize(const FinalArgs& fa, int F, size_t smem, cudaStream_t st) {
    cudaFuncSetAttribute(ransac_finalize_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    ransac_finalize_kernel<G><<<F, 32 * G, smem, st>>>(fa);
    return pb_check_launch("ransac_finalize_kernel");
}

// ------------------------------------------------------------------------------------------
__global__ void resample_inliers_kernel(const uint32_t* bitmap, const int32_t* count, int F, int cap, int n,
                                        uint32_t frame0, uint32_t seed_lo, uint32_t seed_hi, int32_t* out) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    const int words = (cap + 31) / 32;
    const uint32_t* bm = bitmap + (size_t)f * words;
    const int P = min(count[f], cap);
    int32_t* o = out + (size_t)f * n;
    int m = 0;
    for (int w = 0; w < (P + 31) / 32; ++w) {
        uint32_t bits = bm[w];
        if (32 * w + 32 > P) bits &= (1u << (P - 32 * w)) - 1u;  // bits at and beyond count[f] are ignored
        m += __popc(bits);
    }
    if (m == 0) { for (int k = 0; k < n; ++k) o[k] = -1; return; }
    if (m > n) {
        int sel = 0, j = 0;
        uint32_t r[4] = {0, 0, 0, 0};
        for (int i = 0; i < P && sel < n; ++i) {
            if (!((bm[i >> 5] >> (i & 31)) & 1u)) continue;
            if ((j & 3) == 0) philox4x32_10((uint32_t)(j >> 2), frame0 + (uint32_t)f, 0u, PB_RESAMPLE_TAG, seed_lo, seed_hi, r);
            uint32_t pick = __umulhi(r[j & 3], (uint32_t)(m - j));
            if (pick < (uint32_t)(n - sel)) o[sel++] = i;
            ++j;
        }
    } else {
        int k = 0;
        for (int i = 0; i < P; ++i)
            if ((bm[i >> 5] >> (i & 31)) & 1u) o[k++] = i;
        for (; k < n; ++k) o[k] = o[k - m];
    }
}


}  // namespace

extern "C" size_t pb_ransac_workspace_bytes(int F, int cap, const pb_ransac_params* prm) {
    if (F <= 0 || cap <= 0 || !prm || prm->n_hyp <= 0) return 0;
    return make_plan(F, cap, prm, pb_sm_count()).total;
}

extern "C" long long pb_ransac_tc_stats_offset(int F, int cap, const pb_ransac_params* prm) {
    if (F <= 0 || cap <= 0 || !prm || prm->n_hyp <= 0) return -1;
    const Plan p = make_plan(F, cap, prm, pb_sm_count());
    return p.tc ? (long long)p.off_tcstats : -1;
}

extern "C" int pb_pnp_ransac_pts(const float* pts, const int32_t* count, const float* K4, int F, int cap,
                                 uint32_t frame0, const pb_ransac_params* prm, const int32_t* sample_idx,
                                 float* R, float* t, int32_t* n_inl, int32_t* best_hyp, int32_t* n_inl_refined,
                                 int32_t* status, int32_t* hyp_counts, uint32_t* inl_bitmap, float* R0, float* t0,
                                 void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(prm != nullptr, "pb_pnp_ransac_pts: prm is null");
    PB_REQUIRE(F >= 0 && cap > 0, "pb_pnp_ransac_pts: bad sizes F=%d cap=%d", F, cap);
    PB_REQUIRE(prm->n_hyp > 0 && prm->n_hyp <= (1 << 20), "pb_pnp_ransac_pts: n_hyp=%d out of range", prm->n_hyp);
    PB_REQUIRE(prm->thresh_px > 0.0f && prm->gn_iters >= 0 && prm->lo_rounds >= 0, "pb_pnp_ransac_pts: bad params");
    if (F == 0) return 0;
    PB_REQUIRE(pts && count && K4 && R && t && n_inl && best_hyp && n_inl_refined && status,
               "pb_pnp_ransac_pts: null pointer argument");
    Plan p = make_plan(F, cap, prm, pb_sm_count());
    PB_REQUIRE(ws != nullptr && ws_bytes >= p.total, "pb_pnp_ransac_pts: workspace too small (%zu < %zu)", ws_bytes, p.total);
    PB_REQUIRE(p.smem_score <= 227 * 1024 && p.smem_final <= 200 * 1024, "pb_pnp_ransac_pts: shared memory plan too large");
    PB_REQUIRE((size_t)F * p.HS < (1ull << 31), "pb_pnp_ransac_pts: grid too large");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ScoreArgs sa;
    sa.pts = pts; sa.count = count; sa.K4 = K4; sa.sample_idx = sample_idx;
    sa.hypRt = static_cast<float*>(ws);
    sa.hyp_counts = hyp_counts;
    sa.slice_best = reinterpret_cast<unsigned long long*>(static_cast<char*>(ws) + p.off_best);
    if (p.bal && reinterpret_cast<uintptr_t>(pts) % 16 != 0) p.bal = 0;     // the balanced kernel stages with 16-byte async copies
    sa.cnt_all = p.bal ? (hyp_counts ? hyp_counts : reinterpret_cast<int32_t*>(static_cast<char*>(ws) + p.off_cnt)) : nullptr;
    sa.prefix = p.bal ? reinterpret_cast<int32_t*>(static_cast<char*>(ws) + p.off_prefix) : nullptr;
    // K3 with the projection on the tensor cores: the warp-level path (ransac_score_mma.cu: impl 5 / 6, and the automatic
    // choice when PB_SCORE_MMA_DEFAULT) or tcgen05 + TMEM (ransac_score_tc.cu: impl 2 / 3 / 4)
    const int impl = pb_score_impl();
    const bool can_tc = p.tc && p.bal;
    const bool use_tc = can_tc && impl >= 2 && impl <= 4;
    const bool use_mma = can_tc && (impl == 5 || impl == 6 || (impl == 0 && PB_SCORE_MMA_DEFAULT));
    sa.tprefix = use_tc ? reinterpret_cast<int32_t*>(static_cast<char*>(ws) + p.off_tprefix) : nullptr;
    sa.tc_stats = (use_tc || use_mma) ? reinterpret_cast<int32_t*>(static_cast<char*>(ws) + p.off_tcstats) : nullptr;
    sa.fstats = use_mma ? reinterpret_cast<float*>(static_cast<char*>(ws) + p.off_fstats) : nullptr;
    sa.tc_kappa = pb_score_tc_kappa();
    sa.tc_mode = (impl == 3 || impl == 6) ? 1 : impl == 4 ? 2 : 0;
    if (use_tc) sa.prefix = nullptr;
    sa.cap = cap; sa.H = prm->n_hyp; sa.HS = p.HS; sa.Hs = p.Hs; sa.CH = p.CH; sa.PP = p.PP;
    sa.frame0 = frame0; sa.seed_lo = prm->seed_lo; sa.seed_hi = prm->seed_hi;
    sa.inv_thr = 1.0f / prm->thresh_px; sa.F = F;
    if (pb_launch_score(sa, p, F, st)) return -2;
    FinalArgs fa;
    fa.pts = pts; fa.count = count; fa.K4 = K4; fa.hypRt = sa.hypRt; fa.slice_best = sa.slice_best; fa.cnt_all = sa.cnt_all;
    fa.R = R; fa.t = t; fa.n_inl = n_inl; fa.best_hyp = best_hyp; fa.n_inl_refined = n_inl_refined;
    fa.status = status; fa.inl_bitmap = inl_bitmap; fa.R0 = R0; fa.t0 = t0;
    fa.cap = cap; fa.H = prm->n_hyp; fa.HS = p.HS; fa.gn_iters = prm->gn_iters; fa.lo_rounds = prm->lo_rounds;
    fa.min_inliers = prm->min_inliers; fa.inv_thr = sa.inv_thr;
    pb_stage_mark(2, st);
    // warps per frame: the largest G whose CTAs are ALL resident at once (120 registers per thread:
    // 2 / 4 / 8 / 17 CTAs per SM for G = 8 / 4 / 2 / 1), so a batch is one wave with no tail; results
    // do not depend on G.
    int rcf;
    const int sms = pb_sm_count();
    if (F <= sms * 2) rcf = launch_finalize<8>(fa, F, p.smem_final, st);
    else if (F <= sms * 4) rcf = launch_finalize<4>(fa, F, p.smem_final, st);
    else if (F <= sms * 8) rcf = launch_finalize<2>(fa, F, p.smem_final, st);
    else rcf = launch_finalize<1>(fa, F, p.smem_final, st);
    if (rcf) return -2;
    pb_stage_mark(3, st);
    return 0;
}

extern "C" int pb_resample_inliers(const uint32_t* inl_bitmap, const int32_t* count, int F, int cap, int n,
                                   uint32_t frame0, uint32_t seed_lo, uint32_t seed_hi, int32_t* out, void* stream) {
    PB_REQUIRE(F >= 0 && cap > 0 && n > 0, "pb_resample_inliers: bad sizes");
    if (F == 0) return 0;
    PB_REQUIRE(inl_bitmap && count && out, "pb_resample_inliers: null pointer argument");
    resample_inliers_kernel<<<(F + 63) / 64, 64, 0, static_cast<cudaStream_t>(stream)>>>(
        inl_bitmap, count, F, cap, n, frame0, seed_lo, seed_hi, out);
    return pb_check_launch("resample_inliers_kernel") ? -2 : 0;
}
