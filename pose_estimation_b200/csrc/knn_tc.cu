// knn_tc.cu -- K5 on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// pb_knn1_tc: idx[b,i] = argmin_j |x_i - y_j|^2 (the KeOps handle of train.py:126) as a tiled
// tcgen05 contraction with an exact FP32 re-rank, so indices equal the FP32 SqDist argmin (lowest
// j on ties) of the SIMT path / the CPU oracle.
//
//   scores   s~_ij = |q_j|^2 - 2 p_i.q_j  with p = x - c, q = y - c (c = centre of y's bounding box);
//            it differs from |x_i - y_j|^2 by the row constant |p_i|^2, which does not change the argmin.
//   operands every FP32 coordinate is split EXACTLY into three bf16 (8+8+8 mantissa bits); a K = 32
//            row carries the 27 cross products of the three coordinates plus |q|^2 (three bf16),
//            so two K=16 tcgen05.mma (kind::f16, M=128, N=128, FP32 accumulate in TMEM) give s~ to
//            within E ~ 1e-5 R^2 (R = cloud radius), dominated by the accumulation, not by bf16.
//   epilogue one thread per query row reads its 128 scores per tile with tcgen05.ld (32x32b.x32),
//            reduces each chunk of 32 with 3-input minima (FMNMX3, 0.56 ALU op per pair) and only
//            when a chunk dips below  running-min + T  (T = 2^-13 R^2 >> 2E) looks at single
//            columns; those candidates go to a short per-row list in shared memory.
//   re-rank  after the sweep the candidates within T of the final minimum (a provable superset of
//            the FP32 argmin) are re-evaluated with the exact FP32 formula; rows whose list
//            overflowed fall back to an exact scan.
// The MMA of tile n+1 and the bulk copy (1-D TMA) of tile n+2 overlap the epilogue of tile n.
#include <cuda_bf16.h>
#include "common.cuh"
#include "tc.cuh"

namespace {

constexpr int TM = 128;              // rows / columns per operand tile
constexpr int TILE_U4 = 4 * TM;      // uint4 per operand tile: [kc 0..3][row][8 bf16] = 8 KB
constexpr int TILE_BYTES = TILE_U4 * 16;
constexpr float TC_TOL_REL = 1.0f / 8192.0f;   // T = 2^-13 * R^2

struct Split3 { __nv_bfloat16 h, m, l; };
__device__ __forceinline__ Split3 split3(float v) {
    Split3 s;
    s.h = __float2bfloat16_rn(v);
    const float r1 = v - __bfloat162float(s.h);        // exact
    s.m = __float2bfloat16_rn(r1);
    const float r2 = r1 - __bfloat162float(s.m);       // exact
    s.l = __float2bfloat16_rn(r2);
    return s;
}
__device__ __forceinline__ uint32_t pk(__nv_bfloat16 a, __nv_bfloat16 b) {
    return (uint32_t)__bfloat16_as_ushort(a) | ((uint32_t)__bfloat16_as_ushort(b) << 16);
}

struct PrepArgs {
    const float* x; const float* y;
    int N, M, y_shared, ntN, ntM;
    uint4* Aop; uint4* Bop; float* tol; float* centre;   // per chunk-local batch element
    int b0;
};

// K = 32 operand row of one point.  Query rows (IS_A): -2*(h,h,h,m,m,m,l,l,l) per coordinate, then 1,1,1,0,0.
// Database rows: (h,m,l,h,m,l,h,m,l) per coordinate, then |q|^2 as three bf16, 0, 0.
template <bool IS_A>
__device__ __forceinline__ void write_row(uint4* tile_base, int r, float px, float py, float pz, bool pad) {
    __nv_bfloat16 e[32];
    const __nv_bfloat16 zero = __float2bfloat16_rn(0.0f), one = __float2bfloat16_rn(1.0f);
#pragma unroll
    for (int k = 0; k < 32; ++k) e[k] = zero;
    if (!pad) {
        const float pc[3] = {px, py, pz};
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const Split3 s = split3(IS_A ? -2.0f * pc[c] : pc[c]);
            if (IS_A) {
                e[9 * c + 0] = s.h; e[9 * c + 1] = s.h; e[9 * c + 2] = s.h;
                e[9 * c + 3] = s.m; e[9 * c + 4] = s.m; e[9 * c + 5] = s.m;
                e[9 * c + 6] = s.l; e[9 * c + 7] = s.l; e[9 * c + 8] = s.l;
            } else {
#pragma unroll
                for (int g = 0; g < 3; ++g) { e[9 * c + 3 * g] = s.h; e[9 * c + 3 * g + 1] = s.m; e[9 * c + 3 * g + 2] = s.l; }
            }
        }
        if (IS_A) { e[27] = one; e[28] = one; e[29] = one; }
        else {
            const float n = fmaf(pz, pz, fmaf(py, py, px * px));
            const Split3 s = split3(n);
            e[27] = s.h; e[28] = s.m; e[29] = s.l;
        }
    } else if (!IS_A) {
        e[27] = __float2bfloat16_rn(1.0e30f);          // padded database row: never a candidate
    }
#pragma unroll
    for (int kc = 0; kc < 4; ++kc)
        tile_base[kc * TM + r] = make_uint4(pk(e[8 * kc], e[8 * kc + 1]), pk(e[8 * kc + 2], e[8 * kc + 3]),
                                            pk(e[8 * kc + 4], e[8 * kc + 5]), pk(e[8 * kc + 6], e[8 * kc + 7]));
}

// one CTA per batch element: bounding-box centre of y, operand tiles of x and y, tolerance
__global__ void __launch_bounds__(256) knn_tc_prepare_kernel(PrepArgs a) {
    __shared__ float red[6][8];
    __shared__ float cen[3];
    __shared__ float rmax[8];
    const int bl = blockIdx.x, b = a.b0 + bl, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* xb = a.x + (size_t)b * a.N * 3;
    const float* yb = a.y + (a.y_shared ? 0 : (size_t)b * a.M * 3);
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int j = tid; j < a.M; j += 256) {
#pragma unroll
        for (int c = 0; c < 3; ++c) { const float v = yb[3 * (size_t)j + c]; lo[c] = fminf(lo[c], v); hi[c] = fmaxf(hi[c], v); }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], off));
            hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], off));
        }
        if (lane == 0) { red[c][warp] = lo[c]; red[3 + c][warp] = hi[c]; }
    }
    __syncthreads();
    if (tid < 3) {
        float l = red[tid][0], h = red[3 + tid][0];
        for (int w = 1; w < 8; ++w) { l = fminf(l, red[tid][w]); h = fmaxf(h, red[3 + tid][w]); }
        const float c = 0.5f * (l + h);
        cen[tid] = isfinite(c) ? c : 0.0f;
    }
    __syncthreads();
    const float c0 = cen[0], c1 = cen[1], c2 = cen[2];
    float r2 = 0.0f;
    uint4* At = a.Aop + (size_t)bl * a.ntN * TILE_U4;
    uint4* Bt = a.Bop + (size_t)bl * a.ntM * TILE_U4;
    for (int i = tid; i < a.ntN * TM; i += 256) {
        const bool pad = i >= a.N;
        float px = 0.0f, py = 0.0f, pz = 0.0f;
        if (!pad) { px = xb[3 * (size_t)i] - c0; py = xb[3 * (size_t)i + 1] - c1; pz = xb[3 * (size_t)i + 2] - c2; }
        r2 = fmaxf(r2, fmaf(pz, pz, fmaf(py, py, px * px)));
        write_row<true>(At + (size_t)(i / TM) * TILE_U4, i % TM, px, py, pz, pad);
    }
    for (int j = tid; j < a.ntM * TM; j += 256) {
        const bool pad = j >= a.M;
        float px = 0.0f, py = 0.0f, pz = 0.0f;
        if (!pad) { px = yb[3 * (size_t)j] - c0; py = yb[3 * (size_t)j + 1] - c1; pz = yb[3 * (size_t)j + 2] - c2; }
        r2 = fmaxf(r2, fmaf(pz, pz, fmaf(py, py, px * px)));
        write_row<false>(Bt + (size_t)(j / TM) * TILE_U4, j % TM, px, py, pz, pad);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) r2 = fmaxf(r2, __shfl_xor_sync(0xffffffffu, r2, off));
    if (lane == 0) rmax[warp] = r2;
    __syncthreads();
    if (tid == 0) {
        float r = rmax[0];
        for (int w = 1; w < 8; ++w) r = fmaxf(r, rmax[w]);
        a.tol[bl] = isfinite(r) ? r * TC_TOL_REL : INFINITY;   // non-finite input: everything is a candidate
        a.centre[3 * bl] = c0; a.centre[3 * bl + 1] = c1; a.centre[3 * bl + 2] = c2;
    }
}

struct MainArgs {
    const float* x; const float* y;
    int N, M, y_shared, ntN, ntM;
    const uint4* Aop; const uint4* Bop; const float* tol;
    int64_t* idx; float* d2;
    int b0;
    int* err;
};

constexpr int CAND = 24;   // candidate slots per query row

__device__ __forceinline__ float fmin3(float a, float b, float c) {
    float d;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// grid (row tiles, batch elements of the chunk); 128 threads: thread <-> query row of the tile
__global__ void __launch_bounds__(TM) knn_tc_kernel(MainArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint4* sA = reinterpret_cast<uint4*>(smem_raw);                    // [TILE_U4]
    uint4* sB = sA + TILE_U4;                                          // [2][TILE_U4]
    float* cval = reinterpret_cast<float*>(sB + 2 * TILE_U4);          // [CAND][TM] candidate scores
    int* ccol = reinterpret_cast<int*>(cval + CAND * TM);              // [CAND][TM] candidate columns
    __shared__ __align__(8) uint64_t bar_load[2], bar_mma[2], bar_a;
    __shared__ uint32_t tmem_base;

    const int tid = threadIdx.x, warp = tid >> 5;
    const int rt = blockIdx.x, bl = blockIdx.y, b = a.b0 + bl;
    const uint4* Ag = a.Aop + ((size_t)bl * a.ntN + rt) * TILE_U4;
    const uint4* Bg = a.Bop + (size_t)bl * a.ntM * TILE_U4;
    const int ntM = a.ntM;

    if (tid == 0) {
        tc::mbar_init(&bar_load[0], 1); tc::mbar_init(&bar_load[1], 1);
        tc::mbar_init(&bar_mma[0], 1); tc::mbar_init(&bar_mma[1], 1);
        tc::mbar_init(&bar_a, 1);
        tc::fence_mbar_init();
    }
    if (warp == 0) tc::tmem_alloc<256>(&tmem_base);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = tmem_base;
    const uint32_t idesc = tc::idesc_bf16_f32(TM, TM);
    bool ok = true;

    auto issue_mma = [&](int tile) {       // thread 0: D[acc tile&1] = A * B[tile&1]^T  (two K=16 steps)
        const uint32_t acc = tmem + (uint32_t)(tile & 1) * TM;
        const uint32_t sb = tc::smem_u32(sB + (tile & 1) * TILE_U4);
#pragma unroll
        for (int k = 0; k < 2; ++k)
            tc::mma_bf16(acc, tc::smem_desc(tc::smem_u32(sA) + k * 2 * TM * 16, TM * 16, 128),
                         tc::smem_desc(sb + k * 2 * TM * 16, TM * 16, 128), idesc, k > 0);
        tc::mma_commit(&bar_mma[tile & 1]);
    };
    auto issue_load = [&](int tile) {      // thread 0: bulk copy of database tile `tile`
        tc::mbar_expect_tx(&bar_load[tile & 1], TILE_BYTES);
        tc::bulk_g2s(sB + (tile & 1) * TILE_U4, Bg + (size_t)tile * TILE_U4, TILE_BYTES, &bar_load[tile & 1]);
    };

    if (tid == 0) {
        tc::mbar_expect_tx(&bar_a, TILE_BYTES);
        tc::bulk_g2s(sA, Ag, TILE_BYTES, &bar_a);
        issue_load(0);
        if (ntM > 1) issue_load(1);
        ok = tc::mbar_wait(&bar_a, 0) && tc::mbar_wait(&bar_load[0], 0);
        tc::fence_after_sync();
        issue_mma(0);
    }

    const int row = rt * TM + tid;
    const float T = a.tol[bl];
    float m_run = INFINITY, thr = INFINITY;    // running minimum of s~ and m_run + T
    int ncand = 0;
    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);

    for (int n = 0; n < ntM; ++n) {
        ok = tc::mbar_wait(&bar_mma[n & 1], (uint32_t)(n >> 1) & 1u) && ok;
        tc::fence_after_sync();
        if (tid == 0) {
            if (n + 1 < ntM) {
                ok = tc::mbar_wait(&bar_load[(n + 1) & 1], (uint32_t)((n + 1) >> 1) & 1u) && ok;
                tc::fence_after_sync();
                issue_mma(n + 1);
            }
            if (n + 2 < ntM) issue_load(n + 2);    // sB[n&1] is free: MMA n has completed
        }
        // ---- epilogue of tile n: this thread's 128 scores, 32 at a time
        const uint32_t acc = lane_base + (uint32_t)(n & 1) * TM;
#pragma unroll 1
        for (int c0 = 0; c0 < TM; c0 += 32) {
            uint32_t r[32];
            tc::tmem_ld32(acc + c0, r);
            tc::tmem_ld_wait();
            float g[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float* v = reinterpret_cast<const float*>(r) + 8 * q;
                g[q] = fmin3(fmin3(v[0], v[1], v[2]), fmin3(v[3], v[4], v[5]), fminf(v[6], v[7]));
            }
            const float cmin = fminf(fminf(g[0], g[1]), fminf(g[2], g[3]));
            if (cmin < thr) {
                // a chunk reaches below running-min + T: tighten the running minimum first, then keep
                // the single columns below the (new) threshold -- always a superset of "within T of
                // the final minimum" because the running minimum never drops below the final one
                if (cmin < m_run) { m_run = cmin; thr = cmin + T; }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (g[q] < thr) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const float v = __uint_as_float(r[8 * q + k]);
                            if (v < thr) {
                                if (ncand == CAND) {             // full: drop entries the threshold has overtaken
                                    int w = 0;
                                    for (int e = 0; e < CAND; ++e) {
                                        const float cv = cval[e * TM + tid];
                                        if (cv < thr) { cval[w * TM + tid] = cv; ccol[w * TM + tid] = ccol[e * TM + tid]; ++w; }
                                    }
                                    ncand = w;
                                }
                                if (ncand < CAND) { cval[ncand * TM + tid] = v; ccol[ncand * TM + tid] = n * TM + c0 + 8 * q + k; ++ncand; }
                                else ncand = CAND + 1;           // still full: overflow -> exact scan for this row
                            }
                        }
                    }
                }
            }
        }
        tc::fence_before_sync();
        __syncthreads();                          // accumulator n&1 may be overwritten by MMA n+2
    }

    // ---- exact FP32 re-rank of the candidates within T of the minimum (lowest index on ties)
    if (row < a.N) {
        const float* xb = a.x + (size_t)b * a.N * 3;
        const float* yb = a.y + (a.y_shared ? 0 : (size_t)b * a.M * 3);
        const float qx = xb[3 * (size_t)row], qy = xb[3 * (size_t)row + 1], qz = xb[3 * (size_t)row + 2];
        float best = INFINITY;
        int bidx = 0;
        int nexact = 0;
        if (ncand <= CAND && ok) {
            for (int k = 0; k < ncand; ++k) {
                if (!(cval[k * TM + tid] < thr)) continue;      // thr = final minimum + T
                ++nexact;
                const int j = ccol[k * TM + tid];
                if (j >= a.M) continue;
                const float d = sqdist3(qx, qy, qz, yb[3 * (size_t)j], yb[3 * (size_t)j + 1], yb[3 * (size_t)j + 2]);
                if (d < best || (d == best && j < bidx)) { best = d; bidx = j; }
            }
        }
        if (ncand > CAND || !ok || best == INFINITY) {          // list overflow / NaN input: exact scan
            best = INFINITY; bidx = 0;
            atomicAdd(a.err + 1, 1);                            // statistics: rows that fell back
            for (int j = 0; j < a.M; ++j) {
                const float d = sqdist3(qx, qy, qz, yb[3 * (size_t)j], yb[3 * (size_t)j + 1], yb[3 * (size_t)j + 2]);
                if (d < best) { best = d; bidx = j; }
            }
        }
        nexact = __reduce_add_sync(__activemask(), nexact);
        if ((tid & 31) == (__ffs(__activemask()) - 1)) atomicAdd(a.err + 2, nexact);   // statistics: exact evaluations
        a.idx[(size_t)b * a.N + row] = bidx;
        if (a.d2) a.d2[(size_t)b * a.N + row] = best;
    }
    if (!ok && tid == 0) *a.err = 1;
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<256>(tmem);
}

constexpr int TC_CHUNK = 256;      // batch elements per prepare/main pair: operands (~90 MB for C3) stay in L2
constexpr size_t MAIN_SMEM = (size_t)3 * TILE_BYTES + (size_t)CAND * TM * 8;

size_t tc_ws_bytes(int B, int N, int M) {
    const int nb = B < TC_CHUNK ? B : TC_CHUNK;
    const size_t ntN = (N + TM - 1) / TM, ntM = (M + TM - 1) / TM;
    return (size_t)nb * (ntN + ntM) * TILE_BYTES + (size_t)nb * 4 * sizeof(float) + 256;   // tail: int stats[4]
}

// ---- self test: D[128x128] = A[128x32] * B[128x32]^T, operands given in the canonical tile layout
// [kc 0..3][row 0..127][8 bf16]; exercises descriptors, two accumulating K=16 MMAs, commit, tcgen05.ld.
__global__ void __launch_bounds__(128) tc_selftest_kernel(const uint4* __restrict__ A, const uint4* __restrict__ B,
                                                          float* __restrict__ D, int* __restrict__ err) {
    __shared__ __align__(128) uint4 sA[4 * 128];
    __shared__ __align__(128) uint4 sB[4 * 128];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 4 * 128; i += 128) { sA[i] = A[i]; sB[i] = B[i]; }
    if (tid == 0) { tc::mbar_init(&bar, 1); tc::fence_mbar_init(); }
    if (warp == 0) tc::tmem_alloc<128>(&tmem_base);
    tc::fence_proxy_async();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = tmem_base;
    if (tid == 0) {
        const uint32_t idesc = tc::idesc_bf16_f32(128, 128);
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const uint64_t ad = tc::smem_desc(tc::smem_u32(sA) + k * 2 * 2048, 2048, 128);
            const uint64_t bd = tc::smem_desc(tc::smem_u32(sB) + k * 2 * 2048, 2048, 128);
            tc::mma_bf16(tmem, ad, bd, idesc, k > 0);
        }
        tc::mma_commit(&bar);
    }
    if (!tc::mbar_wait(&bar, 0)) { if (tid == 0) *err = 1; }
    tc::fence_after_sync();
#pragma unroll 1
    for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t r[32];
        tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, r);
        tc::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) D[tid * 128 + c0 + j] = __uint_as_float(r[j]);
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<128>(tmem);
}

}  // namespace

extern "C" int pb_tc_selftest(const void* A, const void* B, float* D, int* err, void* stream) {
    PB_REQUIRE(A && B && D && err, "pb_tc_selftest: null pointer argument");
    tc_selftest_kernel<<<1, 128, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const uint4*>(A),
                                                                          static_cast<const uint4*>(B), D, err);
    return pb_check_launch("tc_selftest_kernel") ? -2 : 0;
}

extern "C" size_t pb_knn1_tc_workspace_bytes(int B, int N, int M) {
    if (B <= 0 || N <= 0 || M <= 0) return 0;
    return tc_ws_bytes(B, N, M);
}

extern "C" int pb_knn1_tc(const float* x, const float* y, int B, int N, int M, int y_shared, int64_t* idx, float* d2,
                          void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(B >= 0 && N >= 0 && M > 0, "pb_knn1_tc: bad sizes B=%d N=%d M=%d", B, N, M);
    if (B == 0 || N == 0) return 0;
    PB_REQUIRE(x && y && idx, "pb_knn1_tc: null pointer argument");
    PB_REQUIRE(M <= (1 << 24), "pb_knn1_tc: M too large");
    const size_t need = tc_ws_bytes(B, N, M);
    PB_REQUIRE(ws && ws_bytes >= need, "pb_knn1_tc: workspace too small (%zu < %zu)", ws_bytes, need);
    PB_REQUIRE(reinterpret_cast<uintptr_t>(ws) % 128 == 0, "pb_knn1_tc: workspace must be 128-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nb_max = B < TC_CHUNK ? B : TC_CHUNK;
    const int ntN = (N + TM - 1) / TM, ntM = (M + TM - 1) / TM;
    PB_REQUIRE(ntN <= 65535 * 32, "pb_knn1_tc: N too large");
    char* base = static_cast<char*>(ws);
    uint4* Aop = reinterpret_cast<uint4*>(base);
    uint4* Bop = Aop + (size_t)nb_max * ntN * TILE_U4;
    float* tol = reinterpret_cast<float*>(Bop + (size_t)nb_max * ntM * TILE_U4);
    float* centre = tol + nb_max;
    int* err = reinterpret_cast<int*>(centre + 3 * nb_max);          // last word of the float block (+pad)
    cudaFuncSetAttribute(knn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MAIN_SMEM);
    cudaMemsetAsync(err, 0, 4 * sizeof(int), st);   // [0] barrier timeout, [1] fallback rows, [2] exact evaluations
    for (int b0 = 0; b0 < B; b0 += TC_CHUNK) {
        const int nb = (B - b0) < TC_CHUNK ? (B - b0) : TC_CHUNK;
        PrepArgs pa{x, y, N, M, y_shared, ntN, ntM, Aop, Bop, tol, centre, b0};
        knn_tc_prepare_kernel<<<nb, 256, 0, st>>>(pa);
        if (pb_check_launch("knn_tc_prepare_kernel")) return -2;
        MainArgs ma{x, y, N, M, y_shared, ntN, ntM, Aop, Bop, tol, idx, d2, b0, err};
        knn_tc_kernel<<<dim3(ntN, nb), TM, MAIN_SMEM, st>>>(ma);
        if (pb_check_launch("knn_tc_kernel")) return -2;
    }
    return 0;
}
