// knn_tc.cu -- K5 on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// pb_knn1_tc: idx[b,i] = argmin_j |x_i - y_j|^2 (the KeOps handle of train.py:126) as a tiled
// tcgen05 contraction with an exact FP32 re-rank, so indices equal the FP32 SqDist argmin (lowest
// j on ties) of the SIMT path / the CPU oracle.
//
//   scores   s~_ij = |q_j|^2 - 2 p_i.q_j  with p = x - c, q = y - c (c = centre of y's bounding box);
//            it differs from |x_i - y_j|^2 by the row constant |p_i|^2, which does not change the argmin.
//   operands every centred coordinate is split into TWO bf16, v ~ h + m (h = bf16(v), m = bf16(v - h):
//            |v - (h + m)| <= 2^-16 |v|).  One K = 16 row per point: query rows carry (h,h,m,m) of -2p per
//            coordinate and 1,1,1,0; database rows (h,m,h,m) of q per coordinate and |q|^2 as three bf16
//            (24 bits) and 0 -- the twelve products are the four cross terms of every coordinate.  ONE
//            tcgen05.mma (kind::f16, M = 128, N = 128, K = 16, FP32 accumulate in TMEM) per 128 x 128 scores.
//   error    per coordinate |(-2p)(q) - (-2p)^(q)^| <= 2|p_c| 2^-16 |q_c| + |q_c| 2^-15 |p_c| = 2^-14 |p_c||q_c|,
//            so over the three coordinates <= 2^-14 |p||q| <= 2^-14 R^2 (R = largest centred norm of either
//            cloud, measured by the prepare kernel); |q|^2 adds 2^-24 R^2 and the FP32 accumulation of the 15
//            terms (each partial sum <= 3 R^2) at most 2^-18.4 R^2:  E <= 1.06 * 2^-14 R^2 in the worst case
//            (every rounding at half an ulp, all with the same sign, |p| = |q| = R, p parallel to q).
//   band     T = 2^-12 R^2 = 3.8 E > 2 E: the true FP32 argmin (and every tie) has a score within 2E < T of
//            the smallest score of its row, so it lies in a column group whose minimum is below  row minimum + T.
//   epilogue two epilogue warpgroups, each bound to one TMEM accumulator (thread = query row); per 32-column
//            chunk tcgen05.ld (32x32b.x32) and 3-input minima (FMNMX3) give the minima of the four 8-column
//            groups; a chunk whose minimum is below  running minimum + T  keeps a slot of a 7-deep ring.
//   re-rank  after the sweep every remembered group within T of the final row minimum is re-evaluated with
//            the SIMT kernel's exact FP32 formula (lowest index on ties); a row whose ring dropped a chunk
//            that could matter, a barrier timeout or non-finite input falls back to an exact scan.
// The MMA of tile n+1 and the bulk copy (1-D TMA) of tile n+2 overlap the epilogue of tile n.
#include <atomic>
#include <cuda_bf16.h>
#include "common.cuh"
#include "tc.cuh"

namespace {

constexpr int TM = 128;              // rows / columns per operand tile
constexpr int TILE_U4 = 2 * TM;      // uint4 per operand tile: [kc 0..1][row][8 bf16] = 4 KB (K = 16)
constexpr int TILE_BYTES = TILE_U4 * 16;
constexpr float TC_TOL_REL = 1.0f / 4096.0f;   // T = 2^-12 * R^2 (see the error bound above)

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
    const int* blist;                                     // optional: batch element of list position b0 + bl
    int frag_layout;                                      // database tiles in the mma.sync sweep's fragment order (see knn_mma_kernel)
};

// K = 16 operand row of one point.  Query rows (IS_A): per coordinate (h, h, m, m) of -2p, then 1, 1, 1, 0.
// Database rows: per coordinate (h, m, h, m) of q, then |q|^2 as three bf16, 0.  h + m carries 16 mantissa bits
// of a coordinate, so the twelve cross products give -2 p.q to 2^-16 R^2.
template <bool IS_A>
__device__ __forceinline__ void write_row(uint4* tile_base, int r, float px, float py, float pz, bool pad, bool frag_layout = false) {
    __nv_bfloat16 e[16];
    const __nv_bfloat16 zero = __float2bfloat16_rn(0.0f), one = __float2bfloat16_rn(1.0f);
#pragma unroll
    for (int k = 0; k < 16; ++k) e[k] = zero;
    if (!pad) {
        const float pc[3] = {px, py, pz};
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const Split3 s = split3(IS_A ? -2.0f * pc[c] : pc[c]);
            if (IS_A) { e[4 * c] = s.h; e[4 * c + 1] = s.h; e[4 * c + 2] = s.m; e[4 * c + 3] = s.m; }
            else      { e[4 * c] = s.h; e[4 * c + 1] = s.m; e[4 * c + 2] = s.h; e[4 * c + 3] = s.m; }
        }
        if (IS_A) { e[12] = one; e[13] = one; e[14] = one; }
        else {
            const float n = fmaf(pz, pz, fmaf(py, py, px * px));
            const Split3 s = split3(n);
            e[12] = s.h; e[13] = s.m; e[14] = s.l;
        }
    } else if (!IS_A) {
        e[12] = __float2bfloat16_rn(1.0e30f);          // padded database row: never a candidate
    }
    if (frag_layout) {
        // word (kc, row = 32 chunk + 8 nt + g, t) -> uint32 index 256 chunk + 8 (4 g + t) + 2 nt + kc: thread (g, t) of the mma.sync
        // sweep finds its eight fragment words of a 32-row chunk in 32 consecutive bytes
        uint2* w = reinterpret_cast<uint2*>(tile_base) + ((r >> 5) * 128 + (r & 7) * 16 + ((r >> 3) & 3));
#pragma unroll
        for (int t = 0; t < 4; ++t) w[4 * t] = make_uint2(pk(e[2 * t], e[2 * t + 1]), pk(e[8 + 2 * t], e[8 + 2 * t + 1]));
        return;
    }
#pragma unroll
    for (int kc = 0; kc < 2; ++kc)
        tile_base[kc * TM + r] = make_uint4(pk(e[8 * kc], e[8 * kc + 1]), pk(e[8 * kc + 2], e[8 * kc + 3]),
                                            pk(e[8 * kc + 4], e[8 * kc + 5]), pk(e[8 * kc + 6], e[8 * kc + 7]));
}

constexpr int PREP_SPLIT = 4;   // minimum CTAs per batch element (gridDim.y): each finds the centre, then writes every
                                // gridDim.y-th slab of rows; few batch elements get more CTAs each

// grid (batch elements, PREP_SPLIT): bounding-box centre of y, operand tiles of x and y, squared cloud radius
__global__ void __launch_bounds__(256) knn_tc_prepare_kernel(PrepArgs a) {
    __shared__ float red[6][8];
    __shared__ float cen[3];
    __shared__ float rmax[8];
    const int bl = blockIdx.x, b = a.blist ? a.blist[a.b0 + bl] : a.b0 + bl;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, part = blockIdx.y;
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
    for (int i = part * 256 + tid; i < a.ntN * TM; i += 256 * (int)gridDim.y) {
        const bool pad = i >= a.N;
        float px = 0.0f, py = 0.0f, pz = 0.0f;
        if (!pad) { px = xb[3 * (size_t)i] - c0; py = xb[3 * (size_t)i + 1] - c1; pz = xb[3 * (size_t)i + 2] - c2; }
        r2 = fmaxf(r2, fmaf(pz, pz, fmaf(py, py, px * px)));
        write_row<true>(At + (size_t)(i / TM) * TILE_U4, i % TM, px, py, pz, pad);
    }
    for (int j = part * 256 + tid; j < a.ntM * TM; j += 256 * (int)gridDim.y) {
        const bool pad = j >= a.M;
        float px = 0.0f, py = 0.0f, pz = 0.0f;
        if (!pad) { px = yb[3 * (size_t)j] - c0; py = yb[3 * (size_t)j + 1] - c1; pz = yb[3 * (size_t)j + 2] - c2; }
        r2 = fmaxf(r2, fmaf(pz, pz, fmaf(py, py, px * px)));
        write_row<false>(Bt + (size_t)(j / TM) * TILE_U4, j % TM, px, py, pz, pad, a.frag_layout != 0);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) r2 = fmaxf(r2, __shfl_xor_sync(0xffffffffu, r2, off));
    if (lane == 0) rmax[warp] = r2;
    __syncthreads();
    if (tid == 0) {
        float r = rmax[0];
        for (int w = 1; w < 8; ++w) r = fmaxf(r, rmax[w]);
        // squared radius R^2 >= 0 as ordered bits (zeroed before the launch); non-finite input -> +inf: everything is a candidate
        atomicMax(reinterpret_cast<unsigned int*>(a.tol + bl), __float_as_uint(isfinite(r) ? r : INFINITY));
        if (part == 0) { a.centre[3 * bl] = c0; a.centre[3 * bl + 1] = c1; a.centre[3 * bl + 2] = c2; }
    }
}

// ---- spatially sorted operands (N, M <= SORT_MAX) ------------------------------------------------------------
// Both clouds are put into the order of a 16 x 16 x 16 Morton cell grid over y's bounding box (counting sort in
// shared memory; the order inside a cell is whatever the atomics give -- it only affects which candidates the
// epilogue looks at, never the result).  Consecutive runs of 32 points are then compact, and the main kernel can
// skip a 32-column chunk for a warp of 32 query rows whenever the distance between the two runs' bounding boxes
// exceeds what every one of those rows has already found.  Written per batch element: operand tiles in sorted
// order, perm (sorted position -> original index, -1 = padding), sorted copies of the FP32 coordinates (what the
// exact re-rank reads), and the runs' bounding boxes in centred coordinates.
#ifndef PB_TC_SORTED
#define PB_TC_SORTED 0     // measured on C3 (round 2): 47 % of the (warp, chunk) pairs survive, but the sweep is bound by the
#endif                     // MMA / epilogue hand-off latency, not by the minima: 4.17 ms sorted vs 3.81 ms unsorted -> off
constexpr int SORT_MAX = 4096;
constexpr int SORT_CELLS = 4096;
constexpr float BOX_EMPTY = 1.0e30f;

struct SortArgs {
    const float* x; const float* y;
    int N, M, y_shared, ntN, ntM;
    uint4* Aop; uint4* Bop; float* tol; float* centre;
    int* permX; int* permY; float* xs; float* ys; float* xbox; float* ybox;   // chunk-local, per batch element
    int b0;
    const int* blist;
};

__device__ __forceinline__ uint32_t spread4(uint32_t v) {          // 4 bits -> every third bit
    return (v & 1u) | ((v & 2u) << 2) | ((v & 4u) << 4) | ((v & 8u) << 6);
}

// grid (batch elements, 2): blockIdx.y = 0 sorts and writes x, 1 sorts and writes y
__global__ void __launch_bounds__(256) knn_tc_prepare_sorted_kernel(SortArgs a) {
    __shared__ int s_hist[SORT_CELLS];
    __shared__ int s_perm[SORT_MAX];
    __shared__ float red[6][8];
    __shared__ float s_lo[3], s_inv[3], s_cen[3];
    __shared__ int s_wsum[8];
    __shared__ float rmax[8];
    const int bl = blockIdx.x, b = a.blist ? a.blist[a.b0 + bl] : a.b0 + bl;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool is_y = blockIdx.y == 1;
    const float* xb = a.x + (size_t)b * a.N * 3;
    const float* yb = a.y + (a.y_shared ? 0 : (size_t)b * a.M * 3);
    // ---- y's bounding box: centre of the scores, origin and pitch of the cell grid
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
    for (int i = tid; i < SORT_CELLS; i += 256) s_hist[i] = 0;
    __syncthreads();
    if (tid < 3) {
        float l = red[tid][0], h = red[3 + tid][0];
        for (int w = 1; w < 8; ++w) { l = fminf(l, red[tid][w]); h = fmaxf(h, red[3 + tid][w]); }
        const float c = 0.5f * (l + h);
        s_cen[tid] = isfinite(c) ? c : 0.0f;
        s_lo[tid] = isfinite(l) ? l : 0.0f;
        const float ext = h - l;
        s_inv[tid] = (isfinite(ext) && ext > 0.0f) ? 16.0f / ext : 0.0f;
    }
    __syncthreads();
    const float c0 = s_cen[0], c1 = s_cen[1], c2 = s_cen[2];
    const float* pb = is_y ? yb : xb;
    const int n = is_y ? a.M : a.N, nt = is_y ? a.ntM : a.ntN;
    auto cell_of = [&](int i) -> int {
        uint32_t cc[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float f = (pb[3 * (size_t)i + c] - s_lo[c]) * s_inv[c];
            cc[c] = (uint32_t)min(15, max(0, (int)f));               // NaN -> 0; points outside y's box are clamped
        }
        return (int)(spread4(cc[0]) | (spread4(cc[1]) << 1) | (spread4(cc[2]) << 2));
    };
    // ---- counting sort by Morton cell
    for (int i = tid; i < n; i += 256) atomicAdd(&s_hist[cell_of(i)], 1);
    __syncthreads();
    {
        int v[SORT_CELLS / 256], sum = 0;
#pragma unroll
        for (int k = 0; k < SORT_CELLS / 256; ++k) { v[k] = s_hist[tid * (SORT_CELLS / 256) + k]; sum += v[k]; }
        int incl = sum;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += o;
        }
        if (lane == 31) s_wsum[warp] = incl;
        __syncthreads();
        int base = incl - sum;
        for (int w = 0; w < warp; ++w) base += s_wsum[w];
#pragma unroll
        for (int k = 0; k < SORT_CELLS / 256; ++k) { s_hist[tid * (SORT_CELLS / 256) + k] = base; base += v[k]; }
    }
    __syncthreads();
    for (int i = tid; i < n; i += 256) s_perm[atomicAdd(&s_hist[cell_of(i)], 1)] = i;
    __syncthreads();
    // ---- operand rows, permutation, sorted coordinates, run boxes
    uint4* Ot = (is_y ? a.Bop + (size_t)bl * a.ntM * TILE_U4 : a.Aop + (size_t)bl * a.ntN * TILE_U4);
    int* perm = (is_y ? a.permY : a.permX) + (size_t)bl * nt * TM;
    float* cs = (is_y ? a.ys : a.xs) + (size_t)bl * nt * TM * 3;
    float* box = (is_y ? a.ybox : a.xbox) + (size_t)bl * nt * 4 * 6;
    float r2 = 0.0f;
    for (int pos = tid; pos < nt * TM; pos += 256) {
        const bool pad = pos >= n;
        const int i = pad ? -1 : s_perm[pos];
        float vx = 0.0f, vy = 0.0f, vz = 0.0f, px = 0.0f, py = 0.0f, pz = 0.0f;
        if (!pad) {
            vx = pb[3 * (size_t)i]; vy = pb[3 * (size_t)i + 1]; vz = pb[3 * (size_t)i + 2];
            px = vx - c0; py = vy - c1; pz = vz - c2;
        }
        r2 = fmaxf(r2, fmaf(pz, pz, fmaf(py, py, px * px)));
        if (is_y) write_row<false>(Ot + (size_t)(pos / TM) * TILE_U4, pos % TM, px, py, pz, pad);
        else write_row<true>(Ot + (size_t)(pos / TM) * TILE_U4, pos % TM, px, py, pz, pad);
        perm[pos] = i;
        cs[3 * (size_t)pos] = vx; cs[3 * (size_t)pos + 1] = vy; cs[3 * (size_t)pos + 2] = vz;
        float bl3[3] = {pad ? BOX_EMPTY : px, pad ? BOX_EMPTY : py, pad ? BOX_EMPTY : pz};
        float bh3[3] = {pad ? -BOX_EMPTY : px, pad ? -BOX_EMPTY : py, pad ? -BOX_EMPTY : pz};
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                bl3[c] = fminf(bl3[c], __shfl_xor_sync(0xffffffffu, bl3[c], off));
                bh3[c] = fmaxf(bh3[c], __shfl_xor_sync(0xffffffffu, bh3[c], off));
            }
        if (lane == 0) {
            float* o = box + (size_t)(pos >> 5) * 6;
            o[0] = bl3[0]; o[1] = bl3[1]; o[2] = bl3[2]; o[3] = bh3[0]; o[4] = bh3[1]; o[5] = bh3[2];
        }
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) r2 = fmaxf(r2, __shfl_xor_sync(0xffffffffu, r2, off));
    if (lane == 0) rmax[warp] = r2;
    __syncthreads();
    if (tid == 0) {
        float r = rmax[0];
        for (int w = 1; w < 8; ++w) r = fmaxf(r, rmax[w]);
        atomicMax(reinterpret_cast<unsigned int*>(a.tol + bl), __float_as_uint(isfinite(r) ? r : INFINITY));
        if (is_y) { a.centre[3 * bl] = c0; a.centre[3 * bl + 1] = c1; a.centre[3 * bl + 2] = c2; }
    }
}

struct MainArgs {
    const float* x; const float* y;
    int N, M, y_shared, ntN, ntM;
    const uint4* Aop; const uint4* Bop; const float* tol;
    int64_t* idx; float* d2;
    int b0;
    int* err;
    const int* blist;
    // spatially sorted mode (N, M <= SORT_MAX): NULL otherwise
    const int* permX; const int* permY; const float* xs; const float* ys; const float* xbox; const float* ybox;
    const float* centre;
    uint32_t* cand;            // split re-rank: [rows][NWG][8] = count (or ~0: exact scan) + up to 7 entries (first column | groups << 28)
};

constexpr int CAND = 8;            // ring slots per epilogue thread: the CAND - 1 latest flagged chunks + one scratch slot
constexpr int NWG = 2;             // epilogue warpgroups = TMEM accumulators: warpgroup g owns the tiles v = g (mod NWG)
constexpr int EPI_T = NWG * TM;    // 256 epilogue threads: thread = (query row tid & 127, warpgroup tid >> 7)
constexpr int KT2 = EPI_T + 64;    // + warp 8 (MMA issuer, TMEM owner) + warp 9 (bulk-copy producer)
#ifndef PB_TC_NH
#define PB_TC_NH 1
#endif
constexpr int NH = PB_TC_NH;       // accumulators per warpgroup: the tile is issued as NH MMAs of TM / NH columns, so the
                                   // MMA of the next tile's part overlaps the epilogue of the other part
#ifndef PB_TC_NB
#define PB_TC_NB 1
#endif
constexpr int NB = PB_TC_NB;       // accumulators per warpgroup holding WHOLE tiles: while a warpgroup reads one, the MMA of its
                                   // next tile is already in the other -- the hand-off chain (issue -> complete -> commit ->
                                   // wake -> read -> arrive -> wake) was the sweep's period with one (1200 cycles per tile)
static_assert(NH == 1 || NB == 1, "half-tile accumulators and several whole-tile accumulators are not combined");
#ifndef PB_TC_SPLIT
#define PB_TC_SPLIT 0              // 1: the exact re-rank runs as its own thread-per-row kernel (knn_rerank_kernel) instead of as the
#endif                             // tail of every sweep CTA (unsorted sweep only).  Measured on C3 (round 2): 4.24 ms against 3.85 ms
                                   // with the chunk ring, 6.67 ms with the tile ring (sweep 179 us + re-rank 234 us per 256-element
                                   // pass: a thread per row is unbalanced on surface clouds with many near ties) -> off
#ifndef PB_TC_TILE_RING
#define PB_TC_TILE_RING 0          // measured on C3 (round 2): 6.1 ms against 3.85 ms -> off
#endif
// ring step per 128-column tile instead of per 32-column chunk (unsorted sweep): a slot holds the tile's four chunk minima;
// the exact re-rank then looks at 32-column chunks instead of 8-column groups.  A quarter of the sweep's bookkeeping, but 4x
// the exact evaluations in the latency-bound tail that every CTA runs while it still holds its TMEM columns: slower.
constexpr bool TILE_RING = PB_TC_TILE_RING && NH == 1;
constexpr int NACC = NH * NWG * NB;
constexpr int TH = TM / NH;        // columns per accumulator
constexpr int NSTG = 6;            // shared-memory stages of database tiles (24 KB in flight per CTA)
constexpr int TMEM_COLS = NACC * TH;   // 512 columns with NB = 2: one CTA of an SM sweeps at a time, the others are in prologue / tail

__device__ __forceinline__ float fmin3(float a, float b, float c) {
    float d;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
// minimum of 8 consecutive scores: four 3-input minima (FMNMX3)
__device__ __forceinline__ float group_min(const uint32_t* r) {
    const float* v = reinterpret_cast<const float*>(r);
    return fmin3(fmin3(v[0], v[1], v[2]), fmin3(v[3], v[4], v[5]), fminf(v[6], v[7]));
}

// minimum of 32 consecutive scores: sixteen 3-input minima
__device__ __forceinline__ float chunk_min(const uint32_t* r) {
    const float* v = reinterpret_cast<const float*>(r);
    float m[11];
#pragma unroll
    for (int i = 0; i < 10; ++i) m[i] = fmin3(v[3 * i], v[3 * i + 1], v[3 * i + 2]);
    m[10] = fminf(v[30], v[31]);
    const float a0 = fmin3(m[0], m[1], m[2]), a1 = fmin3(m[3], m[4], m[5]), a2 = fmin3(m[6], m[7], m[8]), a3 = fminf(m[9], m[10]);
    return fmin3(fmin3(a0, a1, a2), a3, a3);
}

#ifdef PB_TC_DEBUG
__device__ unsigned long long g_tcdbg[256 * 8];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TCDBG(slot) do { if (blockIdx.x == 0 && blockIdx.y < 256) g_tcdbg[blockIdx.y * 8 + (slot)] = gtime(); } while (0)
__device__ unsigned long long g_tccyc[256 * 8];
#define TCCLK(var) const long long var = clock64()
#define TCACC(k, t0, t1) cyc[k] += (t1) - (t0)
#else
#define TCCLK(var) do {} while (0)
#define TCACC(k, t0, t1) do {} while (0)
#define TCDBG(slot) do {} while (0)
#endif

// One flagged group (8 consecutive columns from j0) of a query row: exact FP32 distances, lowest index on ties.
// perm != NULL: yb holds the database in sorted order and perm[j] is the original index of position j -- ties
// are broken by the ORIGINAL index, as the exact scan does.
__device__ __forceinline__ void exact_group(const float* __restrict__ yb, const int* __restrict__ perm, int M, int j0,
                                            float qx, float qy, float qz, float& best, int& bidx, int& nexact) {
    const float* __restrict__ yp = yb + 3 * (size_t)j0;
    if (j0 + 8 <= M) {                                   // all 24 loads in flight, immediate offsets
        float p[24];
        int jo[8];
#pragma unroll
        for (int k = 0; k < 24; ++k) p[k] = yp[k];
#pragma unroll
        for (int c = 0; c < 8; ++c) jo[c] = perm ? perm[j0 + c] : j0 + c;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const float d = sqdist3(qx, qy, qz, p[3 * c], p[3 * c + 1], p[3 * c + 2]);
            if (d < best || (d == best && jo[c] < bidx)) { best = d; bidx = jo[c]; }
        }
        nexact += 8;
    } else {
        for (int c = 0; j0 + c < M; ++c) {
            const float d = sqdist3(qx, qy, qz, yp[3 * c], yp[3 * c + 1], yp[3 * c + 2]);
            const int jo = perm ? perm[j0 + c] : j0 + c;
            ++nexact;
            if (d < best || (d == best && jo < bidx)) { best = d; bidx = jo; }
        }
    }
}

// grid (row tiles, batch elements of the chunk), 320 threads, two CTAs per SM.  Warp-specialised:
//   warps 0-7  epilogue: warpgroup g = warp / 4 reads accumulator g (its thread t owns query row t & 127);
//   warp 8     issues the MMAs of tile v into accumulator v % NACC, owns the TMEM allocation;
//   warp 9     streams the database tiles through NSTG shared-memory stages with 1-D TMA bulk copies.
// All hand-offs are mbarriers; the tile loop has no CTA-wide barrier.
template <bool SORTED>
__global__ void __launch_bounds__(KT2, 3) knn_tc_kernel(MainArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint4* sA = reinterpret_cast<uint4*>(smem_raw);                    // [TILE_U4]
    uint4* sB = sA + TILE_U4;                                          // [NSTG][TILE_U4]
    float4* lgrp_all = reinterpret_cast<float4*>(sB + NSTG * TILE_U4);          // [CAND][EPI_T] minima of the four 8-column groups
    int* lid_all = reinterpret_cast<int*>(lgrp_all + CAND * EPI_T);             // [CAND][EPI_T] chunk index
    __shared__ __align__(8) uint64_t bar_load[NSTG], bar_sfree[NSTG], bar_mma[NACC], bar_free[NACC], bar_a;
    __shared__ uint32_t tmem_base;
    __shared__ float s_mrun[EPI_T];
    __shared__ float s_best[EPI_T];
    __shared__ int s_bidx[EPI_T];
    __shared__ int s_rowbad[EPI_T];
    __shared__ int s_bad;
    __shared__ __align__(16) float s_lb[4][SORT_MAX / 32];      // sorted mode: box distance^2 of (row warp, 32-column chunk)
    __shared__ int s_ord[SORT_MAX / TM];                        // sorted mode: sweep order of the column tiles, nearest first

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int rt = blockIdx.x, bl = blockIdx.y, b = a.blist ? a.blist[a.b0 + bl] : a.b0 + bl;
    const uint4* Ag = a.Aop + ((size_t)bl * a.ntN + rt) * TILE_U4;
    const uint4* Bg = a.Bop + (size_t)bl * a.ntM * TILE_U4;
    const int ntM = a.ntM;

    if (tid == 0) {
        TCDBG(0);
        for (int i = 0; i < NSTG; ++i) { tc::mbar_init(&bar_load[i], 1); tc::mbar_init(&bar_sfree[i], 1); }
        for (int i = 0; i < NACC; ++i) { tc::mbar_init(&bar_mma[i], 1); tc::mbar_init(&bar_free[i], 4); }
        tc::mbar_init(&bar_a, 1);
        s_bad = 0;
        tc::fence_mbar_init();
        if (SORTED) {       // both clouds follow one Morton order: start at the column tile facing this row tile, then outwards
            int lo = (int)(((long long)rt * ntM) / a.ntN), hi = lo + 1, i = 0;
            lo = lo >= ntM ? ntM - 1 : lo; hi = lo + 1;
            while (i < ntM) {
                if (lo >= 0) s_ord[i++] = lo--;
                if (hi < ntM && i < ntM) s_ord[i++] = hi++;
            }
        }
    }
    constexpr bool sorted = SORTED;
    if (warp == 8) tc::tmem_alloc<TMEM_COLS>(&tmem_base);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = tmem_base;
    bool ok = true;
    if (tid == 0) TCDBG(1);

    if (warp == 9) {
        // ===== producer
        if (lane == 0) {
            tc::mbar_expect_tx(&bar_a, TILE_BYTES);
            tc::bulk_g2s(sA, Ag, TILE_BYTES, &bar_a);
            for (int v = 0; v < ntM; ++v) {
                const int s = v % NSTG;
                if (v >= NSTG) ok = tc::mbar_wait(&bar_sfree[s], (uint32_t)(v / NSTG - 1) & 1u) && ok;
                tc::mbar_expect_tx(&bar_load[s], TILE_BYTES);
                tc::bulk_g2s(sB + s * TILE_U4, Bg + (size_t)(sorted ? s_ord[v] : v) * TILE_U4, TILE_BYTES, &bar_load[s]);
            }
            if (!ok) s_bad = 1;
        }
    } else if (warp == 8) {
        // ===== MMA issuer: D[acc] = A * B[stage, half]^T, two K=16 steps per half tile.  The whole warp runs
        // the loop (waits and descriptor arithmetic stay warp-uniform); one lane issues.
        const uint32_t idesc = tc::idesc_bf16_f32(TM, TH);
        const uint64_t ad0 = tc::smem_desc(tc::smem_u32(sA), TM * 16, 128);
        const uint64_t bd0 = tc::smem_desc(tc::smem_u32(sB), TM * 16, 128);
        ok = tc::mbar_wait(&bar_a, 0);
        if (lane == 0) TCDBG(2);
#ifdef PB_TC_DEBUG
        long long cyc[5] = {0, 0, 0, 0, 0};
#endif
        int s = 0;
        uint32_t lph = 0;
        for (int v = 0; v < ntM; ++v) {
            const int wg = v % NWG, kk = v / NWG, use = kk / NB;
            TCCLK(t0);
            ok = tc::mbar_wait(&bar_load[s], lph) && ok;
            TCCLK(t1);
            TCACC(0, t0, t1);
#pragma unroll
            for (int h = 0; h < NH; ++h) {                     // database rows TH*h .. of the tile
                const int acc = (wg * NB + kk % NB) * NH + h;
                TCCLK(t2);
                if (use > 0) ok = tc::mbar_wait(&bar_free[acc], (uint32_t)(use - 1) & 1u) && ok;
                tc::fence_after_sync();
                TCCLK(t3);
                const uint64_t bd = bd0 + (uint64_t)((s * TILE_BYTES + h * TH * 16) >> 4);
                if (lane == 0) {
                    tc::mma_bf16(tmem + (uint32_t)acc * TH, ad0, bd, idesc, 0);      // the whole K = 16 in one MMA
                    tc::mma_commit(&bar_mma[acc]);             // accumulator ready -> epilogue
                }
                TCCLK(t4);
                TCACC(1, t2, t3); TCACC(2, t3, t4);
            }
            if (lane == 0) tc::mma_commit(&bar_sfree[s]);      // stage consumed -> producer
            if (v == 0 && lane == 0) TCDBG(3);
            if (++s == NSTG) { s = 0; lph ^= 1u; }
        }
        if (lane == 0) TCDBG(4);
#ifdef PB_TC_DEBUG
        if (lane == 0 && blockIdx.x == 0 && blockIdx.y < 256)
            for (int k = 0; k < 4; ++k) g_tccyc[blockIdx.y * 8 + 4 + k] = (unsigned long long)cyc[k];
#endif
        // the epilogue has read the last tile of every accumulator -> the TMEM columns can go
        for (int acc = 0; acc < NACC; ++acc) {
            const int wga = acc / (NB * NH), slot = (acc / NH) % NB;
            const int kwg = (ntM - wga + NWG - 1) / NWG;                  // tiles of that warpgroup
            const int uses = kwg > slot ? (kwg - slot + NB - 1) / NB : 0; // ... that used this accumulator
            if (uses > 0) ok = tc::mbar_wait(&bar_free[acc], (uint32_t)(uses - 1) & 1u) && ok;
        }
        if (!ok) s_bad = 1;
        __syncwarp();
        tc::fence_after_sync();
        tc::tmem_dealloc<TMEM_COLS>(tmem);
    } else {
        // ===== epilogue: one sweep.  Per 32-column chunk: minimum by 3-input minima; a chunk whose minimum
        // is below  running minimum + T  is remembered with the minima of its 8-column groups.  The ring is
        // written without a branch: every chunk lands in the scratch slot cnt % CAND, a flagged one keeps it.
        const int wg = tid >> 7, rloc = tid & (TM - 1);
        const int row = rt * TM + rloc;
        float4* lgrp = lgrp_all + tid;
        int* lid = lid_all + tid;
        const float T = a.tol[bl] * TC_TOL_REL;         // T = 2^-12 R^2
        const float* xb = a.x + (size_t)b * a.N * 3;
        const float* yb = a.y + (a.y_shared ? 0 : (size_t)b * a.M * 3);
        float qx = 0.0f, qy = 0.0f, qz = 0.0f;
        int orow = row < a.N ? row : -1;                // original index of this query row (-1: padding)
        float pn = 0.0f;                                // sorted mode: |x - c|^2, the row constant the scores leave out
        const int* permY = nullptr;
        const float* ydb = yb;                          // what the exact re-rank reads: y, or its sorted copy
        if (sorted) {
            orow = a.permX[((size_t)bl * a.ntN + rt) * TM + rloc];
            const float* xsr = a.xs + (((size_t)bl * a.ntN + rt) * TM + rloc) * 3;
            qx = xsr[0]; qy = xsr[1]; qz = xsr[2];
            permY = a.permY + (size_t)bl * ntM * TM;
            ydb = a.ys + (size_t)bl * ntM * TM * 3;
            const float cx = a.centre[3 * bl], cy = a.centre[3 * bl + 1], cz = a.centre[3 * bl + 2];
            const float px = qx - cx, py = qy - cy, pz = qz - cz;
            pn = fmaf(pz, pz, fmaf(py, py, px * px));
            // box distance^2 between each of this row tile's four 32-row runs and every 32-column run
            const float* xbx = a.xbox + ((size_t)bl * a.ntN + rt) * 4 * 6;
            const float* ybx = a.ybox + (size_t)bl * ntM * 4 * 6;
            for (int e = tid; e < 4 * ntM * 4; e += EPI_T) {
                const int rw = e / (ntM * 4), ch = e - rw * (ntM * 4);
                float d2 = 0.0f;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float g = fmaxf(fmaxf(xbx[rw * 6 + c] - ybx[ch * 6 + 3 + c], ybx[ch * 6 + c] - xbx[rw * 6 + 3 + c]), 0.0f);
                    d2 = fmaf(g, g, d2);
                }
                s_lb[rw][ch] = d2 * 0.99999f;           // rounded down: a lower bound of every pair's true distance^2
            }
            asm volatile("bar.sync 1, %0;" ::"n"(EPI_T) : "memory");
        } else if (row < a.N) {
            qx = xb[3 * (size_t)row]; qy = xb[3 * (size_t)row + 1]; qz = xb[3 * (size_t)row + 2];
        }
        const bool rvalid = orow >= 0;
        const int wq = warp & 3;                        // this warp's run of 32 rows within the row tile
        int nchunks = 0;                                // statistics: 32-column chunks this warp looked at
        float m_run = INFINITY;        // minimum of s~ over the chunks this thread has seen
        float lost = INFINITY;         // smallest chunk minimum that fell out of the ring
        int cnt = 0;
        const uint32_t taddr0 = tmem + ((uint32_t)((warp & 3) * 32) << 16);
#ifdef PB_TC_DEBUG
        long long cyc[5] = {0, 0, 0, 0, 0};
#endif
        for (int v = wg; v < ntM; v += NWG) {
            if (tid == 0 && v == NWG) TCDBG(5);
            const int vt = sorted ? s_ord[v] : v;          // the column tile behind sweep position v
            const int kk = v / NWG, abase = (wg * NB + kk % NB) * NH;      // this tile's accumulator(s)
            const uint32_t ph = (uint32_t)(kk / NB) & 1u;
            const uint32_t taddr = taddr0 + (uint32_t)abase * TH;
            // sorted mode: what every valid row of this warp has found already (+ the band): a chunk whose box is
            // farther than that from the warp's rows cannot hold a nearest neighbour (nor a tie) of any of them
            float cur = INFINITY;
            float4 lb4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (sorted) {
                float w = rvalid ? m_run + pn : -INFINITY;
#pragma unroll
                for (int off = 16; off >= 1; off >>= 1) w = fmaxf(w, __shfl_xor_sync(0xffffffffu, w, off));
                cur = w + T;                               // NaN scores: comparisons below are false, nothing is skipped
                lb4 = *reinterpret_cast<const float4*>(&s_lb[wq][vt * 4]);
            }
            if (!SORTED && TILE_RING) {
                ok = tc::mbar_wait(&bar_mma[abase], ph) && ok;
                tc::fence_after_sync();
                float cmv[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint32_t r[32];
                    tc::tmem_ld32(taddr + c * 32, r);
                    tc::tmem_ld_wait();
                    cmv[c] = chunk_min(r);
                }
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&bar_free[abase]);              // this warp is done with the accumulator
                const float tm = fminf(fminf(cmv[0], cmv[1]), fminf(cmv[2], cmv[3]));
                const int slot = cnt & (CAND - 1);
                lgrp[slot * EPI_T] = make_float4(cmv[0], cmv[1], cmv[2], cmv[3]);
                lid[slot * EPI_T] = vt;
                if (tm < m_run + T) {
                    ++cnt;                                      // keep it; the next slot becomes scratch
                    if (cnt >= CAND) {                          // ... and the flagged tile it still holds is dropped
                        const float4 o = lgrp[(cnt & (CAND - 1)) * EPI_T];
                        lost = fminf(lost, fminf(fminf(o.x, o.y), fminf(o.z, o.w)));
                    }
                }
                m_run = fminf(m_run, tm);
                continue;
            }
#pragma unroll
            for (int half = 0; half < NH; ++half) {
                TCCLK(t0);
                ok = tc::mbar_wait(&bar_mma[abase + half], ph) && ok;
                TCCLK(t1);
                tc::fence_after_sync();
                TCCLK(t2);
                TCACC(0, t0, t1); TCACC(1, t1, t2);
#pragma unroll
                for (int c = 0; c < 4 / NH; ++c) {
                    const int cc = half * (4 / NH) + c;
                    const float lbc = cc == 0 ? lb4.x : cc == 1 ? lb4.y : cc == 2 ? lb4.z : lb4.w;
                    if (SORTED && lbc > cur) continue;     // warp-uniform
                    if (SORTED) ++nchunks;
                    TCCLK(t3);
                    uint32_t r[32];
                    tc::tmem_ld32(taddr + half * TH + c * 32, r);
                    tc::tmem_ld_wait();
                    TCCLK(t4);
                    float g[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) g[q] = group_min(r + q * 8);
                    const float cm = fminf(fminf(g[0], g[1]), fminf(g[2], g[3]));
                    const int slot = cnt & (CAND - 1);
                    lgrp[slot * EPI_T] = make_float4(g[0], g[1], g[2], g[3]);
                    lid[slot * EPI_T] = vt * 4 + cc;
                    if (cm < m_run + T) {
                        ++cnt;                                  // keep it; the next slot becomes scratch
                        if (cnt >= CAND) {                      // ... and the flagged chunk it still holds is dropped
                            const float4 o = lgrp[(cnt & (CAND - 1)) * EPI_T];
                            lost = fminf(lost, fminf(fminf(o.x, o.y), fminf(o.z, o.w)));
                        }
                    }
                    m_run = fminf(m_run, cm);
                    TCCLK(t5);
                    TCACC(2, t3, t4); TCACC(3, t4, t5);
                }
                TCCLK(t6);
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&bar_free[abase + half]);    // this warp is done with the part
                TCCLK(t7);
                TCACC(4, t6, t7);
            }
        }
#ifdef PB_TC_DEBUG
        if (tid == 0 && blockIdx.x == 0 && blockIdx.y < 256)
            for (int k = 0; k < 4; ++k) g_tccyc[blockIdx.y * 8 + k] = (unsigned long long)cyc[k == 3 ? 4 : k] + (k == 2 ? cyc[3] : 0);
#endif
        if (tid == 0) TCDBG(6);
        if (!ok) s_bad = 1;
        s_mrun[tid] = m_run;
        asm volatile("bar.sync 1, %0;" ::"n"(EPI_T) : "memory");
        const bool sweep_ok = s_bad == 0;      // written before the barrier by any epilogue thread that timed out
        // ---- exact FP32 re-rank of the flagged chunks within T of the row minimum (lowest index on ties)
        float lim = s_mrun[rloc];
#pragma unroll
        for (int g = 1; g < NWG; ++g) lim = fminf(lim, s_mrun[rloc + g * TM]);
        lim += T;
        float best = INFINITY;
        int bidx = 0, nexact = 0;
        const bool rowbad = !sweep_ok || lost < lim;
        if (!SORTED && PB_TC_SPLIT) {
            // ---- split re-rank: this thread's qualifying groups (tile ring: chunks) go to the candidate list of its row
            uint32_t* cd = a.cand + (((size_t)bl * a.ntN + rt) * TM + rloc) * (NWG * 8) + wg * 8;
            uint32_t n = 0;
            bool over = rowbad;
            if (rvalid && !rowbad) {
                for (int i = cnt > CAND - 1 ? cnt - (CAND - 1) : 0; i < cnt; ++i) {
                    const int slot = i & (CAND - 1);
                    const float4 g = lgrp[slot * EPI_T];
                    const int id = lid[slot * EPI_T];
                    const float gv[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (gv[q] < lim) {
                            if (n < 7) cd[1 + n] = TILE_RING ? ((uint32_t)(id * TM + q * 32) | (4u << 28)) : ((uint32_t)(id * 32 + q * 8) | (1u << 28));
                            else over = true;
                            ++n;
                        }
                }
            }
            if (rvalid) cd[0] = over ? 0xFFFFFFFFu : n;
        } else {
        if (rvalid && !rowbad) {
            uint32_t flag = 0;                  // bit 4*slot+q: group q of that ring slot is within T of the row minimum
            for (int i = cnt > CAND - 1 ? cnt - (CAND - 1) : 0; i < cnt; ++i) {
                const int slot = i & (CAND - 1);
                const float4 g = lgrp[slot * EPI_T];
                const uint32_t mask = (g.x < lim ? 1u : 0u) | (g.y < lim ? 2u : 0u) | (g.z < lim ? 4u : 0u) | (g.w < lim ? 8u : 0u);
                flag |= mask << (4 * slot);
            }
            while (flag) {                      // lanes walk their own flagged groups (tile ring: chunks) together
                const int pos = __ffs(flag) - 1;
                flag &= flag - 1;
                if (!SORTED && TILE_RING) {
                    const int j0 = lid[(pos >> 2) * EPI_T] * TM + (pos & 3) * 32;
#pragma unroll 1
                    for (int q = 0; q < 4; ++q)
                        if (j0 + 8 * q < a.M) exact_group(ydb, permY, a.M, j0 + 8 * q, qx, qy, qz, best, bidx, nexact);
                } else {
                    const int chunk = lid[(pos >> 2) * EPI_T];
                    exact_group(ydb, permY, a.M, chunk * 32 + (pos & 3) * 8, qx, qy, qz, best, bidx, nexact);
                }
            }
        }
        s_best[tid] = best; s_bidx[tid] = bidx; s_rowbad[tid] = rowbad ? 1 : 0;
        asm volatile("bar.sync 1, %0;" ::"n"(EPI_T) : "memory");
        if (wg == 0 && rvalid) {
            bool bad = rowbad;
#pragma unroll
            for (int g = 1; g < NWG; ++g) {
                const float d = s_best[rloc + g * TM];
                const int j = s_bidx[rloc + g * TM];
                if (d < best || (d == best && j < bidx)) { best = d; bidx = j; }
                bad = bad || s_rowbad[rloc + g * TM] != 0;
            }
            if (bad || best == INFINITY) {                          // ring overflow / timeout / NaN input: exact scan
                best = INFINITY; bidx = 0;
                atomicAdd(a.err + 1, 1);                            // statistics: rows that fell back
                for (int j = 0; j < a.M; ++j) {
                    const float d = sqdist3(qx, qy, qz, yb[3 * (size_t)j], yb[3 * (size_t)j + 1], yb[3 * (size_t)j + 2]);
                    if (d < best) { best = d; bidx = j; }
                }
            }
            if (a.idx) a.idx[(size_t)b * a.N + orow] = bidx;
            if (a.d2) a.d2[(size_t)b * a.N + orow] = best;
        }
        nexact = __reduce_add_sync(0xffffffffu, nexact);
        if (lane == 0 && nexact) atomicAdd(a.err + 2, nexact);      // statistics: exact evaluations
        if (lane == 0 && sorted) atomicAdd(a.err + 3, nchunks);     // statistics: (warp, chunk) pairs not skipped
        }
    }
    __syncthreads();
    if (tid == 0) {
        if (s_bad) *a.err = 1;
        TCDBG(7);
    }
}

// ---- the same contraction through the warp-level tensor path (mma.sync.m16n8k16, BF16 in, FP32 accumulators in the
// registers of the warp that takes the minima): a measured alternative, opt-in (pb_knn_tc_impl(2)) ---------------------
// With mma.sync the accumulators arrive in the evaluating thread's registers: no TMEM, no hand-off between an issuing
// warp and epilogue warps.  One m16n8k16 costs 8.0 cycles per scheduler (scratch/hmma_lab.cu), 64 scores per clock per
// SM -- a floor of 1.6 ms for C3.  Measured: 6.7 ms against 3.9 ms for the tcgen05 sweep above (profiles/r02_knn_mma.md):
// a thread owns 2 columns of every 8-column tile and 4 rows, so the candidate bookkeeping is per (thread, row) with only
// 8 scores per 32-column chunk behind it (4x the tcgen05 kernel's per-score cost) -- done per 128-column tile instead, the
// exact re-rank has 4x the columns to look at and diverges.  Identical indices and distances (tests/test_gpu_knn_tc.py).
//   grid (row tiles, batch elements), 4 consumer warps + 1 producer warp (1-D TMA bulk copies of the database tiles, which
//   the prepare kernel wrote in fragment order, through 4 shared-memory stages): warp w owns query rows 32 w .. 32 w + 31
//   (two row blocks of 16); a thread holds rows g, g + 8, g + 16, g + 24 (g = lane / 4) and, of every 8-column tile,
//   columns 2t, 2t + 1.  Per 128-column tile and row: minimum of the thread's 32 scores; a tile below  running minimum + T
//   keeps a slot of the row's ring (the thread's own running minimum is >= the row's, so every column within T of the row
//   minimum is kept).  After the sweep the row minimum is the minimum over the quad; every remembered tile within T of it
//   is re-evaluated with the exact FP32 formula on the thread's 32 columns; quad-reduce with the lowest index on ties.
constexpr int KM_T = 128;          // consumer threads (4 warps); a fifth warp streams the database tiles
constexpr int KM_THREADS = KM_T + 32;
constexpr int KM_CAND = 8;
constexpr int KM_STG = 4;          // shared-memory stages of database tiles (16 KB in flight per CTA)

__device__ __forceinline__ void hmma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
        : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(0.0f));
}

__global__ void __launch_bounds__(KM_THREADS, 4) knn_mma_kernel(MainArgs a) {
    __shared__ __align__(128) uint4 sB[KM_STG][TILE_U4];   // database tiles in fragment order
    __shared__ float s_min[KM_CAND][4][KM_T];              // ring: tile minimum ...
    __shared__ uint16_t s_id[KM_CAND][4][KM_T];            // ... and tile index, per (slot, row of the thread, thread)
    __shared__ __align__(8) uint64_t bar_full[KM_STG], bar_empty[KM_STG];
    __shared__ int s_bad;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int rt = blockIdx.x, bl = blockIdx.y, b = a.blist ? a.blist[a.b0 + bl] : a.b0 + bl;
    const int ntM = a.ntM;
    if (tid == 0) {
        for (int i = 0; i < KM_STG; ++i) { tc::mbar_init(&bar_full[i], 1); tc::mbar_init(&bar_empty[i], 4); }
        s_bad = 0;
        tc::fence_mbar_init();
    }
    __syncthreads();
    if (warp == 4) {
        // ===== producer: one bulk copy (1-D TMA) per database tile through KM_STG stages
        if (lane == 0) {
            const uint4* Bg = a.Bop + (size_t)bl * ntM * TILE_U4;
            bool ok = true;
            for (int v = 0; v < ntM && ok; ++v) {
                const int st = v % KM_STG;
                if (v >= KM_STG) ok = tc::mbar_wait(&bar_empty[st], (uint32_t)(v / KM_STG - 1) & 1u);
                tc::mbar_expect_tx(&bar_full[st], TILE_BYTES);
                tc::bulk_g2s(sB[st], Bg + (size_t)v * TILE_U4, TILE_BYTES, &bar_full[st]);
            }
            if (!ok) s_bad = 1;
        }
        return;
    }
    const uint32_t* __restrict__ Aw = reinterpret_cast<const uint32_t*>(a.Aop + ((size_t)bl * a.ntN + rt) * TILE_U4);
    const float T = a.tol[bl] * TC_TOL_REL;
    // ---- A fragments: word (kc, row, t) of the (canonical) tile = uint32 index ((kc * 128 + row) * 4 + t)
    uint32_t A[2][4];
#pragma unroll
    for (int rb = 0; rb < 2; ++rb) {
        const int r = 32 * warp + 16 * rb + g;
        A[rb][0] = Aw[(r) * 4 + t];
        A[rb][1] = Aw[(r + 8) * 4 + t];
        A[rb][2] = Aw[(TM + r) * 4 + t];
        A[rb][3] = Aw[(TM + r + 8) * 4 + t];
    }
    float run[4], lost[4];
    int cnt[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { run[k] = INFINITY; lost[k] = INFINITY; cnt[k] = 0; }
    // ---- sweep: per database tile four chunks of 32 columns (thread (g, t)'s eight fragment words of chunk c are the 32 bytes
    // at uint4 index 64 c + 2 (4 g + t) of the stage); one ring step per tile and row
    bool ok = true;
#pragma unroll 1
    for (int v = 0; v < ntM; ++v) {
        const int st = v % KM_STG;
        ok = tc::mbar_wait(&bar_full[st], (uint32_t)(v / KM_STG) & 1u) && ok;
        const uint4* sq = sB[st] + 2 * (4 * g + t);
        float cm[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const uint4 c0 = sq[64 * c], c1 = sq[64 * c + 1];
            const uint32_t bw[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
                for (int rb = 0; rb < 2; ++rb) {
                    float d[4];
                    hmma_bf16(d, A[rb], bw[2 * nt], bw[2 * nt + 1]);
                    cm[2 * rb] = fmin3(cm[2 * rb], d[0], d[1]);
                    cm[2 * rb + 1] = fmin3(cm[2 * rb + 1], d[2], d[3]);
                }
            }
        }
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&bar_empty[st]);            // this warp has read the stage
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int slot = cnt[k] & (KM_CAND - 1);
            s_min[slot][k][tid] = cm[k];
            s_id[slot][k][tid] = (uint16_t)v;
            if (cm[k] < run[k] + T) {
                ++cnt[k];                                          // keep it; the next slot becomes scratch
                if (cnt[k] >= KM_CAND) lost[k] = fminf(lost[k], s_min[cnt[k] & (KM_CAND - 1)][k][tid]);   // ... and what it held is dropped
            }
            run[k] = fminf(run[k], cm[k]);
        }
    }
    if (!ok) s_bad = 1;
    asm volatile("bar.sync 1, %0;" ::"n"(KM_T) : "memory");
    const bool sweep_ok = s_bad == 0;
    // ---- exact FP32 re-rank of the remembered tiles within T of the row minimum (lowest index on ties).  A thread first
    // collects its (row, slot) pairs that qualify in a bit mask and then walks them, so the warp's trip count is the largest
    // number of qualifying tiles of any lane (2 - 4), not rows x ring slots
    const float* xb = a.x + (size_t)b * a.N * 3;
    const float* yb = a.y + (a.y_shared ? 0 : (size_t)b * a.M * 3);
    int nexact = 0;
    float best[4];
    int bidx[4];
    bool bad[4];
    uint32_t todo = 0u;                                            // bit 8 k + slot
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float m = run[k];
        m = fminf(m, __shfl_xor_sync(0xffffffffu, m, 1));
        m = fminf(m, __shfl_xor_sync(0xffffffffu, m, 2));
        const float lim = m + T;
        const int row = rt * TM + 32 * warp + 16 * (k >> 1) + 8 * (k & 1) + g;
        bad[k] = lost[k] < lim || !(m == m) || !sweep_ok;
        best[k] = INFINITY; bidx[k] = 0x7fffffff;
        if (row < a.N)
            for (int i = cnt[k] > KM_CAND - 1 ? cnt[k] - (KM_CAND - 1) : 0; i < cnt[k]; ++i) {
                const int slot = i & (KM_CAND - 1);
                if (s_min[slot][k][tid] < lim) todo |= 1u << (8 * k + slot);
            }
    }
    while (todo) {
        const int bit = __ffs(todo) - 1;
        todo &= todo - 1;
        const int k = bit >> 3, slot = bit & 7;
        const int row = rt * TM + 32 * warp + 16 * (k >> 1) + 8 * (k & 1) + g;
        const float qx = xb[3 * (size_t)row], qy = xb[3 * (size_t)row + 1], qz = xb[3 * (size_t)row + 2];
        const int c0 = (int)s_id[slot][k][tid] * TM + 2 * t;        // this thread's columns of the tile: 8 nt + 2 t + {0, 1}
        float bb = INFINITY;
        int bi = 0x7fffffff;
#pragma unroll 4
        for (int nt = 0; nt < 16; ++nt) {
            const int j = c0 + 8 * nt;
            if (j + 1 < a.M) {
                const float* yp = yb + 3 * (size_t)j;
                const float d0 = sqdist3(qx, qy, qz, yp[0], yp[1], yp[2]), d1 = sqdist3(qx, qy, qz, yp[3], yp[4], yp[5]);
                nexact += 2;
                if (d0 < bb) { bb = d0; bi = j; }                   // columns ascend: a strict test keeps the lowest index
                if (d1 < bb) { bb = d1; bi = j + 1; }
            } else if (j < a.M) {
                const float d0 = sqdist3(qx, qy, qz, yb[3 * (size_t)j], yb[3 * (size_t)j + 1], yb[3 * (size_t)j + 2]);
                ++nexact;
                if (d0 < bb) { bb = d0; bi = j; }
            }
        }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
            if (kk == k && (bb < best[kk] || (bb == best[kk] && bi < bidx[kk]))) { best[kk] = bb; bidx[kk] = bi; }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int row = rt * TM + 32 * warp + 16 * (k >> 1) + 8 * (k & 1) + g;
        const bool valid = row < a.N;
        float bs = best[k];
        int bx = bidx[k];
        bool bd = bad[k];
#pragma unroll
        for (int off = 1; off <= 2; off <<= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, bs, off);
            const int oi = __shfl_xor_sync(0xffffffffu, bx, off);
            const int obad = __shfl_xor_sync(0xffffffffu, (int)bd, off);   // (unconditionally: every lane takes part)
            if (ob < bs || (ob == bs && oi < bx)) { bs = ob; bx = oi; }
            bd = bd || obad != 0;
        }
        if (valid && t == 0) {
            if (bd || bs == INFINITY) {                              // ring overflow / NaN input: exact scan
                const float qx = xb[3 * (size_t)row], qy = xb[3 * (size_t)row + 1], qz = xb[3 * (size_t)row + 2];
                bs = INFINITY; bx = 0;
                atomicAdd(a.err + 1, 1);
                for (int j = 0; j < a.M; ++j) {
                    const float d = sqdist3(qx, qy, qz, yb[3 * (size_t)j], yb[3 * (size_t)j + 1], yb[3 * (size_t)j + 2]);
                    if (d < bs) { bs = d; bx = j; }
                }
            }
            if (a.idx) a.idx[(size_t)b * a.N + row] = bx;
            if (a.d2) a.d2[(size_t)b * a.N + row] = bs;
        }
    }
    nexact = __reduce_add_sync(0xffffffffu, nexact);
    if (lane == 0 && nexact) atomicAdd(a.err + 2, nexact);
    if (tid == 0 && !sweep_ok) *a.err = 1;
}

std::atomic<int> g_knn_impl{0};     // pb_knn_tc_impl
// Split re-rank: one THREAD per query row walks the candidate lists its two epilogue threads left (a handful of 8-column groups or
// 32-column chunks), exact FP32 distances, lowest index on ties; a row whose list overflowed (or whose sweep failed) is scanned.
// Run as its own kernel the dependent, scattered loads overlap across 10^5 rows instead of stretching every sweep CTA's life.
__global__ void __launch_bounds__(256) knn_rerank_kernel(MainArgs a, int nb) {
    const size_t gid = (size_t)blockIdx.x * 256 + threadIdx.x;
    const int rows = a.ntN * TM;
    if (gid >= (size_t)nb * rows) return;
    const int bl = (int)(gid / rows), row = (int)(gid - (size_t)bl * rows);
    if (row >= a.N) return;
    const int b = a.blist ? a.blist[a.b0 + bl] : a.b0 + bl;
    const float* xb = a.x + (size_t)b * a.N * 3;
    const float* yb = a.y + (a.y_shared ? 0 : (size_t)b * a.M * 3);
    const float qx = xb[3 * (size_t)row], qy = xb[3 * (size_t)row + 1], qz = xb[3 * (size_t)row + 2];
    const uint32_t* cd = a.cand + gid * (NWG * 8);
    float best = INFINITY;
    int bidx = 0, nexact = 0;
    bool bad = false;
#pragma unroll
    for (int wg = 0; wg < NWG; ++wg) {
        const uint4 c0 = reinterpret_cast<const uint4*>(cd + wg * 8)[0], c1 = reinterpret_cast<const uint4*>(cd + wg * 8)[1];
        const uint32_t e[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        if (e[0] == 0xFFFFFFFFu) { bad = true; continue; }
#pragma unroll
        for (int i = 0; i < 7; ++i)
            if ((uint32_t)i < e[0]) {
                const int j0 = (int)(e[1 + i] & 0x0FFFFFFFu), ng = (int)(e[1 + i] >> 28);
                for (int q = 0; q < ng; ++q)
                    if (j0 + 8 * q < a.M) exact_group(yb, nullptr, a.M, j0 + 8 * q, qx, qy, qz, best, bidx, nexact);
            }
    }
    if (bad || best == INFINITY) {                                  // ring overflow / timeout / NaN input: exact scan
        best = INFINITY; bidx = 0;
        atomicAdd(a.err + 1, 1);
        for (int j = 0; j < a.M; ++j) {
            const float d = sqdist3(qx, qy, qz, yb[3 * (size_t)j], yb[3 * (size_t)j + 1], yb[3 * (size_t)j + 2]);
            if (d < best) { best = d; bidx = j; }
        }
    }
    if (a.idx) a.idx[(size_t)b * a.N + row] = bidx;
    if (a.d2) a.d2[(size_t)b * a.N + row] = best;
    nexact = __reduce_add_sync(__activemask(), nexact);
    if ((threadIdx.x & 31) == (__ffs(__activemask()) - 1) && nexact) atomicAdd(a.err + 2, nexact);
}

constexpr int TC_CHUNK = 256;      // batch elements per prepare/main pair: operands (~90 MB for C3) stay in L2
constexpr size_t MAIN_SMEM = (size_t)(1 + NSTG) * TILE_BYTES + (size_t)CAND * EPI_T * 20;

inline size_t al128(size_t x) { return (x + 127) / 128 * 128; }
size_t tc_ws_bytes(int B, int N, int M) {
    const int nb = B < TC_CHUNK ? B : TC_CHUNK;
    const size_t ntN = (N + TM - 1) / TM, ntM = (M + TM - 1) / TM;
    size_t n = 128 + (size_t)nb * (ntN + ntM) * TILE_BYTES + al128((size_t)nb * 4 * sizeof(float)) + 128;   // head: int stats[4]
    if (PB_TC_SORTED && N <= SORT_MAX && M <= SORT_MAX)      // sorted mode: perm (4 B) + sorted coordinates (12 B) per point, a box per 32
        n += al128((size_t)nb * (ntN + ntM) * TM * 16) + al128((size_t)nb * (ntN + ntM) * 4 * 6 * sizeof(float)) + 256;
    n += 128 + (size_t)nb * ntN * TM * NWG * 8 * sizeof(uint32_t);   // split re-rank: candidate lists
    return n;
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

// 0 = automatic (the tcgen05 + TMEM sweep: 3.9 ms on C3), 1 = tcgen05, 2 = the warp-level mma.sync sweep (6.7 ms); identical results
extern "C" int pb_knn_tc_impl(int impl) {
    PB_REQUIRE(impl >= 0 && impl <= 2, "pb_knn_tc_impl: impl=%d (0 automatic, 1 tcgen05, 2 mma.sync)", impl);
    g_knn_impl.store(impl, std::memory_order_relaxed);
    return 0;
}

extern "C" size_t pb_knn1_tc_workspace_bytes(int B, int N, int M) {
    if (B <= 0 || N <= 0 || M <= 0) return 0;
    return tc_ws_bytes(B, N, M);
}

// shared by pb_knn1_tc, pb_add_adds_tc and pb_pose_loss_fwd_tc: nearest y[b] point of every x[b] point for the
// batch elements b = blist[0..nlist) (device array; NULL = all of 0..B)
int pb_internal_knn1_tc(const char* who, const float* x, const float* y, int B, int N, int M, int y_shared, int64_t* idx,
                        float* d2, void* ws, size_t ws_bytes, const int* blist, int nlist, cudaStream_t st) {
    const int nl = blist ? nlist : B;
    PB_REQUIRE(M <= (1 << 23), "%s: M too large", who);         // tile indices are kept in 16 bits by the mma.sync sweep
    const size_t need = tc_ws_bytes(B, N, M);
    PB_REQUIRE(ws && ws_bytes >= need, "%s: workspace too small (%zu < %zu)", who, ws_bytes, need);
    PB_REQUIRE(reinterpret_cast<uintptr_t>(ws) % 128 == 0, "%s: workspace must be 128-byte aligned", who);
    const int nb_max = B < TC_CHUNK ? B : TC_CHUNK;
    const int ntN = (N + TM - 1) / TM, ntM = (M + TM - 1) / TM;
    PB_REQUIRE(ntN <= 65535 * 32, "%s: N too large", who);
    char* base = static_cast<char*>(ws);
    int* err = reinterpret_cast<int*>(base);                          // int stats[4] at the head of the workspace
    uint4* Aop = reinterpret_cast<uint4*>(base + 128);
    uint4* Bop = Aop + (size_t)nb_max * ntN * TILE_U4;
    float* tol = reinterpret_cast<float*>(Bop + (size_t)nb_max * ntM * TILE_U4);
    float* centre = tol + nb_max;
    const bool sorted = PB_TC_SORTED && N <= SORT_MAX && M <= SORT_MAX;
    int *permX = nullptr, *permY = nullptr;
    float *xs = nullptr, *ys = nullptr, *xbox = nullptr, *ybox = nullptr;
    if (sorted) {
        char* q = reinterpret_cast<char*>(tol) + al128((size_t)nb_max * 4 * sizeof(float));
        permX = reinterpret_cast<int*>(q); q += (size_t)nb_max * ntN * TM * 4;
        permY = reinterpret_cast<int*>(q); q += (size_t)nb_max * ntM * TM * 4;
        xs = reinterpret_cast<float*>(q); q += (size_t)nb_max * ntN * TM * 12;
        ys = reinterpret_cast<float*>(q); q += (size_t)nb_max * ntM * TM * 12;
        q = reinterpret_cast<char*>(al128(reinterpret_cast<uintptr_t>(q)));
        xbox = reinterpret_cast<float*>(q); q += (size_t)nb_max * ntN * 4 * 6 * sizeof(float);
        ybox = reinterpret_cast<float*>(q);
    }
    uint32_t* cand = reinterpret_cast<uint32_t*>(al128(reinterpret_cast<uintptr_t>(base) + need - 128 - (size_t)nb_max * ntN * TM * NWG * 8 * sizeof(uint32_t)) );
    const int knn_impl = g_knn_impl.load(std::memory_order_relaxed);     // one reading per call: prepare and sweep must agree
    cudaFuncSetAttribute(knn_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MAIN_SMEM);
    cudaFuncSetAttribute(knn_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MAIN_SMEM);
    cudaMemsetAsync(err, 0, 4 * sizeof(int), st);   // [0] barrier timeout, [1] fallback rows, [2] exact evaluations
    for (int b0 = 0; b0 < nl; b0 += TC_CHUNK) {
        const int nb = (nl - b0) < TC_CHUNK ? (nl - b0) : TC_CHUNK;
        cudaMemsetAsync(tol, 0, (size_t)nb * sizeof(float), st);
        if (sorted) {
            SortArgs sa{x, y, N, M, y_shared, ntN, ntM, Aop, Bop, tol, centre, permX, permY, xs, ys, xbox, ybox, b0, blist};
            knn_tc_prepare_sorted_kernel<<<dim3(nb, 2), 256, 0, st>>>(sa);
            if (pb_check_launch("knn_tc_prepare_sorted_kernel")) return -2;
        } else {
            PrepArgs pa{x, y, N, M, y_shared, ntN, ntM, Aop, Bop, tol, centre, b0, blist, knn_impl == 2 ? 1 : 0};
            int split = (4 * pb_sm_count() + nb - 1) / nb;         // about four CTAs per SM in total
            const int max_split = (max(ntN, ntM) * TM + 255) / 256;
            split = split < PREP_SPLIT ? PREP_SPLIT : (split > 64 ? 64 : split);
            split = split > max_split ? max_split : split;
            knn_tc_prepare_kernel<<<dim3(nb, split), 256, 0, st>>>(pa);
            if (pb_check_launch("knn_tc_prepare_kernel")) return -2;
        }
        MainArgs ma{x, y, N, M, y_shared, ntN, ntM, Aop, Bop, tol, idx, d2, b0, err, blist, permX, permY, xs, ys, xbox, ybox, centre, cand};
        if (!sorted && knn_impl == 2) {
            knn_mma_kernel<<<dim3(ntN, nb), KM_THREADS, 0, st>>>(ma);
            if (pb_check_launch("knn_mma_kernel")) return -2;
            continue;
        }
        if (sorted) knn_tc_kernel<true><<<dim3(ntN, nb), KT2, MAIN_SMEM, st>>>(ma);
        else knn_tc_kernel<false><<<dim3(ntN, nb), KT2, MAIN_SMEM, st>>>(ma);
        if (pb_check_launch("knn_tc_kernel")) return -2;
        if (!sorted && PB_TC_SPLIT) {
            const size_t nrows = (size_t)nb * ntN * TM;
            knn_rerank_kernel<<<(unsigned)((nrows + 255) / 256), 256, 0, st>>>(ma, nb);
            if (pb_check_launch("knn_rerank_kernel")) return -2;
        }
    }
    return 0;
}

extern "C" int pb_knn1_tc(const float* x, const float* y, int B, int N, int M, int y_shared, int64_t* idx, float* d2,
                          void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(B >= 0 && N >= 0 && M > 0, "pb_knn1_tc: bad sizes B=%d N=%d M=%d", B, N, M);
    if (B == 0 || N == 0) return 0;
    PB_REQUIRE(x && y && idx, "pb_knn1_tc: null pointer argument");
    return pb_internal_knn1_tc("pb_knn1_tc", x, y, B, N, M, y_shared, idx, d2, ws, ws_bytes, nullptr, 0,
                               static_cast<cudaStream_t>(stream));
}
