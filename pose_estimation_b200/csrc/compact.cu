// compact.cu -- K0: stream the coordinate map, compact foreground pixels, de-normalise (sm_100a).
//
// Replaces tools/trainer.py:386-421 (whole-map D2H copy, xyz*extent+lfborder, gather at `choose`)
// and the dataset-side mask.flatten().nonzero() (dataset/linemod/batchdataset.py:673).
// HBM-bound: every pixel costs 12 B of xyz + 1 B of mask, read exactly once with 128-bit loads,
// next iteration's loads issued before the current block scan (register double buffer).
// Kernels:
//   compact_stream_kernel (mask given, H*W % 16 == 0): MASK-FIRST, one pass, no frame-sized state.
//     The batch's masks are one stream of 64 KB tiles (65,536 px) handed out by an atomic ticket to a
//     persistent grid; a CTA turns its tile into 4096 16-bit foreground words in shared memory (two
//     waves of eight 128-bit loads per thread); most tiles are empty and end there.  A non-empty tile
//     publishes its count, sums the counts of the earlier tiles of its frame (decoupled look-back over
//     <= 4 words: tickets are taken in execution order, so a predecessor is always held by a running
//     CTA), and emits its pixels BY RANK -- thread r finds the r-th foreground pixel of the tile by a
//     binary search over the per-word prefix -- so the three plane loads and the five stores of
//     consecutive output slots are issued by consecutive threads whatever the mask's density.  The xyz planes are
//     read ONLY at foreground pixels: 0.3-0.4 MB per 480x640 frame instead of 4.0 MB.  The pointers may
//     be device memory or pinned (mapped) host memory: with host maps the same kernel pulls just the
//     mask and the touched sectors across PCIe (zero-copy gather).
//   compact_sparse_kernel: the round-1 mask-first kernel (CTA per frame, banded bitmap), kept for
//     callers that pass no workspace.
//   compact_kernel (no mask: foreground iff any channel != 0): streams all three planes.
// Output order is row-major pixel order (deterministic): per-iteration block scan, one
// __syncthreads per 1024 pixels.  When there are too few frames to fill the GPU a frame is split
// into S row bands; each band writes its own segment and a tiny merge kernel concatenates them.
#include "common.cuh"

namespace {

constexpr int CT = 256;  // threads per CTA

struct CompactArgs {
    const float* xyz;
    const uint8_t* mask;
    const int32_t* roi;
    const float* extent;
    const float* lfborder;
    float* pts;        // S==1: [F,5,cap] ; S>1: segments [F*S,5,segcap]
    int32_t* count;    // S==1: [F]
    int32_t* total;    // S==1: [F] optional
    int32_t* seg_total;  // S>1: [F*S]
    int HW, Ww, cap, S, quads_per_split, segcap;
};

__device__ __forceinline__ float4 ldg_stream_f4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t ldg_stream_u32(const uint32_t* p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}

template <bool HAS_MASK>
__global__ void __launch_bounds__(CT) compact_kernel(CompactArgs a) {
    const int f = blockIdx.x / a.S, s = blockIdx.x % a.S;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nq = a.HW >> 2;
    const int q0 = s * a.quads_per_split;
    const int q1 = min(nq, q0 + a.quads_per_split);
    const float4* px = reinterpret_cast<const float4*>(a.xyz + (size_t)f * 3 * a.HW);
    const float4* py = px + nq;
    const float4* pz = py + nq;
    const uint32_t* pm = HAS_MASK ? reinterpret_cast<const uint32_t*>(a.mask + (size_t)f * a.HW) : nullptr;
    const float e0 = a.extent[3 * f], e1 = a.extent[3 * f + 1], e2 = a.extent[3 * f + 2];
    const float l0 = a.lfborder[3 * f], l1 = a.lfborder[3 * f + 1], l2 = a.lfborder[3 * f + 2];
    const int rmin = a.roi ? a.roi[2 * f] : 0, cmin = a.roi ? a.roi[2 * f + 1] : 0;
    const int ocap = a.S == 1 ? a.cap : a.segcap;
    float* oX = a.pts + (size_t)blockIdx.x * 5 * ocap;
    float* oY = oX + ocap; float* oZ = oY + ocap; float* oU = oZ + ocap; float* oV = oU + ocap;

    __shared__ int wtot[2][CT / 32];
    int running = 0;
    int buf = 0;

    float4 vx, vy, vz; uint32_t vm = 0;
    int q = q0 + tid;
    bool have = q < q1;
    if (have) {
        vx = ldg_stream_f4(px + q); vy = ldg_stream_f4(py + q); vz = ldg_stream_f4(pz + q);
        if (HAS_MASK) vm = ldg_stream_u32(pm + q);
    }
    for (int qb = q0; qb < q1; qb += CT) {
        // current registers
        const float4 cxv = vx, cyv = vy, czv = vz; const uint32_t cm = vm; const bool chave = have;
        const int cq = qb + tid;
        // prefetch next
        q = qb + CT + tid;
        have = q < q1;
        if (have) {
            vx = ldg_stream_f4(px + q); vy = ldg_stream_f4(py + q); vz = ldg_stream_f4(pz + q);
            if (HAS_MASK) vm = ldg_stream_u32(pm + q);
        }
        unsigned bits = 0;
        if (chave) {
            if (HAS_MASK) {
                bits = ((cm & 0xFFu) ? 1u : 0u) | ((cm & 0xFF00u) ? 2u : 0u) | ((cm & 0xFF0000u) ? 4u : 0u) |
                       ((cm & 0xFF000000u) ? 8u : 0u);
            } else {
                bits = ((cxv.x != 0.0f || cyv.x != 0.0f || czv.x != 0.0f) ? 1u : 0u) |
                       ((cxv.y != 0.0f || cyv.y != 0.0f || czv.y != 0.0f) ? 2u : 0u) |
                       ((cxv.z != 0.0f || cyv.z != 0.0f || czv.z != 0.0f) ? 4u : 0u) |
                       ((cxv.w != 0.0f || cyv.w != 0.0f || czv.w != 0.0f) ? 8u : 0u);
            }
        }
        const int c = __popc(bits);
        int incl = c;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            int o = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += o;
        }
        if (lane == 31) wtot[buf][warp] = incl;
        __syncthreads();
        int woff = 0, btot = 0;
#pragma unroll
        for (int w = 0; w < CT / 32; ++w) {
            int v = wtot[buf][w];
            if (w < warp) woff += v;
            btot += v;
        }
        if (bits) {
            int pos = running + woff + incl - c;
            const float xs[4] = {cxv.x, cxv.y, cxv.z, cxv.w};
            const float ys[4] = {cyv.x, cyv.y, cyv.z, cyv.w};
            const float zs[4] = {czv.x, czv.y, czv.z, czv.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (bits & (1u << k)) {
                    if (pos < ocap) {
                        const int p = 4 * cq + k;
                        const int row = p / a.Ww, col = p - row * a.Ww;
                        oX[pos] = fmaf(xs[k], e0, l0);
                        oY[pos] = fmaf(ys[k], e1, l1);
                        oZ[pos] = fmaf(zs[k], e2, l2);
                        oU[pos] = (float)(col + cmin);
                        oV[pos] = (float)(row + rmin);
                    }
                    ++pos;
                }
            }
        }
        running += btot;
        buf ^= 1;
    }
    if (tid == 0) {
        if (a.S == 1) {
            a.count[f] = min(running, a.cap);
            if (a.total) a.total[f] = running;
        } else {
            a.seg_total[blockIdx.x] = running;
        }
    }
}

// ---- mask-first sparse variant ------------------------------------------------------------------
__device__ __forceinline__ uint4 ldg_stream_u4(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float ldg_stream_f32(const float* p) {
    float r;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
// 4 mask bytes -> 4 bits (byte != 0): per-byte compare (0xFF per non-zero byte), keep one bit per
// byte, gather the four bits with a multiply (bit k of the result's top byte <- byte k)
__device__ __forceinline__ uint32_t nz4(uint32_t w) {
    return ((__vcmpne4(w, 0u) & 0x01010101u) * 0x10204080u) >> 28;
}

constexpr int SP_UNROLL = 4;   // 128-bit mask loads in flight per thread

// One CTA per frame, the frame walked in bands of SP_BAND_WORDS bitmap words (76,800 px) so the
// bitmap is ~10 KB and 8 CTAs fit per SM: a 1024-frame batch is ONE resident wave.  Per band:
// phase 1 mask -> shared bitmap (1 bit/px); phase 2 per-tile (32 words) counts, exclusive prefix,
// then every warp emits (pixel index, u, v) for its round-robin tiles at the tile's offset, i.e.
// in pixel order.  After the last band, phase 3: all threads gather xyz at the emitted pixels and
// de-normalise.
constexpr int SP_BAND_WORDS = 2400;
constexpr int SP_BAND_TILES = (SP_BAND_WORDS + 31) / 32;

__global__ void __launch_bounds__(CT) compact_sparse_kernel(CompactArgs a) {
    __shared__ __align__(16) uint32_t sbm[SP_BAND_WORDS];
    __shared__ int stile[SP_BAND_TILES];
    __shared__ int stotal;
    const int f = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwords = a.HW >> 5;
    const uint4* pm = reinterpret_cast<const uint4*>(a.mask + (size_t)f * a.HW);
    uint16_t* sb16 = reinterpret_cast<uint16_t*>(sbm);
    const int rmin = a.roi ? a.roi[2 * f] : 0, cmin = a.roi ? a.roi[2 * f + 1] : 0;
    float* oX = a.pts + (size_t)f * 5 * a.cap;
    float* oY = oX + a.cap; float* oZ = oY + a.cap; float* oU = oZ + a.cap; float* oV = oU + a.cap;

    int total = 0;                                              // foreground pixels before this band
    for (int wb0 = 0; wb0 < nwords; wb0 += SP_BAND_WORDS) {
        const int bw = min(SP_BAND_WORDS, nwords - wb0);        // words in this band
        const int nv = bw * 2;                                  // 16-px groups (one uint4 of mask each)
        const uint4* pmb = pm + (size_t)wb0 * 2;
        // ---- phase 1
        for (int g0 = 0; g0 < nv; g0 += CT * SP_UNROLL) {
            uint4 m[SP_UNROLL];
#pragma unroll
            for (int k = 0; k < SP_UNROLL; ++k) {
                const int g = g0 + k * CT + tid;
                m[k] = g < nv ? ldg_stream_u4(pmb + g) : make_uint4(0u, 0u, 0u, 0u);
            }
#pragma unroll
            for (int k = 0; k < SP_UNROLL; ++k) {
                const int g = g0 + k * CT + tid;
                if (g < nv) {
                    uint32_t bits = 0u;
                    if (m[k].x | m[k].y | m[k].z | m[k].w)      // most of a frame is background
                        bits = nz4(m[k].x) | (nz4(m[k].y) << 4) | (nz4(m[k].z) << 8) | (nz4(m[k].w) << 12);
                    sb16[g] = (uint16_t)bits;
                }
            }
        }
        __syncthreads();
        // ---- phase 2: tiles of 32 bitmap words (1024 px) dealt round-robin to the warps, so a
        // compact foreground blob (a few dozen image rows) is spread over all warps
        const int ntiles = (bw + 31) >> 5;
        for (int k = warp; k < ntiles; k += CT / 32) {
            const int w = (k << 5) + lane;
            const int c = __reduce_add_sync(0xffffffffu, w < bw ? __popc(sbm[w]) : 0);
            if (lane == 0) stile[k] = c;
        }
        __syncthreads();
        if (warp == 0) {   // exclusive prefix over the tile counts (in place), running total -> stotal
            int run = total;
            for (int k0 = 0; k0 < ntiles; k0 += 32) {
                const int k = k0 + lane;
                const int c = k < ntiles ? stile[k] : 0;
                int incl = c;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) {
                    const int o = __shfl_up_sync(0xffffffffu, incl, off);
                    if (lane >= off) incl += o;
                }
                if (k < ntiles) stile[k] = run + incl - c;
                run += __shfl_sync(0xffffffffu, incl, 31);
            }
            if (lane == 0) stotal = run;
        }
        __syncthreads();
        const int btotal = stotal;
        if (btotal > total && total < a.cap) {                  // uniform: this band has foreground below cap
            for (int k = warp; k < ntiles; k += CT / 32) {
                const int base = stile[k];
                const int next = (k + 1 < ntiles) ? stile[k + 1] : btotal;
                if (next == base || base >= a.cap) continue;    // empty tile / beyond cap (warp-uniform)
                const int w = (k << 5) + lane;
                uint32_t bits = w < bw ? sbm[w] : 0u;
                const int n = __popc(bits);
                int incl = n;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) {
                    const int o = __shfl_up_sync(0xffffffffu, incl, off);
                    if (lane >= off) incl += o;
                }
                // one non-empty word at a time, its 32 pixels across the lanes (a blob row is a few words:
                // walking each lane's own bits would serialise up to 32 steps on one or two live lanes)
                const int wbase = base + incl - n;
                uint32_t live = __ballot_sync(0xffffffffu, n > 0);
                while (live) {
                    const int src = __ffs(live) - 1;
                    live &= live - 1u;
                    const uint32_t wbits = __shfl_sync(0xffffffffu, bits, src);
                    const int wpos = __shfl_sync(0xffffffffu, wbase, src);
                    if ((wbits >> lane) & 1u) {
                        const int pos = wpos + __popc(wbits & ((1u << lane) - 1u));
                        if (pos < a.cap) {
                            const int p = ((wb0 + (k << 5) + src) << 5) + lane;
                            const int row = p / a.Ww, col = p - row * a.Ww;
                            oX[pos] = __int_as_float(p);
                            oU[pos] = (float)(col + cmin);
                            oV[pos] = (float)(row + rmin);
                        }
                    }
                }
            }
        }
        total = btotal;
        __syncthreads();   // bitmap / tile offsets are rewritten by the next band; also publishes the
                           // pixel indices written by this CTA to all of its threads
    }

    // ---- phase 3: gather + de-normalise
    const int n = min(total, a.cap);
    const float* px = a.xyz + (size_t)f * 3 * a.HW;
    const float* py = px + a.HW;
    const float* pz = py + a.HW;
    const float e0 = a.extent[3 * f], e1 = a.extent[3 * f + 1], e2 = a.extent[3 * f + 2];
    const float l0 = a.lfborder[3 * f], l1 = a.lfborder[3 * f + 1], l2 = a.lfborder[3 * f + 2];
    for (int i = tid; i < n; i += CT) {
        const int p = __float_as_int(oX[i]);
        const float x = ldg_stream_f32(px + p), y = ldg_stream_f32(py + p), z = ldg_stream_f32(pz + p);
        oX[i] = fmaf(x, e0, l0);
        oY[i] = fmaf(y, e1, l1);
        oZ[i] = fmaf(z, e2, l2);
    }
    if (tid == 0) {
        a.count[f] = n;
        if (a.total) a.total[f] = total;
    }
}

// ---- mask-first, streaming: tiles by ticket + decoupled look-back ---------------------------------
#ifndef PB_SM_G
#define PB_SM_G 8
#endif
#ifndef PB_SM_W
#define PB_SM_W 2
#endif
#ifndef PB_SM_OCC
#define PB_SM_OCC 6
#endif
#ifndef PB_SM_PF
#define PB_SM_PF 0
#endif
// Measured on a C2-shaped batch (scratch/k0_lab.cu, 1024 frames): G x W x OCC = 8 x 2 x 6 0.086 ms, 4 x 2 x 8 0.088,
// 4 x 4 x 6 0.092, 4 x 4 x 8 0.107, 16 KB tiles (4 x 1 x 8) 0.098; prefetching the next wave (PB_SM_PF) does not help.
constexpr int SM_G = PB_SM_G;                    // 128-bit mask loads per thread per wave (all in flight together)
constexpr int SM_W = PB_SM_W;                    // waves per tile
constexpr int SM_OCC = PB_SM_OCC;                // resident CTAs per SM the kernel is compiled for
constexpr int SM_WGROUPS = CT * SM_G;            // 16-px groups per wave (1024)
constexpr int SM_GROUPS = SM_WGROUPS * SM_W;     // 16-px groups per tile (4096)
constexpr int SM_TILE_PX = SM_GROUPS * 16;       // 65,536 px = 64 KB of mask
constexpr int SM_GPT = SM_GROUPS / CT;           // groups per thread in the rank scan (16 consecutive ones)
constexpr uint32_t SM_FLAG = 0x80000000u;        // tile state word: FLAG | foreground count
static_assert(SM_GPT % 8 == 0, "the rank scan reads its groups' 16-bit words as 128-bit shared-memory loads");

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t r;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// position of the n-th (0-based) set bit of a 16-bit word that has more than n bits set
__device__ __forceinline__ int nth_bit16(uint32_t x, int n) {
    int b = 0, c = __popc(x & 0xFFu);
    if (n >= c) { n -= c; x >>= 8; b = 8; }
    c = __popc(x & 0xFu);
    if (n >= c) { n -= c; x >>= 4; b += 4; }
    c = __popc(x & 0x3u);
    if (n >= c) { n -= c; x >>= 2; b += 2; }
    return b + ((n >= (int)(x & 1u)) ? 1 : 0);
}

__global__ void __launch_bounds__(CT, SM_OCC) compact_stream_kernel(CompactArgs a, uint32_t* __restrict__ state,
                                                                    unsigned int* __restrict__ ticket, int tpf, int ntiles) {
    __shared__ int s_gpre[SM_GROUPS];                    // exclusive prefix of the groups' foreground counts (tile-local)
    __shared__ __align__(16) uint16_t s_gbits[SM_GROUPS];
    __shared__ int s_wcnt[CT / 32];
    __shared__ int s_next;
    __shared__ int s_fbase;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gpf = a.HW >> 4;                           // 16-px groups per frame

    for (;;) {
        // A ticket is taken only when the CTA is ready to process the tile: its count is then published a few load
        // latencies later, whatever other tiles are waiting for -- a ticket held across another tile's look-back
        // (to prefetch) chains the CTAs into a wavefront (measured: 10x slower).  Latency is hidden by occupancy.
        if (tid == 0) s_next = (int)atomicAdd(ticket, 1u);
        __syncthreads();
        const int tile = s_next;
        if (tile >= ntiles) break;
        const int f = tile / tpf, ti = tile - f * tpf;
        const uint4* pm = reinterpret_cast<const uint4*>(a.mask + (size_t)f * a.HW) + (size_t)ti * SM_GROUPS;
        const int ng = min(SM_GROUPS, gpf - ti * SM_GROUPS);
        // ---- mask -> 16-bit foreground words in shared memory, SM_W waves of SM_G 128-bit loads per thread
        uint32_t seen = 0u;
        auto load_wave = [&](int w, uint4 (&m)[SM_G]) {
#pragma unroll
            for (int k = 0; k < SM_G; ++k) {
                const int g = w * SM_WGROUPS + warp * (32 * SM_G) + k * 32 + lane;   // a load instruction covers 512 contiguous bytes
                m[k] = g < ng ? ldg_stream_u4(pm + g) : make_uint4(0u, 0u, 0u, 0u);
            }
        };
        auto bits_wave = [&](int w, const uint4 (&m)[SM_G]) {
#pragma unroll
            for (int k = 0; k < SM_G; ++k) {
                const int g = w * SM_WGROUPS + warp * (32 * SM_G) + k * 32 + lane;
                uint32_t bits = 0u;
                if (m[k].x | m[k].y | m[k].z | m[k].w)                  // most of a frame is background
                    bits = nz4(m[k].x) | (nz4(m[k].y) << 4) | (nz4(m[k].z) << 8) | (nz4(m[k].w) << 12);
                s_gbits[g] = (uint16_t)bits;
                seen |= bits;
            }
        };
#if PB_SM_PF
        {   // wave w + 1 is in flight while wave w is turned into bits (two register buffers)
            uint4 ma[SM_G], mb[SM_G];
            load_wave(0, ma);
#pragma unroll
            for (int w = 0; w < SM_W; w += 2) {
                if (w + 1 < SM_W) load_wave(w + 1, mb);
                bits_wave(w, ma);
                if (w + 2 < SM_W) load_wave(w + 2, ma);
                if (w + 1 < SM_W) bits_wave(w + 1, mb);
            }
        }
#else
#pragma unroll 1
        for (int w = 0; w < SM_W; ++w) {
            uint4 m[SM_G];
            load_wave(w, m);
            bits_wave(w, m);
        }
#endif
        const bool last = ti == tpf - 1;                                // the frame's last tile writes count[f]
        const int any = __syncthreads_or((int)seen);                    // (s_next is rewritten only after this barrier)
        if (!any && !last) {
            if (tid == 0) st_volatile_u32(state + tile, SM_FLAG);
            continue;
        }
        // ---- tile-local ranks: thread t scans the SM_GPT consecutive groups t*SM_GPT ..
        uint32_t wbits[SM_GPT / 2];                                     // two 16-bit words per register
        {
            const uint4* sb = reinterpret_cast<const uint4*>(s_gbits) + tid * (SM_GPT / 8);
#pragma unroll
            for (int q = 0; q < SM_GPT / 8; ++q) {
                const uint4 v = sb[q];
                wbits[4 * q] = v.x; wbits[4 * q + 1] = v.y; wbits[4 * q + 2] = v.z; wbits[4 * q + 3] = v.w;
            }
        }
        int mine = 0;
#pragma unroll
        for (int j = 0; j < SM_GPT / 2; ++j) mine += __popc(wbits[j]);
        int incl = mine;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += o;
        }
        if (lane == 31) s_wcnt[warp] = incl;
        __syncthreads();
        int total = 0, wbase = 0;
#pragma unroll
        for (int w = 0; w < CT / 32; ++w) {
            const int v = s_wcnt[w];
            wbase += w < warp ? v : 0;
            total += v;
        }
        if (tid == 0) st_volatile_u32(state + tile, SM_FLAG | (uint32_t)total);
        {
            int run = wbase + incl - mine;
#pragma unroll
            for (int j = 0; j < SM_GPT / 2; ++j) {
                s_gpre[tid * SM_GPT + 2 * j] = run;
                run += __popc(wbits[j] & 0xFFFFu);
                s_gpre[tid * SM_GPT + 2 * j + 1] = run;
                run += __popc(wbits[j] >> 16);
            }
        }
        // ---- look-back: foreground pixels of the frame's earlier tiles (warp 0; lanes poll one word each)
        if (warp == 0) {
            int sum = 0;
            for (int j0 = 0; j0 < ti; j0 += 32) {
                const int j = j0 + lane;
                uint32_t v = SM_FLAG;
                if (j < ti) {
                    const uint32_t* sp = state + (tile - ti + j);
                    do { v = ld_volatile_u32(sp); } while (!(v & SM_FLAG));
                }
                sum += (int)(v & ~SM_FLAG);
            }
            sum = __reduce_add_sync(0xffffffffu, sum);
            if (lane == 0) s_fbase = sum;
        }
        __syncthreads();
        const int fbase = s_fbase;
        // ---- emit by rank: consecutive threads -> consecutive output slots
        const int nout = min(total, a.cap - fbase);                     // <= 0: the frame is already at cap
        if (nout > 0) {
            const float* px = a.xyz + (size_t)f * 3 * a.HW;
            const float* py = px + a.HW;
            const float* pz = py + a.HW;
            const float e0 = a.extent[3 * f], e1 = a.extent[3 * f + 1], e2 = a.extent[3 * f + 2];
            const float l0 = a.lfborder[3 * f], l1 = a.lfborder[3 * f + 1], l2 = a.lfborder[3 * f + 2];
            const int rmin = a.roi ? a.roi[2 * f] : 0, cmin = a.roi ? a.roi[2 * f + 1] : 0;
            float* oX = a.pts + (size_t)f * 5 * a.cap + fbase;
            float* oY = oX + a.cap; float* oZ = oY + a.cap; float* oU = oZ + a.cap; float* oV = oU + a.cap;
            const int pix0 = ti * SM_TILE_PX;
            for (int r = tid; r < nout; r += CT) {
                int lo = 0, hi = SM_GROUPS;                             // last group with s_gpre <= r
#pragma unroll 1
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (s_gpre[mid] <= r) lo = mid; else hi = mid;
                }
                const int p = pix0 + (lo << 4) + nth_bit16(s_gbits[lo], r - s_gpre[lo]);
                const float x = ldg_stream_f32(px + p), y = ldg_stream_f32(py + p), z = ldg_stream_f32(pz + p);
                const int row = p / a.Ww, col = p - row * a.Ww;
                oX[r] = fmaf(x, e0, l0);
                oY[r] = fmaf(y, e1, l1);
                oZ[r] = fmaf(z, e2, l2);
                oU[r] = (float)(col + cmin);
                oV[r] = (float)(row + rmin);
            }
        }
        if (last && tid == 0) {
            a.count[f] = min(fbase + total, a.cap);
            if (a.total) a.total[f] = fbase + total;
        }
        __syncthreads();                                                // s_gpre / s_gbits / s_wcnt / s_fbase are reused
    }
}

// generic (any HW, any alignment) scalar variant: one pixel per thread per iteration
template <bool HAS_MASK>
__global__ void __launch_bounds__(CT) compact_kernel_scalar(CompactArgs a) {
    const int f = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* px = a.xyz + (size_t)f * 3 * a.HW;
    const float* py = px + a.HW;
    const float* pz = py + a.HW;
    const uint8_t* pm = HAS_MASK ? a.mask + (size_t)f * a.HW : nullptr;
    const float e0 = a.extent[3 * f], e1 = a.extent[3 * f + 1], e2 = a.extent[3 * f + 2];
    const float l0 = a.lfborder[3 * f], l1 = a.lfborder[3 * f + 1], l2 = a.lfborder[3 * f + 2];
    const int rmin = a.roi ? a.roi[2 * f] : 0, cmin = a.roi ? a.roi[2 * f + 1] : 0;
    float* oX = a.pts + (size_t)f * 5 * a.cap;
    float* oY = oX + a.cap; float* oZ = oY + a.cap; float* oU = oZ + a.cap; float* oV = oU + a.cap;
    __shared__ int wtot[2][CT / 32];
    int running = 0, buf = 0;
    for (int pb = 0; pb < a.HW; pb += CT) {
        const int p = pb + tid;
        float x = 0.0f, y = 0.0f, z = 0.0f;
        bool m = false;
        if (p < a.HW) {
            x = px[p]; y = py[p]; z = pz[p];
            m = HAS_MASK ? (pm[p] != 0) : (x != 0.0f || y != 0.0f || z != 0.0f);
        }
        unsigned b = __ballot_sync(0xffffffffu, m);
        if (lane == 0) wtot[buf][warp] = __popc(b);
        __syncthreads();
        int woff = 0, btot = 0;
#pragma unroll
        for (int w = 0; w < CT / 32; ++w) {
            int v = wtot[buf][w];
            if (w < warp) woff += v;
            btot += v;
        }
        if (m) {
            int pos = running + woff + __popc(b & ((1u << lane) - 1u));
            if (pos < a.cap) {
                const int row = p / a.Ww, col = p - row * a.Ww;
                oX[pos] = fmaf(x, e0, l0); oY[pos] = fmaf(y, e1, l1); oZ[pos] = fmaf(z, e2, l2);
                oU[pos] = (float)(col + cmin); oV[pos] = (float)(row + rmin);
            }
        }
        running += btot;
        buf ^= 1;
    }
    if (tid == 0) {
        a.count[f] = min(running, a.cap);
        if (a.total) a.total[f] = running;
    }
}

__global__ void __launch_bounds__(CT) merge_segments_kernel(const float* seg, const int32_t* seg_total, int S,
                                                            int segcap, int cap, float* pts, int32_t* count,
                                                            int32_t* total) {
    const int f = blockIdx.x;
    int base = 0;
    for (int s = 0; s < S; ++s) {
        const int n_all = seg_total[f * S + s];
        const int n = min(n_all, segcap);
        const float* src = seg + (size_t)(f * S + s) * 5 * segcap;
        for (int i = threadIdx.x; i < n; i += CT) {
            const int o = base + i;
            if (o < cap) {
#pragma unroll
                for (int c = 0; c < 5; ++c) pts[((size_t)f * 5 + c) * cap + o] = src[(size_t)c * segcap + i];
            }
        }
        base += n_all;
    }
    if (threadIdx.x == 0) {
        count[f] = min(base, cap);
        if (total) total[f] = base;
    }
}

// choose-list gather (tools/trainer.py:406-421): one thread per list entry.  IDX = int64_t (the reference's
// `choose`) or int32_t; xmap / ymap NULL -> pixel coordinates from the index itself (col + cmin, row + rmin);
// n_valid NULL -> every list has NP entries (the reference's wrap-padded lists), else list f has n_valid[f].
// xyz may be device memory or pinned host memory (then only the touched sectors cross PCIe).
template <typename IDX>
__global__ void __launch_bounds__(128) gather_kernel(const float* __restrict__ xyz, const IDX* __restrict__ choose,
                                                     const float* __restrict__ xmap, const float* __restrict__ ymap,
                                                     const int32_t* __restrict__ n_valid, const int32_t* __restrict__ roi,
                                                     const float* __restrict__ extent, const float* __restrict__ lfborder,
                                                     int HW, int Ww, int NP, int cap, float* __restrict__ pts,
                                                     int32_t* __restrict__ count) {
    const int f = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    int n = n_valid ? n_valid[f] : NP;
    n = max(0, min(n, min(NP, cap)));
    if (i == 0) count[f] = n;
    if (i >= n) return;
    long long p = (long long)choose[(size_t)f * NP + i];
    if (p < 0) p = 0;
    if (p >= HW) p = HW - 1;
    const float* base = xyz + (size_t)f * 3 * HW;
    const float x = ldg_stream_f32(base + p), y = ldg_stream_f32(base + HW + p), z = ldg_stream_f32(base + 2 * (size_t)HW + p);
    float u, v;
    if (xmap) { u = xmap[(size_t)f * NP + i]; v = ymap[(size_t)f * NP + i]; }
    else {
        const int row = (int)(p / Ww), col = (int)(p - (long long)row * Ww);
        u = (float)(col + (roi ? roi[2 * f + 1] : 0));
        v = (float)(row + (roi ? roi[2 * f] : 0));
    }
    float* o = pts + (size_t)f * 5 * cap;
    o[i] = fmaf(x, extent[3 * f], lfborder[3 * f]);
    o[cap + i] = fmaf(y, extent[3 * f + 1], lfborder[3 * f + 1]);
    o[2 * cap + i] = fmaf(z, extent[3 * f + 2], lfborder[3 * f + 2]);
    o[3 * cap + i] = u;
    o[4 * cap + i] = v;
}

struct CPlan { int S, quads_per_split, segcap; bool vec, sparse; size_t off_tot, total, smem_sparse; };

// workspace of the streaming mask-first kernel: ticket (256 B) + one state word per tile
inline int stream_tiles_per_frame(long long HW) { return (int)((HW + SM_TILE_PX - 1) / SM_TILE_PX); }
inline size_t stream_ws_bytes(int F, long long HW) { return 256 + (size_t)F * stream_tiles_per_frame(HW) * sizeof(uint32_t); }

CPlan make_cplan(int F, int Hh, int Ww, int cap, bool has_mask, int sms = 148) {
    CPlan p;
    const long long HW = (long long)Hh * Ww;
    p.vec = (HW % 4) == 0;
    p.smem_sparse = 0;
    p.sparse = has_mask && (HW % 32) == 0;
    if (p.sparse) { p.S = 1; p.quads_per_split = 0; p.segcap = cap; p.off_tot = 0; p.total = 0; return p; }
    int S = 1;
    if (p.vec && F < 4 * sms) {
        S = (4 * sms + F - 1) / F;                  // ~4 CTAs per SM
        int maxS = (int)((HW / 4 + 4095) / 4096);   // keep >= 16k pixels per band
        if (S > maxS) S = maxS;
        if (S < 1) S = 1;
    }
    const int nq = (int)(HW / 4);
    p.quads_per_split = ((nq + S - 1) / S + CT - 1) / CT * CT;
    p.S = (nq + p.quads_per_split - 1) / p.quads_per_split;
    if (p.S < 1) p.S = 1;
    long long sc = (long long)p.quads_per_split * 4;
    p.segcap = (int)(sc < cap ? sc : cap);
    if (p.S == 1) { p.off_tot = 0; p.total = 0; }
    else {
        size_t seg = (size_t)F * p.S * 5 * p.segcap * sizeof(float);
        p.off_tot = (seg + 255) / 256 * 256;
        p.total = p.off_tot + (size_t)F * p.S * sizeof(int32_t);
    }
    return p;
}

}  // namespace

extern "C" size_t pb_compact_workspace_bytes(int F, int Hh, int Ww, int cap) {
    if (F <= 0 || Hh <= 0 || Ww <= 0 || cap <= 0) return 0;
    const size_t dense = make_cplan(F, Hh, Ww, cap, false, pb_sm_count()).total;      // maskless stream (row bands + merge)
    const size_t stream = stream_ws_bytes(F, (long long)Hh * Ww);      // mask-first stream (ticket + tile states)
    return dense > stream ? dense : stream;
}

extern "C" int pb_compact_map(const float* xyz, const uint8_t* mask, const int32_t* roi, const float* extent,
                              const float* lfborder, int F, int Hh, int Ww, int cap, float* pts, int32_t* count,
                              int32_t* total, void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(F >= 0 && Hh > 0 && Ww > 0 && cap > 0, "pb_compact_map: bad sizes F=%d H=%d W=%d cap=%d", F, Hh, Ww, cap);
    PB_REQUIRE((long long)Hh * Ww < (1ll << 30), "pb_compact_map: map too large");
    if (F == 0) return 0;
    PB_REQUIRE(xyz && extent && lfborder && pts && count, "pb_compact_map: null pointer argument");
    CPlan p = make_cplan(F, Hh, Ww, cap, mask != nullptr && reinterpret_cast<uintptr_t>(mask) % 16 == 0, pb_sm_count());
    PB_REQUIRE(p.total == 0 || (ws && ws_bytes >= p.total), "pb_compact_map: workspace too small (%zu < %zu)", ws_bytes, p.total);
    PB_REQUIRE((long long)F * p.S < (1ll << 31), "pb_compact_map: grid too large");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CompactArgs a;
    a.xyz = xyz; a.mask = mask; a.roi = roi; a.extent = extent; a.lfborder = lfborder;
    a.count = count; a.total = total;
    a.HW = Hh * Ww; a.Ww = Ww; a.cap = cap; a.S = p.S; a.quads_per_split = p.quads_per_split; a.segcap = p.segcap;
    const long long HWl = (long long)Hh * Ww;
    if (mask && HWl % 16 == 0 && reinterpret_cast<uintptr_t>(mask) % 16 == 0 && ws && ws_bytes >= stream_ws_bytes(F, HWl) &&
        (long long)F * stream_tiles_per_frame(HWl) < (1ll << 30)) {
        a.pts = pts; a.S = 1; a.seg_total = nullptr;
        const int tpf = stream_tiles_per_frame(HWl), ntiles = F * tpf;
        if (cudaMemsetAsync(ws, 0, stream_ws_bytes(F, HWl), st) != cudaSuccess) return pb_check_launch("compact_stream memset") ? -2 : -2;
        int per_sm = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, compact_stream_kernel, CT, 0);
        const int grid = min(ntiles, pb_sm_count() * (per_sm > 0 ? per_sm : 1));
        compact_stream_kernel<<<grid, CT, 0, st>>>(a, reinterpret_cast<uint32_t*>(static_cast<char*>(ws) + 256),
                                                     static_cast<unsigned int*>(ws), tpf, ntiles);
        return pb_check_launch("compact_stream_kernel") ? -2 : 0;
    }
    if (p.sparse) {
        a.pts = pts; a.S = 1; a.seg_total = nullptr;
        compact_sparse_kernel<<<F, CT, 0, st>>>(a);
        return pb_check_launch("compact_sparse_kernel") ? -2 : 0;
    }
    const bool aligned = (reinterpret_cast<uintptr_t>(xyz) % 16 == 0) &&
                         (!mask || reinterpret_cast<uintptr_t>(mask) % 4 == 0);
    if (!p.vec || !aligned) {
        a.pts = pts; a.S = 1; a.seg_total = nullptr;
        if (mask) compact_kernel_scalar<true><<<F, CT, 0, st>>>(a);
        else compact_kernel_scalar<false><<<F, CT, 0, st>>>(a);
        return pb_check_launch("compact_kernel_scalar") ? -2 : 0;
    }
    if (p.S == 1) {
        a.pts = pts; a.seg_total = nullptr;
    } else {
        a.pts = static_cast<float*>(ws);
        a.seg_total = reinterpret_cast<int32_t*>(static_cast<char*>(ws) + p.off_tot);
    }
    if (mask) compact_kernel<true><<<F * p.S, CT, 0, st>>>(a);
    else compact_kernel<false><<<F * p.S, CT, 0, st>>>(a);
    if (pb_check_launch("compact_kernel")) return -2;
    if (p.S > 1) {
        merge_segments_kernel<<<F, CT, 0, st>>>(a.pts, a.seg_total, p.S, p.segcap, cap, pts, count, total);
        if (pb_check_launch("merge_segments_kernel")) return -2;
    }
    return 0;
}

extern "C" int pb_gather_choose(const float* xyz, const void* choose, int choose_is_i32, const float* xmap,
                                const float* ymap, const int32_t* n_valid, const int32_t* roi, const float* extent,
                                const float* lfborder, int F, int Hh, int Ww, int NP, int cap, float* pts,
                                int32_t* count, void* stream) {
    PB_REQUIRE(F >= 0 && Hh > 0 && Ww > 0 && NP > 0 && cap > 0, "pb_gather_choose: bad sizes");
    PB_REQUIRE(F <= 65535, "pb_gather_choose: F too large (<= 65535)");
    PB_REQUIRE((long long)Hh * Ww < (1ll << 30), "pb_gather_choose: map too large");
    if (F == 0) return 0;
    PB_REQUIRE(xyz && choose && extent && lfborder && pts && count, "pb_gather_choose: null pointer argument");
    PB_REQUIRE((xmap == nullptr) == (ymap == nullptr), "pb_gather_choose: xmap and ymap go together");
    const int n = NP < cap ? NP : cap;
    dim3 grid((n + 127) / 128, F);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (choose_is_i32)
        gather_kernel<int32_t><<<grid, 128, 0, st>>>(xyz, static_cast<const int32_t*>(choose), xmap, ymap, n_valid, roi,
                                                     extent, lfborder, Hh * Ww, Ww, NP, cap, pts, count);
    else
        gather_kernel<int64_t><<<grid, 128, 0, st>>>(xyz, static_cast<const int64_t*>(choose), xmap, ymap, n_valid, roi,
                                                     extent, lfborder, Hh * Ww, Ww, NP, cap, pts, count);
    return pb_check_launch("gather_kernel") ? -2 : 0;
}

extern "C" int pb_gather_correspondences(const float* xyz, const int64_t* choose, const float* xmap,
                                         const float* ymap, const float* extent, const float* lfborder, int F,
                                         int Hh, int Ww, int NP, int cap, float* pts, int32_t* count, void* stream) {
    PB_REQUIRE(xmap && ymap, "pb_gather_correspondences: null pointer argument");
    return pb_gather_choose(xyz, choose, 0, xmap, ymap, nullptr, nullptr, extent, lfborder, F, Hh, Ww, NP, cap, pts, count, stream);
}

// ---- pb_compact_map_noalloc: the mask-first streaming K0 with its scratch in a file-scope device array ------------
// No counterpart in the reference; added for callers that cannot allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same kernel as pb_compact_map's
// mask-first path (compact_stream_kernel); its scratch -- one ticket word and one state word per 64 KB mask tile --
// lives in g_noalloc_scratch instead of a caller workspace.  The array is shared by every call on a device and keeps
// the last call's contents: each call clears the part it uses on its stream first, and two calls that overlap on the
// GPU (different streams) RACE -- serialise them.
constexpr int PB_NOALLOC_MAX_TILES = 65536;        // 256 KB of tile states (+ 256 B ticket) per GPU that loads the library
static __device__ __align__(256) uint32_t g_noalloc_scratch[64 + PB_NOALLOC_MAX_TILES];

namespace {
// device address of g_noalloc_scratch per device: in host code the array's name is only a host-side stand-in
void* g_noalloc_ptr[64];
int noalloc_scratch_ptr(uint32_t** out) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { pb_set_error("pb_compact_map_noalloc: cudaGetDevice failed"); return 1; }
    if (!g_noalloc_ptr[dev]) {                    // once per device, with that device current
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) { pb_set_error("pb_compact_map_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return 1; }
        g_noalloc_ptr[dev] = p;
    }
    *out = static_cast<uint32_t*>(g_noalloc_ptr[dev]);
    return 0;
}
}  // namespace

extern "C" int pb_compact_map_noalloc_max_tiles(void) { return PB_NOALLOC_MAX_TILES; }

extern "C" int pb_compact_map_noalloc(const float* xyz, const uint8_t* mask, const int32_t* roi, const float* extent,
                                      const float* lfborder, int F, int Hh, int Ww, int cap, float* pts, int32_t* count,
                                      int32_t* total, void* stream) {
    PB_REQUIRE(F >= 0 && Hh > 0 && Ww > 0 && cap > 0, "pb_compact_map_noalloc: bad sizes F=%d H=%d W=%d cap=%d", F, Hh, Ww, cap);
    const long long HW = (long long)Hh * Ww;
    PB_REQUIRE(HW < (1ll << 30), "pb_compact_map_noalloc: map too large");
    if (F == 0) return 0;
    PB_REQUIRE(xyz && mask && extent && lfborder && pts && count, "pb_compact_map_noalloc: null pointer argument (a mask is required)");
    PB_REQUIRE(HW % 16 == 0 && reinterpret_cast<uintptr_t>(mask) % 16 == 0,
               "pb_compact_map_noalloc: H*W must be a multiple of 16 and the mask 16-byte aligned");
    const int tpf = stream_tiles_per_frame(HW);
    const long long ntiles = (long long)F * tpf;
    PB_REQUIRE(ntiles <= PB_NOALLOC_MAX_TILES,
               "pb_compact_map_noalloc: %lld mask tiles (F=%d x %d tiles of 65536 px) exceed the static scratch of %d tiles; "
               "split the batch or use pb_compact_map with a workspace", ntiles, F, tpf, PB_NOALLOC_MAX_TILES);
    uint32_t* scratch = nullptr;
    if (noalloc_scratch_ptr(&scratch)) return -2;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (cudaMemsetAsync(scratch, 0, (64 + (size_t)ntiles) * sizeof(uint32_t), st) != cudaSuccess) {   // ticket + this call's states
        pb_check_launch("pb_compact_map_noalloc memset");
        return -2;
    }
    CompactArgs a;
    a.xyz = xyz; a.mask = mask; a.roi = roi; a.extent = extent; a.lfborder = lfborder;
    a.pts = pts; a.count = count; a.total = total; a.seg_total = nullptr;
    a.HW = Hh * Ww; a.Ww = Ww; a.cap = cap; a.S = 1; a.quads_per_split = 0; a.segcap = cap;
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, compact_stream_kernel, CT, 0);
    const int grid = (int)(ntiles < (long long)pb_sm_count() * (per_sm > 0 ? per_sm : 1) ? ntiles : (long long)pb_sm_count() * (per_sm > 0 ? per_sm : 1));
    compact_stream_kernel<<<grid, CT, 0, st>>>(a, scratch + 64, scratch, tpf, (int)ntiles);
    return pb_check_launch("compact_stream_kernel (noalloc)") ? -2 : 0;
}
