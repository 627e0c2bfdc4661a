// synthgen.cu -- device-side synthetic frame generator for the sharded sweep, sm_100a.
//
// Measurement infrastructure, not a reference call site: config C5 evaluates 65,536 DISTINCT LineMOD-shaped frames,
// 262 GB of dense maps that cannot be kept (or built on the host) -- they are generated on the device, chunk by
// chunk, into a slab that stays zero outside the objects' silhouettes: a chunk's foreground pixels are written
// before it is evaluated and erased (from the list of written pixels) afterwards, so generating a 2048-frame chunk
// costs ~2 M pixel writes instead of 8 GB.
//
// A frame shows the ellipsoid inscribed in its class box under a pose given by the host (pose_estimation_b200/
// synth.py::c5_params: 48 floats per frame); per pixel of the silhouette's bounding box the ray / ellipsoid hit
// gives the exact model coordinate; a pixel is kept with probability P / area (the silhouette ellipse's analytic
// area, from the host; Philox4x32-10 keyed by the seed,
// counter = (pixel, frame)), a kept pixel is an outlier (uniform in the box) with probability `outlier` and
// otherwise gets ~N(0, sigma) noise (sum of four uniforms per coordinate).  Every operation is IEEE float32
// add / mul / div / sqrt in a written order (-fmad=false), so synth.py's NumPy twin reproduces every byte.
#include "common.cuh"

namespace {

constexpr int PARAM_STRIDE = 48;
constexpr int LIST_CAP = 4096;      // written pixels remembered per frame for erase (P <= 2048 is checked on the host)
// params[f]: 0-8 R (object -> camera, row major) | 9-11 t | 12-15 fx fy cx cy | 16-18 o = -R^T t | 19-21 oo = (o - centre) / semi
// | 22-24 semi | 25-27 lfb | 28-30 ext | 31 qc = oo.oo - 1 | 32-35 bbox r0 r1 c0 c1 | 36 P | 37 outlier | 38 sigma * sqrt(3)
// | 39 keep probability = P / (analytic area of the silhouette ellipse)
// | 40-41 1/fx 1/fy | 42-44 1/semi   (reciprocals rounded once on the host: the per-pixel test has no division)
constexpr uint32_t GEN_TAG = 0x47454e35u;   // "GEN5"

struct GenArgs {
    const float* params;
    int F, Hh, Ww, bbox_max;
    uint32_t seed_lo, seed_hi, frame0;
    float* xyz; uint8_t* mask; int32_t* area; int32_t* nlist; int32_t* list;
};

__device__ __forceinline__ float u01(uint32_t r) { return (float)(r >> 8) * 5.9604644775390625e-08f; }   // 2^-24: exact

// ray / ellipsoid test of pixel (row, col).  The pixel is on the silhouette iff the ray's quadratic has real roots
// in front of the camera: disc >= 0 and qb < 0 (the camera is outside the ellipsoid, qc > 0) -- no sqrt, no division.
struct Ray { float d[3], qa, qb, disc; };
__device__ __forceinline__ bool hit_pixel(const float* __restrict__ p, int row, int col, Ray& r) {
    const float dcx = ((float)col - p[14]) * p[40];
    const float dcy = ((float)row - p[15]) * p[41];
#pragma unroll
    for (int k = 0; k < 3; ++k) r.d[k] = (p[k] * dcx + p[3 + k] * dcy) + p[6 + k];      // R^T (dcx, dcy, 1)
    const float ddx = r.d[0] * p[42], ddy = r.d[1] * p[43], ddz = r.d[2] * p[44];
    r.qa = (ddx * ddx + ddy * ddy) + ddz * ddz;
    r.qb = (ddx * p[19] + ddy * p[20]) + ddz * p[21];
    r.disc = r.qb * r.qb - r.qa * p[31];
    return r.disc >= 0.0f && r.qb < 0.0f;
}
// exact model point of a silhouette pixel: o + tau d, tau the near root
__device__ __forceinline__ void hit_point(const float* __restrict__ p, const Ray& r, float Xo[3]) {
    const float tau = (-r.qb - sqrtf(r.disc)) / r.qa;
#pragma unroll
    for (int k = 0; k < 3; ++k) Xo[k] = p[16 + k] + tau * r.d[k];
}

constexpr int GEN_PX = 2048;         // bounding-box pixels per CTA (8 per thread): few, fat CTAs -- the launch of
                                     // one 256-pixel CTA per tile of the LARGEST box (~750 k CTAs per pass) was the cost

// one kept pixel: model coordinate with noise / outlier, written into the slab and remembered for erase
__device__ __forceinline__ void write_pixel(const GenArgs& a, const float* __restrict__ p, int f, int pix, const float Xo[3],
                                            const uint32_t r[4]) {
    const uint32_t frame = a.frame0 + (uint32_t)f;
    const size_t plane = (size_t)a.Hh * a.Ww;
    float* px = a.xyz + (size_t)f * 3 * plane + pix;
    const int slot = atomicAdd(a.nlist + f, 1);              // remembered for erase (order does not matter)
    if (slot < LIST_CAP) a.list[(size_t)f * LIST_CAP + slot] = pix;
    float X[3];
    if (u01(r[1]) < p[37]) {                     // outlier: uniform in the box
        uint32_t q[4];
        philox4x32_10((uint32_t)pix, frame, 1u, GEN_TAG, a.seed_lo, a.seed_hi, q);
#pragma unroll
        for (int c = 0; c < 3; ++c) X[c] = p[25 + c] + u01(q[c]) * p[28 + c];
    } else {                                     // inlier: + sigma * sqrt(3) * (u1 + u2 + u3 + u4 - 2)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            uint32_t q[4];
            philox4x32_10((uint32_t)pix, frame, 2u + (uint32_t)c, GEN_TAG, a.seed_lo, a.seed_hi, q);
            const float s4 = ((u01(q[0]) + u01(q[1])) + (u01(q[2]) + u01(q[3]))) - 2.0f;
            X[c] = Xo[c] + p[38] * s4;
        }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float v = (X[c] - p[25 + c]) / p[28 + c];
        v = fminf(fmaxf(v, 1.0e-3f), 1.0f);
        px[c * plane] = v;
    }
    a.mask[(size_t)f * plane + pix] = 1;
}

// Phase A: every pixel of the CTA's part of the bounding box -- hit test, keep draw; kept pixels go to a list in
// shared memory.  Phase B: the list is worked off densely (the ~5 % kept pixels would otherwise drag whole warps
// through the four Philox calls and the divisions of the write path: measured 0.44 -> see profiles/r02_summary.md).
__global__ void __launch_bounds__(256) synth_kernel(GenArgs a) {
    __shared__ float sp[PARAM_STRIDE];
    __shared__ int s_list[GEN_PX];
    __shared__ int s_n;
    const int f = blockIdx.y;
    const float* gp = a.params + (size_t)f * PARAM_STRIDE;
    const int r0 = (int)gp[32], r1 = (int)gp[33], c0 = (int)gp[34], c1 = (int)gp[35];
    const int bw = c1 - c0, npx = bw > 0 ? bw * (r1 - r0) : 0;
    const int k0 = blockIdx.x * GEN_PX;
    if (k0 >= npx) return;                       // this frame's box is smaller than the largest one (uniform exit)
    if (threadIdx.x < PARAM_STRIDE) sp[threadIdx.x] = gp[threadIdx.x];
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
    const float* p = sp;
    const float pk = p[39];                      // P / (analytic silhouette area), from the host
    const uint32_t frame = a.frame0 + (uint32_t)f;
#pragma unroll 2
    for (int j = 0; j < GEN_PX / 256; ++j) {
        const int k = k0 + j * 256 + threadIdx.x;
        if (k >= npx) break;
        const int rr = k / bw;
        const int row = r0 + rr, col = c0 + (k - rr * bw);
        Ray ray;
        if (!hit_pixel(p, row, col, ray)) continue;
        uint32_t r[4];
        philox4x32_10((uint32_t)(row * a.Ww + col), frame, 0u, GEN_TAG, a.seed_lo, a.seed_hi, r);
        if (u01(r[0]) < pk) s_list[atomicAdd(&s_n, 1)] = k;
    }
    __syncthreads();
    const int n = s_n;
    for (int i = threadIdx.x; i < n; i += 256) {
        const int k = s_list[i];
        const int rr = k / bw;
        const int row = r0 + rr, col = c0 + (k - rr * bw);
        const int pix = row * a.Ww + col;
        Ray ray;
        float Xo[3];
        hit_pixel(p, row, col, ray);             // the same arithmetic as in phase A
        hit_point(p, ray, Xo);
        uint32_t r[4];
        philox4x32_10((uint32_t)pix, frame, 0u, GEN_TAG, a.seed_lo, a.seed_hi, r);
        write_pixel(a, p, f, pix, Xo, r);
    }
}

__global__ void __launch_bounds__(256) synth_erase_kernel(GenArgs a) {
    const int f = blockIdx.y, i = blockIdx.x * 256 + threadIdx.x;
    if (i >= min(a.nlist[f], LIST_CAP)) return;
    const int pix = a.list[(size_t)f * LIST_CAP + i];
    const size_t plane = (size_t)a.Hh * a.Ww;
    float* px = a.xyz + (size_t)f * 3 * plane + pix;
    px[0] = 0.0f; px[plane] = 0.0f; px[2 * plane] = 0.0f;
    a.mask[(size_t)f * plane + pix] = 0;
}

}  // namespace

extern "C" size_t pb_synth_workspace_bytes(int F) {
    return F <= 0 ? 0 : (size_t)F * (2 + LIST_CAP) * sizeof(int32_t);
}

extern "C" int pb_synth_frames(const float* params, int F, int Hh, int Ww, int bbox_max, uint32_t seed_lo,
                               uint32_t seed_hi, uint32_t frame0, float* xyz, uint8_t* mask, void* ws, size_t ws_bytes,
                               int erase, void* stream) {
    PB_REQUIRE(F >= 0 && Hh > 0 && Ww > 0 && bbox_max > 0, "pb_synth_frames: bad sizes");
    PB_REQUIRE(F <= 65535, "pb_synth_frames: F too large (<= 65535)");
    if (F == 0) return 0;
    PB_REQUIRE(params && xyz && mask && ws && ws_bytes >= pb_synth_workspace_bytes(F), "pb_synth_frames: null pointer / workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int32_t* area = static_cast<int32_t*>(ws);
    GenArgs a{params, F, Hh, Ww, bbox_max, seed_lo, seed_hi, frame0, xyz, mask, area, area + F, area + 2 * F};
    if (erase) {
        synth_erase_kernel<<<dim3(LIST_CAP / 256, F), 256, 0, st>>>(a);
        return pb_check_launch("synth_erase_kernel") ? -2 : 0;
    }
    dim3 grid((bbox_max + GEN_PX - 1) / GEN_PX, F);
    if (cudaMemsetAsync(area, 0, (size_t)F * 2 * sizeof(int32_t), st) != cudaSuccess) { pb_check_launch("pb_synth_frames memset"); return -2; }
    synth_kernel<<<grid, 256, 0, st>>>(a);
    return pb_check_launch("synth_kernel") ? -2 : 0;
}
