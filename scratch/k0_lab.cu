// Developer probe: the mask-first K0 kernels in isolation on a C2-shaped batch (1024 frames of 480x640, ~1000
// foreground pixels per frame drawn from a disc of ~75 px radius).  Builds compact.cu with -DPB_SM_G / -DPB_SM_OCC.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false -DPB_SM_G=4 -DPB_SM_OCC=8 scratch/k0_lab.cu -o scratch/k0_lab
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../pose_estimation_b200/csrc/compact.cu"

void pb_set_error(const char* fmt, ...) { va_list ap; va_start(ap, fmt); vfprintf(stderr, fmt, ap); va_end(ap); fprintf(stderr, "\n"); }
int pb_check_launch(const char* what) { cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", what, cudaGetErrorString(e)); return 1; } return 0; }
int pb_sm_count() { int n = 0; cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, 0); return n; }

int main(int argc, char** argv) {
    const int F = argc > 1 ? atoi(argv[1]) : 1024, H = 480, W = 640, cap = 1280;
    const int dense = argc > 2 ? atoi(argv[2]) : 0;          // 1: every pixel of the disc (radius 18 -> ~1000 px)
    const size_t HW = (size_t)H * W;
    std::vector<uint8_t> mask(F * HW, 0);
    unsigned long long st = 88172645463325252ull;
    auto rnd = [&]() { st ^= st << 13; st ^= st >> 7; st ^= st << 17; return (unsigned)(st >> 32); };
    long long npx = 0;
    for (int f = 0; f < F; ++f) {
        const int rad = dense ? 18 : 40 + rnd() % 60;
        const int cr = rad + rnd() % (H - 2 * rad), cc = rad + rnd() % (W - 2 * rad);
        int want = 800 + rnd() % 401, got = 0;
        if (dense) {
            for (int r = -rad; r <= rad; ++r) for (int c = -rad; c <= rad; ++c)
                if (r * r + c * c <= rad * rad) { mask[f * HW + (size_t)(cr + r) * W + cc + c] = 1; ++got; }
        } else {
            while (got < want) {
                int r = (int)(rnd() % (2 * rad + 1)) - rad, c = (int)(rnd() % (2 * rad + 1)) - rad;
                if (r * r + c * c > rad * rad) continue;
                uint8_t& m = mask[f * HW + (size_t)(cr + r) * W + cc + c];
                if (!m) { m = 1; ++got; }
            }
        }
        npx += got;
    }
    float* xyz; uint8_t* dmask; float *ext, *lfb, *pts; int32_t* count; void* ws;
    cudaMalloc(&xyz, F * 3 * HW * 4); cudaMemset(xyz, 0, F * 3 * HW * 4);
    cudaMalloc(&dmask, F * HW); cudaMemcpy(dmask, mask.data(), F * HW, cudaMemcpyHostToDevice);
    cudaMalloc(&ext, F * 12); cudaMalloc(&lfb, F * 12); cudaMemset(ext, 0, F * 12); cudaMemset(lfb, 0, F * 12);
    cudaMalloc(&pts, (size_t)F * 5 * cap * 4); cudaMalloc(&count, F * 4);
    const size_t wsb = pb_compact_workspace_bytes(F, H, W, cap);
    cudaMalloc(&ws, wsb);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 2; ++mode) {                  // 0: streaming kernel (workspace given); 1: round-1 kernel (no workspace)
        for (int i = 0; i < 3; ++i) pb_compact_map(xyz, dmask, nullptr, ext, lfb, F, H, W, cap, pts, count, nullptr, mode ? nullptr : ws, mode ? 0 : wsb, 0);
        cudaDeviceSynchronize();
        const int n = 20;
        cudaEventRecord(e0);
        for (int i = 0; i < n; ++i) pb_compact_map(xyz, dmask, nullptr, ext, lfb, F, H, W, cap, pts, count, nullptr, mode ? nullptr : ws, mode ? 0 : wsb, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= n;
        std::vector<int32_t> hc(F); cudaMemcpy(hc.data(), count, F * 4, cudaMemcpyDeviceToHost);
        long long tot = 0; for (int f = 0; f < F; ++f) tot += hc[f];
        const double strict = (double)F * HW + (double)npx * 32.0;
        printf("%s G=%d W=%d OCC=%d PF=%d F=%d dense=%d: %.4f ms  (%.0f GB/s strict bytes; count sum %lld of %lld) %s\n",
               mode ? "compact_sparse_kernel (r1)" : "compact_stream_kernel", PB_SM_G, PB_SM_W, PB_SM_OCC, PB_SM_PF, F, dense, ms, strict / ms / 1e6, tot, npx,
               cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
