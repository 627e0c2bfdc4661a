// abi.cu -- error reporting and the composed map -> pose entry point of libposeb200.so.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include "common.cuh"

namespace {
thread_local char g_err[512] = "";
}

void pb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int pb_check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        pb_set_error("%s: %s", what, cudaGetErrorString(e));
        return 1;
    }
    return 0;
}

int pb_sm_count() {
    static int cached[64];                         // 0 = not queried yet; benign race (same value written)
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

// ---- choice of the scoring kernel (process-wide, read at every pb_pnp_ransac_* call) ----
namespace {
std::atomic<int> g_score_impl{0};
std::atomic<float> g_score_kappa{1.0f};
}
int pb_score_impl() { return g_score_impl.load(std::memory_order_relaxed); }
float pb_score_tc_kappa() { return g_score_kappa.load(std::memory_order_relaxed); }
extern "C" int pb_ransac_score_impl(int impl, float band_scale) {
    PB_REQUIRE(impl >= 0 && impl <= 6, "pb_ransac_score_impl: impl=%d (0 automatic, 1 FP32 pipe, 2 tensor cores, 3 tensor cores + calibration)", impl);
    PB_REQUIRE(band_scale > 0.0f && band_scale <= 1024.0f, "pb_ransac_score_impl: band_scale out of range");
    g_score_impl.store(impl, std::memory_order_relaxed);
    g_score_kappa.store(band_scale, std::memory_order_relaxed);
    return 0;
}

// ---- optional per-kernel timing of the RANSAC call (CUDA events recorded on the caller's stream) ----
namespace {
constexpr int TIM_SLOTS = 256, TIM_MARKS = 4;     // marks: before hyp | before score | before finalize | after
// Per THREAD: a thread that has not called pb_stage_timing(1) never touches any of it, and two threads that both
// measure do not share slots.  Off by default.
thread_local bool g_tim_on = false;
thread_local int g_tim_calls = 0;
thread_local cudaEvent_t g_tim_ev[TIM_SLOTS][TIM_MARKS];
thread_local bool g_tim_made = false;
}

void pb_stage_mark(int mark, cudaStream_t st) {
    if (!g_tim_on || g_tim_calls >= TIM_SLOTS) return;
    cudaEventRecord(g_tim_ev[g_tim_calls][mark], st);
    if (mark == TIM_MARKS - 1) ++g_tim_calls;
}

extern "C" int pb_stage_timing(int enable) {
    if (enable && !g_tim_made) {
        for (int i = 0; i < TIM_SLOTS; ++i)
            for (int k = 0; k < TIM_MARKS; ++k)
                if (cudaEventCreate(&g_tim_ev[i][k]) != cudaSuccess) { pb_set_error("pb_stage_timing: cudaEventCreate failed"); return -2; }
        g_tim_made = true;
    }
    g_tim_on = enable != 0;
    g_tim_calls = 0;
    return 0;
}

extern "C" int pb_stage_timing_read(float* ms3, int* calls) {
    PB_REQUIRE(ms3 && calls, "pb_stage_timing_read: null pointer argument");
    double acc[TIM_MARKS - 1] = {0, 0, 0};
    const int n = g_tim_calls;
    for (int i = 0; i < n; ++i) {
        if (cudaEventSynchronize(g_tim_ev[i][TIM_MARKS - 1]) != cudaSuccess) { pb_set_error("pb_stage_timing_read: sync failed"); return -2; }
        for (int k = 0; k + 1 < TIM_MARKS; ++k) {
            float ms = 0.0f;
            cudaEventElapsedTime(&ms, g_tim_ev[i][k], g_tim_ev[i][k + 1]);
            acc[k] += ms;
        }
    }
    for (int k = 0; k + 1 < TIM_MARKS; ++k) ms3[k] = n ? (float)(acc[k] / n) : 0.0f;
    *calls = n;
    return 0;
}

extern "C" const char* pb_last_error(void) { return g_err; }
extern "C" int pb_abi_version(void) { return PB_ABI_VERSION; }

namespace {
inline size_t align256(size_t x) { return (x + 255) / 256 * 256; }
}

extern "C" size_t pb_pnp_ransac_map_workspace_bytes(int F, int Hh, int Ww, int cap, const pb_ransac_params* prm) {
    if (F <= 0 || Hh <= 0 || Ww <= 0 || cap <= 0 || !prm) return 0;
    size_t pts = align256((size_t)F * 5 * cap * sizeof(float));
    size_t cnt = align256((size_t)F * sizeof(int32_t));
    return pts + cnt + align256(pb_compact_workspace_bytes(F, Hh, Ww, cap)) + align256(pb_ransac_workspace_bytes(F, cap, prm));
}

extern "C" int pb_pnp_ransac_map(const float* xyz, const uint8_t* mask, const int32_t* roi, const float* extent,
                                 const float* lfborder, const float* K4, int F, int Hh, int Ww, int cap, uint32_t frame0,
                                 const pb_ransac_params* prm, float* R, float* t, int32_t* n_inl, int32_t* best_hyp,
                                 int32_t* n_inl_refined, int32_t* status, void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(prm != nullptr, "pb_pnp_ransac_map: prm is null");
    PB_REQUIRE(F >= 0 && Hh > 0 && Ww > 0 && cap > 0, "pb_pnp_ransac_map: bad sizes");
    if (F == 0) return 0;
    const size_t need = pb_pnp_ransac_map_workspace_bytes(F, Hh, Ww, cap, prm);
    PB_REQUIRE(ws != nullptr && ws_bytes >= need, "pb_pnp_ransac_map: workspace too small (%zu < %zu)", ws_bytes, need);
    char* base = static_cast<char*>(ws);
    float* pts = reinterpret_cast<float*>(base);
    size_t off = align256((size_t)F * 5 * cap * sizeof(float));
    int32_t* count = reinterpret_cast<int32_t*>(base + off);
    off += align256((size_t)F * sizeof(int32_t));
    void* cws = base + off;
    const size_t cbytes = align256(pb_compact_workspace_bytes(F, Hh, Ww, cap));
    off += cbytes;
    void* rws = base + off;
    const size_t rbytes = align256(pb_ransac_workspace_bytes(F, cap, prm));
    int rc = pb_compact_map(xyz, mask, roi, extent, lfborder, F, Hh, Ww, cap, pts, count, nullptr, cws, cbytes, stream);
    if (rc) return rc;
    return pb_pnp_ransac_pts(pts, count, K4, F, cap, frame0, prm, nullptr, R, t, n_inl, best_hyp, n_inl_refined, status,
                             nullptr, nullptr, nullptr, nullptr, rws, rbytes, stream);
}
