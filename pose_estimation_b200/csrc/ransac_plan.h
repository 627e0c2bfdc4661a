// ransac_plan.h -- launch plan + argument block shared by ransac_score.cu and ransac.cu.
#pragma once
#include "common.cuh"

struct ScoreArgs {
    const float* pts;
    const int32_t* count;
    const float* K4;
    const int32_t* sample_idx;
    float* hypRt;                    // [F,H,12] R|t of every hypothesis (t_z = NaN: no model)
    int32_t* hyp_counts;             // optional [F,H]
    unsigned long long* slice_best;  // [F,HS] packed (count+1, ~hyp)
    int32_t* cnt_all;                // balanced mode: [F,H] inlier counts accumulated with atomics (-1 = no model)
    int32_t* prefix;                 // balanced mode: [F+1] exclusive prefix of the frames' point counts rounded up to 8
                                     // (written by block 0 of the hypothesis kernel, read by every scoring CTA)
    int32_t* tprefix;                // tensor-core mode: [F+1] exclusive prefix of the frames' PB_TC_TILE-point tile counts (same writer)
    int32_t* tc_stats;               // tensor-core mode: [0] barrier timeout, [1] exact re-evaluations (groups of 4 tests x 32 lanes),
                                     // [2] calibration: largest observed |s_tc - s_fp32| / band as float bits, [3] verdicts that differ outside the band
    float tc_kappa;                  // tensor-core mode: multiplier of the re-evaluation band (1 = the derived bound)
    int tc_mode;                     // tensor-core modes: 0 product, 1 calibration pass, 2 clock breakdown (tcgen05 only)
    float* fstats;                   // mma.sync mode: [F][8] frame constants (written by frame_stats_kernel before the scoring kernel)
    int cap, H, HS, Hs, CH, PP;      // Hs hypotheses per slice (power of two >= 32); PP point parts
    uint32_t frame0, seed_lo, seed_hi;
    float inv_thr;
    int F;
};

struct Plan {
    int HS, Hs, CH, PP, NHL;
    size_t smem_score, smem_final, off_best, total;
    int bal, bal_grid, bal_nhl;      // balanced scoring (see ransac_score_bal_kernel): CTAs of equal point ranges
    size_t smem_bal, off_cnt, off_prefix;
    int tc;                          // tensor-core scoring (ransac_score_tc.cu) is possible for this shape
    size_t off_tprefix, off_tcstats, off_fstats;
};

// Hypotheses of a frame are cut into HS slices of Hs (power of two) so that few-frame launches
// still fill 148 SMs; a slice narrower than the CTA splits the points PP ways instead.
#define PB_SCORE_THREADS 128     /* CTA size of the scoring kernel (ST in ransac_score.cu) */
#define PB_TC_TILE 64            /* points per work unit of the tensor-core scoring kernel */
#define PB_TC_STATS_BYTES 128     /* int32 stats[4] + twelve 64-bit clock sums of the profiling mode */
#define PB_BAL_SUB 256           /* points per staging buffer of the balanced scoring kernel (two buffers) */
#ifndef PB_BAL_DYN_PCT
#define PB_BAL_DYN_PCT 0         /* per cent of the points the balanced scoring kernel deals out through a counter */
#endif
#ifndef PB_BAL_WAVES
#define PB_BAL_WAVES 1           /* CTAs of the balanced scoring kernel = SMs x resident CTAs x this */
#endif

static inline Plan make_plan(int F, int cap, const pb_ransac_params* prm, int sms = 148) {
    if (sms <= 0) sms = 148;
    const int two_waves = 2 * sms;                     // fewer frames than this: slice the hypotheses to fill the GPU
    Plan p;
    const int H = prm->n_hyp;
    int Hs = 32;
    while (Hs < H && Hs < 2 * PB_SCORE_THREADS) Hs <<= 1;   // up to 2 hypotheses per thread
    if (F < two_waves) {                               // few frames: more, narrower slices
        int want = (two_waves + F - 1) / F;
        while (Hs > 64 && (H + Hs - 1) / Hs < want) Hs >>= 1;
    }
    p.Hs = Hs;
    p.HS = (H + Hs - 1) / Hs;
    p.NHL = Hs > PB_SCORE_THREADS ? Hs / PB_SCORE_THREADS : 1;
    p.PP = Hs < PB_SCORE_THREADS ? PB_SCORE_THREADS / Hs : 1;
    int CH = (cap + 127) / 128 * 128;
    if (CH > 2048) CH = 2048;
    p.CH = CH;
    p.smem_score = (size_t)CH * 5 * 4 + (size_t)Hs * 4;
    p.smem_final = (size_t)((cap + PB_THREADS - 1) / PB_THREADS) * (PB_THREADS / 32) * 4;
    size_t hyp = (size_t)F * H * 12 * sizeof(float);
    p.off_best = (hyp + 255) / 256 * 256;
    p.total = p.off_best + (size_t)F * p.HS * sizeof(unsigned long long);
    // A batch of about one wave of CTAs cannot be rebalanced by the hardware scheduler: frames carry different
    // numbers of points and the launch lasts as long as the SM with the largest sum.  Then CTAs take equal
    // ranges of the concatenated point list instead (H <= 256 hypotheses held by one CTA).
    p.smem_bal = 0;                                    // the balanced kernel's two 256-point buffers are static (10 KB)
    const int per_sm = 6;                              // <= 85 registers x 128 threads: six CTAs per SM, one wave (measured with
                                                       // 5 / 4 / 3: 0.1526 / 0.1542 / 0.1581 ms against 0.1518, and the three-stream
                                                       // sweep gains nothing from the freed registers: 0.302 / 0.304 / 0.314 ms per step)
    // also for fewer frames when they can hold many points (dense masks: 2 k .. 50 k pixels per frame, very uneven)
    const bool many = F >= two_waves || (long long)F * cap >= 4ll * sms * per_sm * PB_BAL_SUB;
    p.bal = (H <= 2 * PB_SCORE_THREADS && many && cap % 4 == 0) ? 1 : 0;
    p.bal_nhl = H > PB_SCORE_THREADS ? 2 : 1;          // hypotheses per thread of the balanced kernel (the slice plan above is its fallback's)
    p.bal_grid = sms * per_sm * PB_BAL_WAVES;
    p.off_cnt = (p.total + 255) / 256 * 256;
    p.off_prefix = p.off_cnt;
    if (p.bal) {
        p.off_prefix = (p.off_cnt + (size_t)F * H * sizeof(int32_t) + 255) / 256 * 256;
        p.total = p.off_prefix + (size_t)(F + 2) * sizeof(int32_t);    // prefix[0..F] + the unit counter
    }
    // tensor-core scoring: same accumulation target (cnt_all) as the balanced kernel, H <= 256 hypotheses per frame
    p.tc = (p.bal && H <= 256) ? 1 : 0;
    p.off_tprefix = p.off_tcstats = p.off_fstats = p.total;
    if (p.tc) {
        p.off_tprefix = (p.total + 255) / 256 * 256;
        p.off_tcstats = (p.off_tprefix + (size_t)(F + 1) * sizeof(int32_t) + 255) / 256 * 256;
        p.off_fstats = (p.off_tcstats + PB_TC_STATS_BYTES + 255) / 256 * 256;
        p.total = p.off_fstats + (size_t)F * 8 * sizeof(float);
    }
    return p;
}

// defined in ransac_score.cu
int pb_launch_score(const ScoreArgs& a, const Plan& p, int F, cudaStream_t st);
// defined in ransac_score_tc.cu: K3 on the tensor cores (a.tprefix, a.tc_stats set; a.cnt_all initialised by the hypothesis kernel)
int pb_launch_score_tc(const ScoreArgs& a, int mode /*0 product, 1 calibration, 2 clock breakdown*/, cudaStream_t st);
// process-wide choice of the scoring kernel (abi: pb_ransac_score_impl): 0 = automatic, 1 = FP32 pipe only, 2 = tensor cores
// whenever the shape allows, 3 = tensor cores with the calibration pass (slow; fills tc_stats[2..3]), 4 = tensor cores with the clock breakdown,
// 5 = the warp-level tensor path (mma.sync) whenever the shape allows, 6 = the same with the calibration pass
// defined in ransac_score_mma.cu: K3 with the projection on the warp-level tensor path (a.prefix, a.fstats, a.tc_stats set)
int pb_launch_score_mma(const ScoreArgs& a, int calib, cudaStream_t st);
int pb_score_impl();
float pb_score_tc_kappa();
#ifndef PB_SCORE_MMA_DEFAULT
#define PB_SCORE_MMA_DEFAULT 0   /* 1: the automatic choice (impl 0) takes the mma.sync kernel whenever the plan allows */
#endif
