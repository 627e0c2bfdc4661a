/* oracle/pose_ref.h -- CPU oracle for the pose-recovery + evaluation hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under pose_estimation_b200/ may include, link or call
 * this.  It is a plain-C, single-threaded, bit-defined restatement of the algorithm that the
 * sm_100a kernels implement (same counter-based RNG, same minimal solver, same FP32 operation
 * order with explicit fmaf, same tie-breaks, same reduction trees), so tests can demand
 * bit-exact inlier counts / chosen hypotheses / KNN indices from the GPU.
 *
 * Behaviour it stands in for in the reference (/root/reference, read-only):
 *   tools/trainer.py:383-438   Trainer.get_pose  (xyz map -> correspondences -> PnP-RANSAC -> R,t)
 *   tools/trainer.py:370-381   Trainer.cal_dis   (model@R^T+t -> ADD(-S), angular, translation)
 *   lib/utils/metric.py:17-36  Metric.cal_adds_cuda
 *   lib/utils/metric.py:67-84  angular_distance / translation_distance
 *   lib/network/loss.py:25-41  PoseLoss.forward  (symmetric-KNN loss)
 *   train.py:126               knn = generic_argkmin('SqDist(x, y)', ...)   (pykeops, not vendored)
 *   lib/network/krrn.py:153    T-offset voting
 *   lib/utils/utlis.py:55-74   choose_ransac_batch
 * The arithmetic of PnP-RANSAC in the reference lives in OpenCV (cv2.solvePnPRansac, un-vendored
 * third-party, no version pin in the reference; 4.13.0 in this image).  OpenCV's internal RNG /
 * EPnP are not reproducible from outside, so parity against cv2 is pinned at the pose / metric
 * level by oracle/ref_callsites.py (real cv2) and tests/golden/, while THIS file pins the
 * bit-level behaviour of our own RANSAC.  Parity status: pinned against cv2-generated golden
 * vectors (tests/golden/), not against reference-held tests (the reference has none).
 */
#ifndef POSE_REF_H
#define POSE_REF_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int32_t n_hyp;        /* hypotheses per frame                       */
    float   thresh_px;    /* reprojection threshold (px)                */
    int32_t gn_iters;     /* Gauss-Newton iterations per refit          */
    int32_t lo_rounds;    /* extra re-score -> refit rounds             */
    int32_t min_inliers;  /* below this the frame is reported as failed */
    uint32_t seed_lo, seed_hi;
} ref_ransac_params;

/* Philox4x32-10 */
void ref_philox4x32(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
/* 4 distinct indices in [0,P); returns 0 if none could be drawn */
int ref_sample4(uint32_t seed_lo, uint32_t seed_hi, uint32_t frame, uint32_t hyp, int P, int32_t idx[4]);

/* P3P on three correspondences: x = object points, y = unit bearings.  Returns #solutions (<=4). */
int ref_p3p(const float x[3][3], const float y[3][3], float R[4][9], float t[4][3]);

/* One hypothesis from 4 correspondence indices. pts = SoA X,Y,Z,U,V each [P].
 * Writes M[12]=K[R|t], R[9], t[3]; returns 1 if valid. */
int ref_hypothesis(const float* X, const float* Y, const float* Z, const float* U, const float* V,
                   const float K4[4], const int32_t idx[4], float M[12], float R[9], float t[3]);

/* inlier count of one projection matrix (bitmap optional, ceil(P/32) words); inv_thr = 1/thresh_px */
int ref_score(const float M[12], const float* X, const float* Y, const float* Z, const float* U,
              const float* V, int P, float inv_thr, uint32_t* bitmap);

/* Full per-frame RANSAC.  sample_idx optional [H,4].  hyp_counts optional [H] (-1 = invalid).
 * status: 0 ok, 1 too few points, 2 no valid hypothesis / too few inliers. */
int ref_ransac_frame(const float* X, const float* Y, const float* Z, const float* U, const float* V,
                     int P, const float K4[4], const ref_ransac_params* prm, uint32_t frame,
                     const int32_t* sample_idx, float R[9], float t[3], int32_t* n_inl,
                     int32_t* best_hyp, int32_t* n_inl_refined, int32_t* hyp_counts,
                     uint32_t* bitmap, float R0[9], float t0[3]);

/* Mask compaction + de-normalisation of one frame (tools/trainer.py:415-421).
 * xyz: [3,H,W] planar f32 (normalised), mask: [H,W] u8 or NULL (-> any channel != 0).
 * roi: rmin,cmin offsets added to row/col to give v,u.  Returns count (clamped to cap). */
int ref_compact_map(const float* xyz, const uint8_t* mask, int Hh, int Ww, const float extent[3],
                    const float lfborder[3], int rmin, int cmin, int cap,
                    float* X, float* Y, float* Z, float* U, float* V, int32_t* total);

/* argmin_j SqDist(x_i, y_j), first index on ties; FP32 direct-difference formula. */
void ref_knn1(const float* x, const float* y, int N, int M, int64_t* idx, float* d2);

/* ADD = mean ||pred_i - tgt_i||; ADD-S = mean_i min_j ||pred_j - tgt_i|| (lib/utils/metric.py:23-31) */
void ref_add_adds(const float* pred, const float* tgt, int N, int M, float* add, float* adds);

/* pred = model @ R^T + t */
void ref_transform(const float* model, int N, const float R[9], const float t[3], float* out);

/* rotation matrix -> unit quaternion wxyz (trig-free) */
void ref_rot_to_quat(const float R[9], float q[4]);
/* angular distance in degrees (lib/utils/metric.py:67-80) */
float ref_angular_deg(const float R1[9], const float R2[9]);

/* T-offset vote (lib/network/krrn.py:153): cloud [NP,3], t_res [3,NP] -> t[3] */
void ref_vote_translation(const float* cloud, const float* t_res, int NP, float t[3]);

/* order-preserving subset / wrap-pad of an inlier bitmap to exactly n indices
 * (lib/utils/utlis.py:64-71 semantics, Philox instead of numpy's global RNG) */
void ref_resample_inliers(const uint32_t* bitmap, int P, int n, uint32_t seed_lo, uint32_t seed_hi,
                          uint32_t frame, int32_t* out);

/* ---- 3D-3D RANSAC + Umeyama (sim3_ref.c; lib/transform/umeyama.py:8-97) ---- */
int ref_umeyama_core(const double ms[3], const double mt[3], const double C[3][3], double varp, double* scale,
                     double R[9], double T[3]);
double ref_sim3_inlier_t(const float* src, int n);
int ref_sim3_hypothesis(const float* src, const float* dst, int n, const int32_t idx[5], double inlier_t, float hyp[16]);
void ref_sim3_sample5(uint32_t seed_lo, uint32_t seed_hi, uint32_t prob, uint32_t hyp, int n, int32_t idx[5]);
int ref_sim3_ransac(const float* src, const float* dst, int n, int H, float confidence, float min_ratio, uint32_t seed_lo,
                    uint32_t seed_hi, uint32_t prob, const int32_t* sample_idx, float* scale_out, float R_out[9],
                    float t_out[3], int32_t* n_inl, int32_t* best_hyp, int32_t* hyp_counts, uint32_t* bitmap);
void ref_fps(const float* pts, int n, int n_samples, int32_t* out);
int ref_umeyama_fit(const float* src, const float* dst, int n, float* scale_out, float R_out[9], float t_out[3]);

#ifdef __cplusplus
}
#endif
#endif
