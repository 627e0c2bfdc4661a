/* include/poseb200.h -- C ABI of libposeb200.so (sm_100a kernels for pose recovery + evaluation).
 *
 * Drop-in boundary for the pose-recovery / evaluation hot path of yaomy533/pose_estimation.
 * The reference has no FFI of its own: the path is a handful of Python callables backed by
 * third-party native code (OpenCV, KeOps, torch ops).  Each entry point below names the
 * reference call site (file:line under /root/reference) whose native work it replaces.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; every pointer is a DEVICE pointer unless marked host.
 *   - every function returns int: 0 = OK, negative = argument / launch error
 *     (pb_last_error() gives a thread-local message).  No exception crosses the ABI.
 *   - per-frame numerical failure (too few points, no valid hypothesis, too few inliers -- what
 *     OpenCV reports as ok=False and the reference ignores, tools/trainer.py:423) is reported in
 *     status[F], never through the return code.
 *   - the caller owns every buffer including the workspace (pb_*_workspace_bytes()); the library
 *     allocates no device memory and keeps no state between calls except, per calling thread, the error
 *     string and -- only after pb_stage_timing(1) -- that thread's own timing events (a measurement aid,
 *     off by default), plus a per-device cache of the multiprocessor count.
 *   - all work is asynchronous on the given stream (a cudaStream_t passed as void*).
 *   - results are a pure function of (inputs, seed): no dependence on launch geometry.
 */
#ifndef POSEB200_H
#define POSEB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PB_ABI_VERSION 2

/* status codes written to status[F] */
#define PB_STATUS_OK 0
#define PB_STATUS_TOO_FEW_POINTS 1
#define PB_STATUS_NO_MODEL 2

/* rows of the accuracy table (tools/trainer.py:479-509 `result`), each [n_cls] doubles */
#define PB_ROW_ALL_NUM 0
#define PB_ROW_OBJ_NUM 1
#define PB_ROW_VARIANT0 2      /* + 6*variant (0 base, 1 reg, 2 final): succ_rt, dis_rt, succ_r, dis_r, succ_t, dis_t */
#define PB_ROW_DIS_MASK 20
#define PB_ROW_DIS_NORMAL 21
#define PB_ROW_DIS_XYZ 22
#define PB_ROW_FAILED 23       /* frames whose RANSAC status != 0 (new: the reference drops cv2's ok flag) */
#define PB_TABLE_ROWS 24

/* Minimal solver: P3P on 3 correspondences + a 4th to pick the root, ONLY.  SURVEY.md 8(b) sketched a
 * `min_set (4|5)` field for an EPnP-5 option like cv2's flags=SOLVEPNP_EPNP (tools/trainer.py:426); it was not
 * built -- every hypothesis is a P3P solve and the final pose a Gauss-Newton refit on the inliers. */
typedef struct pb_ransac_params {
    int32_t  n_hyp;        /* hypotheses per frame (reference: iterationsCount=100 default of cv2; ours 256) */
    float    thresh_px;    /* reprojectionError (tools/trainer.py:426: 1)                                     */
    int32_t  gn_iters;     /* Gauss-Newton iterations of the inlier refit (cv2: EPnP re-solve on inliers)     */
    int32_t  lo_rounds;    /* extra re-score -> refit rounds ("iterative inlier refine"), 0 = cv2-like        */
    int32_t  min_inliers;  /* fewer inliers than this -> PB_STATUS_NO_MODEL                                   */
    uint32_t seed_lo, seed_hi; /* Philox key; replaces cv::RNG inside solvePnPRansac                          */
} pb_ransac_params;

typedef struct pb_sim3_params {
    int32_t  n_iter;           /* maxIter = 128 (lib/transform/umeyama.py:21)                                     */
    float    confidence;       /* 0.99 (:22): stop once 1 - (1 - ratio^5)^i exceeds it                            */
    float    min_inlier_ratio; /* 0.1 (:50): below it the reference returns None -> PB_STATUS_NO_MODEL            */
    uint32_t seed_lo, seed_hi; /* Philox key; replaces np.random.randint(nPoints, size=5) (:32)                   */
    int32_t  rule;             /* 0 = lib/transform/umeyama.py (best inlier ratio, confidence break);
                                  1 = the older lib/transform/trans.py:47-90 (smallest total residual wins, stop below
                                  PassThreshold / 100, thresholds from the clouds' mean norms; n_iter = 100 there;
                                  confidence is unused; hyp_counts = its nInliers, which skips index 0)            */
} pb_sim3_params;

const char* pb_last_error(void);
int pb_abi_version(void);

/* ---- K0: coordinate-map stream -> compacted correspondences --------------------------------
 * Replaces tools/trainer.py:386-421 (D2H of the map, de-normalisation xyz*extent+lfborder,
 * gather at `choose`) and the dataset-side mask.flatten().nonzero() (batchdataset.py:673).
 * xyz [F,3,Hh,Ww] f32 planar (pred['xyz']); mask [F,Hh,Ww] u8 or NULL (-> pixel is foreground iff
 * any channel != 0, the map's "0 = background" convention, batchdataset.py:688-693);
 * roi [F,2] (rmin,cmin) added to row/col to give image v,u, or NULL; extent/lfborder [F,3].
 * pts [F,5,cap] SoA (X,Y,Z,U,V) in row-major pixel order; count[F] = min(total, cap);
 * total[F] optional = number of foreground pixels.                                              */
size_t pb_compact_workspace_bytes(int F, int Hh, int Ww, int cap);
int pb_compact_map(const float* xyz, const uint8_t* mask, const int32_t* roi, const float* extent,
                   const float* lfborder, int F, int Hh, int Ww, int cap, float* pts,
                   int32_t* count, int32_t* total, void* ws, size_t ws_bytes, void* stream);

/* pb_compact_map_noalloc -- NO COUNTERPART IN THE REFERENCE; added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The mask-first path of pb_compact_map (same kernel, same output bits) with its scratch -- a ticket and one state
 * word per 65,536-pixel mask tile -- in a file-scope device array of the library instead of a caller workspace: the
 * call allocates nothing and takes no workspace.  Limits: a mask is required, Hh*Ww % 16 == 0, mask 16-byte aligned,
 * and F * ceil(Hh*Ww / 65536) <= pb_compact_map_noalloc_max_tiles() (65,536 tiles = 256 KB per GPU, held from
 * library load to exit; 13,107 frames of 480 x 640); larger batches return -1 with a message.  The scratch is SHARED
 * by all calls on a device and keeps the last call's contents: every call clears what it uses on its own stream
 * first, and calls that overlap on the GPU (different streams) race -- serialise them.                          */
int pb_compact_map_noalloc_max_tiles(void);
int pb_compact_map_noalloc(const float* xyz, const uint8_t* mask, const int32_t* roi, const float* extent,
                           const float* lfborder, int F, int Hh, int Ww, int cap, float* pts, int32_t* count,
                           int32_t* total, void* stream);

/* gather variant for the reference's `choose` index list (tools/trainer.py:408-421):
 * choose [F,NP] int64 flat indices into Hh*Ww, xmap/ymap [F,NP] f32 pixel coordinates.        */
int pb_gather_correspondences(const float* xyz, const int64_t* choose, const float* xmap,
                              const float* ymap, const float* extent, const float* lfborder, int F,
                              int Hh, int Ww, int NP, int cap, float* pts, int32_t* count, void* stream);

/* The same gather for every form of the list the reference's data pipeline produces (batchdataset.py:673-687,
 * 712-716): choose [F,NP] int64 (choose_is_i32 = 0) or int32; xmap / ymap NULL -> the pixel coordinates are taken
 * from the index itself (col + cmin, row + rmin with roi [F,2] = (rmin, cmin), or 0); n_valid [F] NULL -> every
 * list has NP entries (the reference wrap-pads to NUM_POINTS), else list f has n_valid[f] <= NP.  xyz may be
 * device memory or PINNED HOST memory read in place: only the sectors of the chosen pixels cross PCIe.        */
int pb_gather_choose(const float* xyz, const void* choose, int choose_is_i32, const float* xmap,
                     const float* ymap, const int32_t* n_valid, const int32_t* roi, const float* extent,
                     const float* lfborder, int F, int Hh, int Ww, int NP, int cap, float* pts,
                     int32_t* count, void* stream);

/* ---- region-coded maps: xyz = xyz_off + base(region) ------------------------------------------
 * KRRN predicts an offset map plus per-pixel region logits; the map the solver consumes is rebuilt from both.
 * soft = 0: tools/script/eval.py:93-103 (tools/trainer.py:391-398, commented out there) -- class = arg-max of the
 *           softmax, i.e. of the logits (lowest index on ties), base = region_point[class];
 * soft = 1: lib/network/loss_utils.py:18-28 get_xyz -- base = mean_k softmax(region)[k] * region_point[k].
 * xyz_off [F,3,Hh,Ww], region [F,n_region,Hh,Ww], region_point [F,n_region,3] f32.
 * pb_region_decode_points rewrites X,Y,Z of compacted correspondences pts [F,5,cap] / count [F] in place
 * (X = (xyz_off + base) * extent + lfborder): the pixel of entry i is choose[f,i] (choose [F,NP] int64, the list
 * pb_gather_correspondences took) or, with choose NULL, (V - rmin) * Ww + (U - cmin) from its own U,V and roi
 * (pb_compact_map output).  Only foreground pixels touch the logit planes.
 * pb_region_decode_map is the dense form: xyz [F,3,Hh,Ww] = xyz_off + base at every pixel.                       */
int pb_region_decode_points(const float* xyz_off, const float* region, const float* region_point,
                            const int64_t* choose, int NP, const int32_t* roi, const float* extent,
                            const float* lfborder, int F, int Hh, int Ww, int n_region, int cap, int soft,
                            float* pts, const int32_t* count, void* stream);
int pb_region_decode_map(const float* xyz_off, const float* region, const float* region_point, int F, int Hh,
                         int Ww, int n_region, int soft, float* xyz, void* stream);

/* ---- K1-K4: PnP-RANSAC --------------------------------------------------------------------
 * Replaces cv2.solvePnPRansac(obj, img, K, EPNP, conf .9999, 1 px) at tools/trainer.py:423-427,
 * lib/utils/utlis.py:59-63, tools/script/eval.py:131-135, plus Rodrigues (trainer.py:433).
 * pts [F,5,cap], count [F], K4 [F,4] (fx,fy,cx,cy; lib/utils/utlis.py:12-34 layout).
 * sample_idx optional [F,H,4] int32 (forces the minimal sets; else Philox(seed,frame0+f,hyp)).
 * Outputs: R [F,9] row-major, t [F,3], n_inl [F] (inliers of the best minimal hypothesis),
 * best_hyp [F], n_inl_refined [F] (inliers of the returned pose), status [F];
 * optional hyp_counts [F,H] (-1 = no model), inl_bitmap [F,ceil(cap/32)] (inliers of the best
 * minimal hypothesis -- the set cv2 returns as `inliers`), R0 [F,9] / t0 [F,3] (its pose).      */
size_t pb_ransac_workspace_bytes(int F, int cap, const pb_ransac_params* prm);
int pb_pnp_ransac_pts(const float* pts, const int32_t* count, const float* K4, int F, int cap,
                      uint32_t frame0, const pb_ransac_params* prm /*host*/, const int32_t* sample_idx,
                      float* R, float* t, int32_t* n_inl, int32_t* best_hyp, int32_t* n_inl_refined,
                      int32_t* status, int32_t* hyp_counts, uint32_t* inl_bitmap, float* R0, float* t0,
                      void* ws, size_t ws_bytes, void* stream);

/* Which kernel scores the hypotheses (K3; the iteration loop of cv2.solvePnPRansac, tools/trainer.py:423-427).
 * Process-wide setting (not thread-safe; set it before launching work) read at every pb_pnp_ransac_* call:
 *   0 = automatic (the FP32-pipe kernels), 1 = FP32-pipe kernels only,
 *   2 = projection by tcgen05 + TMEM (csrc/ransac_score_tc.cu), 3 = the same with a calibration pass, 4 = with a clock
 *       breakdown of its warps' waits,
 *   5 = projection by warp-level mma.sync (csrc/ransac_score_mma.cu), 6 = the same with a calibration pass.
 * 2 - 6 apply whenever the shape allows (n_hyp <= 256 and a batch large enough for the balanced plan), otherwise the
 * FP32-pipe kernels run.  All of them return identical integers and poses; 2 and 5 are measured alternatives that are
 * not faster on B200 (profiles/r02_k3_tc.md), which is why 0 does not choose them.
 * band_scale multiplies the band inside which the tensor-core kernels re-evaluate a test with the FP32 contract
 * (1 = the derived bound).  pb_ransac_tc_stats_offset: byte offset inside the pb_pnp_ransac_pts workspace of
 * int32 stats[4] = {barrier timeout, re-evaluated blocks, calibration: largest |s_tc - s_fp32| / band as float
 * bits, calibration: verdicts that differ outside the band} followed by twelve 64-bit clock sums (mode 4), or -1 when
 * the shape never takes those kernels.                                                                             */
int pb_ransac_score_impl(int impl, float band_scale);
long long pb_ransac_tc_stats_offset(int F, int cap, const pb_ransac_params* prm);

/* map -> pose in one call (K0 + K1-K4); workspace holds the compacted points.                 */
size_t pb_pnp_ransac_map_workspace_bytes(int F, int Hh, int Ww, int cap, const pb_ransac_params* prm);
int pb_pnp_ransac_map(const float* xyz, const uint8_t* mask, const int32_t* roi, const float* extent,
                      const float* lfborder, const float* K4, int F, int Hh, int Ww, int cap,
                      uint32_t frame0, const pb_ransac_params* prm /*host*/, float* R, float* t,
                      int32_t* n_inl, int32_t* best_hyp, int32_t* n_inl_refined, int32_t* status,
                      void* ws, size_t ws_bytes, void* stream);

/* inlier bitmap -> exactly n indices (order-preserving random subset / wrap pad);
 * replaces lib/utils/utlis.py:64-71.  out [F,n] int32 (-1 if the frame has no inliers).       */
int pb_resample_inliers(const uint32_t* inl_bitmap, const int32_t* count, int F, int cap, int n,
                        uint32_t frame0, uint32_t seed_lo, uint32_t seed_hi, int32_t* out, void* stream);

/* ---- 3D-3D RANSAC + Umeyama similarity fit (lib/transform/umeyama.py:8-97; trans.py:9-123) --------
 * For each of B problems: target ~ scale * R * source + t from `count[b]` corresponding pairs
 * src/dst [B,cap,3].  n_iter hypotheses of 5 pairs drawn WITH replacement (sample_idx [B,n_iter,5]
 * overrides the RNG), Umeyama per hypothesis, inlier iff |dst - (sR src + t)| < scale * diameter/10,
 * the reference's sequential best-ratio / confidence-break rule replayed over the counts, then the
 * Umeyama refit on the inliers.  Outputs: scale [B], R [B,9], t [B,3], n_inl, best_hyp (iteration
 * index), status; optional hyp_counts [B,n_iter], inl_bitmap [B,ceil(cap/32)].                   */
size_t pb_sim3_ransac_workspace_bytes(int B, const pb_sim3_params* prm);
int pb_sim3_ransac(const float* src, const float* dst, const int32_t* count, int B, int cap,
                   uint32_t prob0, const pb_sim3_params* prm /*host*/, const int32_t* sample_idx,
                   float* scale, float* R, float* t, int32_t* n_inl, int32_t* best_hyp, int32_t* status,
                   int32_t* hyp_counts, uint32_t* inl_bitmap, void* ws, size_t ws_bytes, void* stream);
/* batched estimateSimilarityUmeyama (umeyama.py:65-97): least-squares similarity over all pairs. */
int pb_umeyama_fit(const float* src, const float* dst, const int32_t* count, int B, int cap, float* scale,
                   float* R, float* t, int32_t* status, void* stream);

/* ---- batched rigid fit + ICP-style refinement (version/transparent/lib/declarative/leastsquares.py:10-48,
 *      version/transparent/scripts/train_declarative.py:80-100) ------------------------------------------------
 * pb_kabsch_fit = best_batch_fit_transform: for each of B problems the rotation R [B,9] and translation t [B,3]
 * with  Bm ~ R A + t  in the least-squares sense (proper rotation: the reflection case of :38-43 included), and
 * the two centroids (optional).  A / Bm [B,cap,3], count [B] or NULL (-> cap points each).  status optional:
 * PB_STATUS_NO_MODEL where the points have no spread (then R = I, t = centroid shift).
 * pb_icp_refine: n_steps times { pred = model R^T + t; idx = nearest target point of every pred point
 * (pb_knn1 / the tensor-core path when use_tc); (opt_R, opt_t) = rigid fit pred -> target[idx]; compose } with
 * R [B,9] / t [B,3] updated in place.  compose 0 is the reference's update as written (:98-99: R <- R opt_R,
 * t <- t + opt_t); compose 1 is the rigid-motion composition (R <- opt_R R, t <- opt_R t + opt_t).
 * model [B or 1,N,3], target [B or 1,M,3]; ws: pb_icp_refine_workspace_bytes, 128-byte aligned.              */
int pb_kabsch_fit(const float* A, const float* Bm, const int32_t* count, int B, int cap, float* R, float* t,
                  float* centroid_A, float* centroid_B, int32_t* status, void* stream);
size_t pb_icp_refine_workspace_bytes(int B, int N, int M);
int pb_icp_refine(const float* model, int model_shared, const float* target, int target_shared, float* R, float* t,
                  int B, int N, int M, int n_steps, int compose, int use_tc, int32_t* status, void* ws,
                  size_t ws_bytes, void* stream);

/* ---- depth back-projection (dataset/linemod/batchdataset.py:712-721, lib/transform/coordinate.py:91-101) ------
 * pb_backproject_cloud: cloud[f,i,:] = ((x - cx) d / fx, (y - cy) d / fy, d) with d = depth[f, choose[f,i]] * depth_scale
 * -- the `cloud` the T-offset vote consumes, made on the device from the depth crop and the same choose /
 * x_map / y_map lists as pb_gather_choose (xmap / ymap NULL -> from the index and roi).  depth [F,Hh,Ww] f32.
 * Float32 operation order of the reference: (x - cx) * d / fx.
 * pb_uvd2cloud: the list form, uvd [B,n,3] (u, v, depth) and one camera matrix K [9] (row major) -> xyz [B,n,3]. */
int pb_backproject_cloud(const float* depth, const void* choose, int choose_is_i32, const float* xmap,
                         const float* ymap, const int32_t* roi, const float* K4, float depth_scale, int F, int Hh,
                         int Ww, int NP, float* cloud, void* stream);
int pb_uvd2cloud(const float* uvd, const float* K9, int B, int n, float* xyz, void* stream);

/* ---- farthest point sampling (tools/script/sample_model.py:33-63) ----------------------------------
 * For each of B clouds pts [B,cap,3] (count[b] valid points, NULL -> cap): idx[b, 0] = 0, then
 * repeatedly the point farthest from the chosen set; float32 sqrt((dx*dx+dy*dy)+dz*dz) distances
 * and np.argmax's first-maximum rule, so the indices equal the reference's.  idx [B,n_samples]
 * int32 (-1 rows for empty clouds).  ws only for cap > 57344 (pb_fps_workspace_bytes).           */
size_t pb_fps_workspace_bytes(int B, int cap);
int pb_fps(const float* pts, const int32_t* count, int B, int cap, int n_samples, int32_t* idx, void* ws,
           size_t ws_bytes, void* stream);

/* ---- back-propagatable PnP (lib/network/dnn/BPnP.py) -----------------------------------------------
 * pb_pnp_refine replaces the cv2.solvePnP(SOLVEPNP_ITERATIVE, useExtrinsicGuess=True) call of
 * BPnP.forward (BPnP.py:43-44): Levenberg-Marquardt over (angle-axis, t) on ALL n correspondences
 * pts2d [B,n,2] <-> pts3d ([n,3], or [B,n,3] when pts3d_batched), camera matrix K [9] row major
 * (general 3x3), from pose6_init [B,6] or, when that is NULL, from R_init [B,9] + t_init [B,3]
 * (the pb_pnp_ransac_pts pose stands in for BPnP.py:36-38).  FP64 inside.  pose6 [B,6] =
 * (angle-axis, t); rmse [B] optional (px); status: 0 converged, 1 iteration limit, 2 degenerate. */
int pb_pnp_refine(const float* pts2d, const float* pts3d, int pts3d_batched, const float* K,
                  const float* pose6_init, const float* R_init, const float* t_init, int B, int n,
                  int max_iter, float* pose6, float* rmse, int32_t* status, void* stream);
/* pb_bpnp_backward replaces BPnP.backward (BPnP.py:54-117): grad_pts2d [B,n,2], grad_pts3d ([n,3]
 * summed over the batch, or [B,n,3] when pts3d_batched), grad_K [9] summed over the batch, from
 * grad_out [B,6] by the implicit function theorem on f_j = sum_i coef_ij . r_i (get_coefs
 * BPnP.py:127-138; rotation = the angle-axis formula the reference differentiates, eps 1e-6).
 * status[b] = 1 (and zero contribution) where J_fy is singular -- the reference raises there.  */
size_t pb_bpnp_backward_workspace_bytes(int B, int n);
int pb_bpnp_backward(const float* pts2d, const float* pts3d, int pts3d_batched, const float* K,
                     const float* pose6, const float* grad_out, int B, int n, float* grad_pts2d,
                     float* grad_pts3d, float* grad_K, int32_t* status, void* ws, size_t ws_bytes,
                     void* stream);

/* ---- per-hypothesis pose loss (version/transparent/lib/networks/loss.py:197-267, PossLoss) -----------
 * H = bs*num_p hypotheses: pred_r [H,4] quaternion (w,x,y,z, any norm), pred_t [H,3], pred_c [H];
 * model / target [bs,N,3]; axis3 [3] and target_r9 [9] are device arrays (the reference's `axis`
 * [1,3] and `target_r` [1,3,3]); w = the confidence regulariser.  nn_mode != 0 is the symmetric, non-refine
 * branch (loss.py:224-233): every predicted point is matched to its nearest point of target[0]
 * (FP32 direct-difference SqDist, lowest index on ties -- the KeOps argmin, fused here).
 * vals3 (device) = {loss, dis[0][argmax pred_c[0]], loss_rotation}.  The workspace keeps what
 * pb_poss_loss_bwd needs (pass the same buffer, untouched); grad_vals3 is a DEVICE array.        */
size_t pb_poss_loss_workspace_bytes(int bs, int num_p, int N);
int pb_poss_loss_fwd(const float* pred_r, const float* pred_t, const float* pred_c, const float* model,
                     const float* target, int bs, int num_p, int N, int nn_mode, float w,
                     const float* axis3, const float* target_r9, float* vals3, void* ws, size_t ws_bytes,
                     void* stream);
int pb_poss_loss_bwd(const float* pred_r, const float* pred_t, const float* pred_c, const float* model,
                     const float* target, int bs, int num_p, int N, int nn_mode, float w,
                     const float* axis3, const float* target_r9, const float* grad_vals3, const void* ws,
                     size_t ws_bytes, float* grad_r, float* grad_t, float* grad_c, void* stream);

/* ---- K5: nearest neighbour / ADD-S -----------------------------------------------------------
 * pb_knn1 replaces the KeOps handle generic_argkmin('SqDist(x, y)','a = Vi(1)','x = Vi(3)',
 * 'y = Vj(3)') (train.py:126; used lib/network/loss.py:34, version/transparent/lib/utils.py:24):
 * idx[b,i] = argmin_j sum_c (x[b,i,c]-y[b,j,c])^2 in FP32, lowest j on ties.
 * x [B,N,3], y [B,M,3] (y_batch_stride 0 shares one y), idx [B,N] int64, d2 [B,N] optional.    */
int pb_knn1(const float* x, const float* y, int B, int N, int M, int y_shared, int64_t* idx, float* d2,
            void* stream);

/* Tensor-core path of the same handle: tiled tcgen05 contraction (|q|^2 - 2 p.q from two-way bf16
 * splits of the centred coordinates, K = 16, FP32 accumulate in TMEM) + exact FP32 re-rank of every
 * column group within the proven error bound of the row minimum, so idx / d2 are identical to
 * pb_knn1's (csrc/knn_tc.cu).  ws: pb_knn1_tc_workspace_bytes(B,N,M), 128-byte aligned; after the
 * call its first 16 bytes hold int stats[4] = {barrier timeout, rows that fell back to the exact
 * scan, exact evaluations, 0}.                                                                  */
size_t pb_knn1_tc_workspace_bytes(int B, int N, int M);
int pb_knn1_tc(const float* x, const float* y, int B, int N, int M, int y_shared, int64_t* idx, float* d2,
               void* ws, size_t ws_bytes, void* stream);
/* Which sweep pb_knn1_tc / pb_add_adds_tc / pb_pose_loss_fwd_tc use (process-wide): 0 = automatic (tcgen05 + TMEM),
 * 1 = tcgen05 + TMEM, 2 = warp-level mma.sync (a measured alternative, slower on B200).  Identical indices and
 * distances; no counterpart in the reference (one KeOps handle).                                                  */
int pb_knn_tc_impl(int impl);

/* replaces Metric.cal_adds_cuda (lib/utils/metric.py:17-36): add[b] = mean_i |pred_i - tgt_i|;
 * adds[b] = mean_i min_j |pred_j - tgt_i| computed where want_adds[b] != 0 (NULL -> all),
 * else adds[b] = add[b].  pred/target [B,N,3].                                                  */
int pb_add_adds(const float* pred, const float* target, int B, int N, const uint8_t* want_adds,
                float* add, float* adds, void* stream);
/* The same metric with the N x N search on the tensor cores (pb_knn1_tc's kernels); add / adds are
 * bit-identical to pb_add_adds.  sym_list: device list of the n_sym batch elements with
 * want_adds[b] != 0 (want_adds NULL -> every element, list ignored).  ws: 128-byte aligned.     */
size_t pb_add_adds_tc_workspace_bytes(int B, int N);
int pb_add_adds_tc(const float* pred, const float* target, int B, int N, const uint8_t* want_adds,
                   const int* sym_list, int n_sym, float* add, float* adds, void* ws, size_t ws_bytes,
                   void* stream);

/* out[b,i,:] = R[b] * model[b or 0,i,:] + t[b]   (tools/trainer.py:376, lib/network/loss.py:26) */
int pb_transform_points(const float* model, int model_shared, const float* R, const float* t, int B,
                        int N, float* out, void* stream);

/* ---- K6: per-frame metrics + accuracy table ---------------------------------------------------
 * Replaces Trainer.cal_dis (tools/trainer.py:370-381) + Metric.angular_distance /
 * translation_distance (lib/utils/metric.py:67-84).  model [C,N,3] per-class clouds, cls [F];
 * target [F,N,3] or NULL (-> model*R_gt^T + t_gt, batchdataset.py:707).  For classes with
 * is_sym[c] != 0 add[f] is ADD-S (metric.py:24-32).  Outputs all [F].                          */
int pb_pose_metrics(const float* R, const float* t, const float* R_gt, const float* t_gt,
                    const float* model, const float* target, const int32_t* cls,
                    const uint8_t* is_sym, int F, int N, int C, float* add, float* r_deg, float* t_m,
                    void* stream);

/* success tests + per-object accumulation (tools/trainer.py:156-157,174-221): table [24,C] doubles
 * += this batch (deterministic, one CTA per class).  add < add_frac*diameter[c], r < r_thr, t < t_thr.
 * A class index outside [0,C) is clamped into range by pb_pose_metrics AND here (same column in both), and the
 * frame is counted in PB_ROW_FAILED, so all_num always sums to the number of frames seen.                  */
int pb_accumulate_results(const float* add, const float* r_deg, const float* t_m, const int32_t* cls,
                          const int32_t* status, const float* diameter, int F, int C, int variant,
                          float add_frac, float r_thr_deg, float t_thr_m, int count_frames,
                          double* table, void* stream);

/* ---- symmetric-KNN pose loss (lib/network/loss.py:25-41) --------------------------------------
 * forward: pred = model*R^T + t; tgt_i <- tgt[knn(pred_i)] for symmetric samples;
 * loss_b = mean_i |pred_i - tgt_i|; loss = mean_b loss_b.  nn_idx [B,N] int64 scratch/out
 * (identity rows for non-symmetric samples).  backward: closed form wrt R [B,9] and t [B,3].   */
int pb_pose_loss_fwd(const float* R, const float* t, const float* model, const float* target,
                     const uint8_t* sym /*[B]*/, int B, int N, float* pred /*[B,N,3]*/, int64_t* nn_idx,
                     float* loss_b /*[B]*/, void* stream);
/* forward with the symmetric samples' search on the tensor cores; sym_list = device list of the
 * n_sym elements with sym[b] != 0 (sym NULL -> all symmetric, list ignored);
 * ws = pb_knn1_tc_workspace_bytes(B,N,N), 128-byte aligned.                                     */
int pb_pose_loss_fwd_tc(const float* R, const float* t, const float* model, const float* target,
                        const uint8_t* sym, const int* sym_list, int n_sym, int B, int N, float* pred,
                        int64_t* nn_idx, float* loss_b, void* ws, size_t ws_bytes, void* stream);
int pb_pose_loss_bwd(const float* pred, const float* model, const float* target, const int64_t* nn_idx,
                     int B, int N, float gscale, float* grad_R /*[B,9]*/, float* grad_t /*[B,3]*/,
                     void* stream);

/* ---- K7: T-offset voting (lib/network/krrn.py:153) --------------------------------------------
 * t[b,:] = mean_i (cloud[b,i,:] + t_res[b,:,i]);  cloud [B,NP,3], t_res [B,3,NP].               */
int pb_vote_translation(const float* cloud, const float* t_res, int B, int NP, float* t, void* stream);

/* ---- measurement probe (no reference counterpart) ----------------------------------------------
 * Sustained FP32 FMA rate: ctas CTAs x 256 threads x iters x 8 independent FMAs (x2 lanes when
 * packed != 0, FFMA2).  Timed by the caller with CUDA events; bench.py uses it as the denominator
 * of the scoring kernel's FP32 roofline (MEASURED_PEAKS.json has HBM and bf16 peaks only).        */
int pb_probe_fma(int ctas, int iters, int packed, float* out, void* stream);

/* Device-side generator of synthetic LineMOD-shaped frames (measurement infrastructure for the sharded sweep,
 * SURVEY.md 8(d) config C5; csrc/synthgen.cu).  params [F,48] f32 per-frame pose / camera / ellipsoid / bounding box
 * (pose_estimation_b200.synth.c5_params), bbox_max = largest bounding-box area.  erase = 0 writes the foreground
 * pixels of F frames into xyz [F,3,Hh,Ww] / mask [F,Hh,Ww], which must be ZERO elsewhere, and remembers them in ws
 * (pb_synth_workspace_bytes(F); int32 [F] unused, then [F] numbers of written pixels); erase != 0 zeroes exactly
 * those pixels.  Philox4x32-10 keyed (seed), counter (pixel, frame0 + f); IEEE float32 throughout, so the NumPy
 * twin (synth.c5_frame_twin) gives identical bytes.                                                            */
size_t pb_synth_workspace_bytes(int F);
int pb_synth_frames(const float* params, int F, int Hh, int Ww, int bbox_max, uint32_t seed_lo, uint32_t seed_hi,
                    uint32_t frame0, float* xyz, uint8_t* mask, void* ws, size_t ws_bytes, int erase, void* stream);

/* tcgen05 self test: D[128,128] f32 = A[128,32] * B[128,32]^T for bf16 operands given in the canonical
 * no-swizzle K-major tile layout [kc 0..3][row 0..127][8 bf16] (the layout pb_knn1's tensor-core
 * path stages); err (device int) is set when the completion barrier timed out.                     */
int pb_tc_selftest(const void* A, const void* B, float* D, int* err, void* stream);

/* Per-kernel timing of pb_pnp_ransac_pts, PER CALLING THREAD: while enabled, every call made by this thread
 * records CUDA events on its stream around its three launches (hypotheses | scoring | finalize; up to 256 calls).
 * pb_stage_timing_read waits for them and returns the mean milliseconds per call (host pointers). */
int pb_stage_timing(int enable);
int pb_stage_timing_read(float* ms3 /*host [3]*/, int* calls /*host*/);

#ifdef __cplusplus
}
#endif
#endif
