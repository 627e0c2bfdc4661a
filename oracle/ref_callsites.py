"""oracle/ref_callsites.py -- reference-behaviour oracle (TEST INFRASTRUCTURE ONLY).

CPU restatement of the reference's Python call sites for the pose-recovery + evaluation path,
around the REAL third-party arithmetic where it is available in this image:

* ``cv2.solvePnPRansac`` (OpenCV 4.13.0 here; the reference pins no version) is called exactly as
  ``tools/trainer.py:423-427`` / ``lib/utils/utlis.py:59-63`` call it.
* kornia (absent) is replaced by ``cv2.Rodrigues`` for ``angle_axis_to_rotation_matrix``
  (``tools/trainer.py:433``) and a trace-based R->quaternion for ``rotation_matrix_to_quaternion``
  (``lib/utils/metric.py:70-72``; the sign ambiguity is irrelevant because ``angular_distance``
  takes ``|q1.q2|``, ``metric.py:78``).
* pykeops (absent) ``generic_argkmin('SqDist(x, y)', 'a = Vi(1)', ...)`` (``train.py:126``) is
  replaced by an FP32 brute-force argmin with first-index tie-break.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import this module.  Nothing here is used by the product path.  Parity status: pinned by
running the reference's OWN functions in this container (tests/golden/gen_golden_callsites.py executes
Trainer.get_pose / cal_dis, lib/utils/metric.py, lib/utils/utlis.py and PoseLoss from /root/reference on
seeded inputs -> callsites_ref.npz) and checking every function here against those outputs
(tests/test_oracle_golden.py); the stand-ins for kornia / KeOps named above are what stays unpinned.
tests/golden/pnp_cv2.npz / metric.npz are further vectors produced by this file itself.
"""
from __future__ import annotations

import numpy as np

try:  # cv2 is present in this image; keep the import failure explicit for other hosts
    import cv2
except Exception as e:  # pragma: no cover
    cv2 = None
    _cv2_err = e


def _need_cv2():
    if cv2 is None:  # pragma: no cover
        raise RuntimeError(f"cv2 unavailable: {_cv2_err}")


# ----------------------------------------------------------------------------- intrinsics
def batch_intrinsic_transform(K):
    """lib/utils/utlis.py:12-34 -- [bs,4] (fx,fy,cx,cy) <-> [bs,3,3]."""
    K = np.asarray(K)
    if K.ndim == 2 and K.shape[-1] == 4:
        out = np.zeros((K.shape[0], 3, 3), dtype=K.dtype)
        out[:, 0, 0] = K[:, 0]
        out[:, 1, 1] = K[:, 1]
        out[:, 0, 2] = K[:, 2]
        out[:, 1, 2] = K[:, 3]
        out[:, 2, 2] = 1
        return out
    if K.ndim == 3 and K.shape[-1] == 3:
        return np.stack([K[:, 0, 0], K[:, 1, 1], K[:, 0, 2], K[:, 1, 2]], axis=1)
    raise Exception('Wrong Shape!')


# ----------------------------------------------------------------------------- get_pose
def solve_pnp_ransac(obj, img, K33):
    """The exact cv2 call of tools/trainer.py:423-427 (flags/confidence/threshold are the
    reference's hard-coded constants; iterationsCount is OpenCV's default 100)."""
    _need_cv2()
    ok, rvec, tvec, inl = cv2.solvePnPRansac(
        objectPoints=np.ascontiguousarray(obj), imagePoints=np.ascontiguousarray(img),
        cameraMatrix=np.ascontiguousarray(K33), distCoeffs=None,
        flags=cv2.SOLVEPNP_EPNP, confidence=0.9999, reprojectionError=1)
    return ok, rvec, tvec, inl


def get_pose(xyz, extent, lfborder, choose, x_map_choosed, y_map_choosed, intrinsic,
             num_points=256, rng=None, pick=None):
    """tools/trainer.py:383-438 for one frame (the reference asserts bs == 1).

    xyz [3,h,w] f32 normalised map; extent/lfborder [3] float64; choose [NP] flat indices into
    h*w; x/y_map_choosed [NP] pixel coordinates; intrinsic [4].  ``rng`` replaces the global torch
    CPU RNG behind ``torch.randperm`` (:407); ``pick`` gives the drawn positions outright (how the
    golden vectors recorded from the reference's own get_pose are replayed).
    Returns (R [3,3] f32, t [3] f32, ok, inliers)."""
    choose = np.asarray(choose).reshape(-1)
    if pick is None:
        rng = rng or np.random.default_rng()
        pick = rng.permutation(choose.shape[0])[:num_points]                  # :407
    pick = np.asarray(pick).reshape(-1)
    chosen = choose[pick]                                                     # :408
    pts2d = np.stack([np.asarray(x_map_choosed).reshape(-1)[pick],
                      np.asarray(y_map_choosed).reshape(-1)[pick]], axis=-1).astype(np.float32)  # :410-413
    K33 = batch_intrinsic_transform(np.asarray(intrinsic, dtype=np.float32).reshape(1, 4))[0]     # :414
    # :415-419 -- float32 map times float64 extent/lfborder promotes to float64
    ext = np.asarray(extent, dtype=np.float64).reshape(3, 1, 1)
    lfb = np.asarray(lfborder, dtype=np.float64).reshape(3, 1, 1)
    coord = (np.asarray(xyz, dtype=np.float32) * ext + lfb).transpose(1, 2, 0).reshape(-1, 3)
    obj = coord[chosen]                                                       # :421
    ok, rvec, tvec, inl = solve_pnp_ransac(obj, pts2d, K33)
    if not ok:
        # the reference ignores `ok` and would read uninitialised rvec
        return np.eye(3, dtype=np.float32), np.zeros(3, np.float32), False, None
    rv = np.asarray(rvec, dtype=np.float32).reshape(3)                        # :429
    R, _ = cv2.Rodrigues(rv.astype(np.float64))                               # :433 (kornia)
    return R.astype(np.float32), np.asarray(tvec, dtype=np.float32).reshape(3), True, inl


# ----------------------------------------------------------------------------- region-coded maps
def _softmax0(region):
    r = np.asarray(region, np.float32)
    e = np.exp(r - r.max(axis=0, keepdims=True), dtype=np.float32)
    return e / e.sum(axis=0, keepdims=True, dtype=np.float32)


def region_xyz_hard(xyz_off, region, region_point):
    """tools/script/eval.py:93-103 for one frame: xyz_off [3,h,w], region [n,h,w] logits, region_point [n,3]
    -> xyz [3,h,w] = xyz_off + region_point[argmax_k softmax(region)_k] (float32)."""
    cls = _softmax0(region).argmax(axis=0)                                      # :94-95
    base = np.asarray(region_point, np.float32)[cls]                            # :99-100 (gather)  [h,w,3]
    return np.asarray(xyz_off, np.float32) + base.transpose(2, 0, 1)            # :102


def get_xyz(xyz_off, region, region_point):
    """lib/network/loss_utils.py:18-28 for one frame: softmax-weighted base, MEAN over the n regions."""
    sm = _softmax0(region)                                                      # :24
    rp = np.asarray(region_point, np.float32)
    base = (sm[:, None, :, :] * rp[:, :, None, None]).mean(axis=0, dtype=np.float32)   # :27
    return np.asarray(xyz_off, np.float32) + base


# ----------------------------------------------------------------------------- metric
def rot_to_quat_wxyz(R):
    """Trace-based rotation matrix -> quaternion (w,x,y,z); stands in for kornia
    rotation_matrix_to_quaternion(order=WXYZ) used by lib/utils/metric.py:70-72."""
    R = np.asarray(R, dtype=np.float64)
    tr = R[0, 0] + R[1, 1] + R[2, 2]
    if tr > 0:
        s = np.sqrt(tr + 1.0) * 2
        q = [0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s]
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = np.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = [(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s]
    elif R[1, 1] > R[2, 2]:
        s = np.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = [(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s]
    else:
        s = np.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = [(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s]
    q = np.asarray(q)
    return q / max(np.linalg.norm(q), 1e-12)


class Metric:
    """lib/utils/metric.py:13-84."""

    def __init__(self, sys):
        self.sys = list(sys)

    def cal_adds_cuda(self, pred, target, idx):
        """metric.py:17-36.  ADD = mean_i ||pred_i - target_i||; for a symmetric class
        ADD-S = mean_i min_j ||pred_j - target_i|| replaces it (FP32, like the torch code)."""
        pred = np.asarray(pred, dtype=np.float32)
        target = np.asarray(target, dtype=np.float32)
        assert pred.ndim == target.ndim == 2
        add = np.linalg.norm(pred - target, axis=1).astype(np.float32).mean(dtype=np.float32)
        if idx in self.sys:
            best = np.full(target.shape[0], np.inf, dtype=np.float32)
            for s in range(0, pred.shape[0], 512):  # [N,N,3] of :27-29 in slabs
                d = np.linalg.norm(pred[None, s:s + 512, :] - target[:, None, :], axis=2).astype(np.float32)
                best = np.minimum(best, d.min(axis=1))
            adds = best.mean(dtype=np.float32)
            add = adds
        else:
            adds = add
        return float(add), float(adds)

    def cal_auc(self, add_dis, max_dis=0.1):
        """metric.py:38-50."""
        D = np.array(add_dis, dtype=np.float64)
        D[D > max_dis] = np.inf
        D = np.sort(D)
        n = len(add_dis)
        acc = np.cumsum(np.ones((1, n)), dtype=np.float32) / n
        return self.voc_ap(D, acc) * 100.0

    @staticmethod
    def voc_ap(rec, prec):
        """metric.py:52-65."""
        keep = np.where(rec != np.inf)
        if len(keep[0]) == 0:
            return 0
        rec = rec[keep]
        prec = prec[keep]
        mrec = np.array([0.0] + list(rec) + [0.1])
        mpre = np.array([0.0] + list(prec) + [prec[-1]])
        for i in range(1, prec.shape[0]):
            mpre[i] = max(mpre[i], mpre[i - 1])
        i = np.where(mrec[1:] != mrec[0:-1])[0] + 1
        return np.sum((mrec[i] - mrec[i - 1]) * mpre[i]) * 10

    @staticmethod
    def angular_distance(R1, R2, eps=1e-7):
        """metric.py:67-80 for one pair of rotation matrices -> degrees."""
        q1 = rot_to_quat_wxyz(R1)
        q2 = rot_to_quat_wxyz(R2)
        dot = abs(float(np.dot(q1, q2)))
        dot = min(max(dot, -1.0 + eps), 1.0 - eps)
        return 2.0 * np.arccos(dot) / np.pi * 180.0

    @staticmethod
    def translation_distance(t1, t2):
        """metric.py:82-84."""
        return float(np.linalg.norm(np.asarray(t1, np.float32) - np.asarray(t2, np.float32)))


def cal_dis(metric, pred_r, pred_t, target_r, target_t, target, model_points, cls_id):
    """tools/trainer.py:370-381 for one frame."""
    pred_points = np.asarray(model_points, np.float32) @ np.asarray(pred_r, np.float32).T \
        + np.asarray(pred_t, np.float32).reshape(1, 3)
    add_dis, _ = metric.cal_adds_cuda(pred_points, target, int(cls_id))
    r_dis = metric.angular_distance(pred_r, target_r)
    t_dis = metric.translation_distance(np.asarray(pred_t).reshape(3), np.asarray(target_t).reshape(3))
    return add_dis, r_dis, t_dis


def cal_adds_torch_nxn(pred, target):
    """lib/utils/metric.py:24-32 in the reference's own form, on the host: the two [N,N,3] tensors made with
    ``repeat``, ``norm(dim=2)``, ``min(dim=1)``, ``mean`` (the reference moves them to the GPU first; the CPU-baseline
    leg of the benchmarks keeps them on the host cores).  pred, target: torch [N,3] float32."""
    import torch
    N = pred.size(0)
    pd = pred.view(1, N, 3).repeat(N, 1, 1)                  # :27
    gt = target.view(N, 1, 3).repeat(1, N, 1)                # :28
    return torch.mean(torch.min(torch.norm(pd - gt, dim=2), dim=1)[0]).item()      # :30-31


# ----------------------------------------------------------------------------- KNN / loss
def knn(x, y):
    """Stand-in for the KeOps handle of train.py:126: argmin_j SqDist(x_i, y_j), K=1 -> [N,1] int64."""
    x = np.asarray(x, np.float32)
    y = np.asarray(y, np.float32)
    out = np.empty((x.shape[0], 1), dtype=np.int64)
    for s in range(0, x.shape[0], 1024):
        d = x[s:s + 1024, None, :] - y[None, :, :]
        d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
        out[s:s + 1024, 0] = d2.argmin(axis=1)
    return out


def pose_loss(sym_list, pred_r, pred_t, targets, model_points, idxs):
    """lib/network/loss.py:25-41 forward value (FP32)."""
    pred_r = np.asarray(pred_r, np.float32)
    model_points = np.asarray(model_points, np.float32)
    targets = np.asarray(targets, np.float32)
    pred_points = model_points @ pred_r.transpose(0, 2, 1) + np.asarray(pred_t, np.float32).reshape(-1, 1, 3)
    per = []
    for b in range(pred_points.shape[0]):
        tgt = targets[b]
        if int(np.asarray(idxs).reshape(-1)[b]) in sym_list:
            tgt = tgt[knn(pred_points[b], tgt).reshape(-1)]
        per.append(np.linalg.norm(pred_points[b] - tgt, axis=1).astype(np.float32).mean(dtype=np.float32))
    return float(np.mean(np.asarray(per, np.float32), dtype=np.float32))


# ----------------------------------------------------------------------------- misc
def vote_translation(p_emb, t_res):
    """lib/network/krrn.py:153 -- p_emb [bs,NP,3], t_res [bs,3,NP] -> [bs,3]."""
    return (np.asarray(p_emb, np.float32) + np.asarray(t_res, np.float32).transpose(0, 2, 1)).mean(axis=1)


def choose_ransac_batch(coor_emb, uv_emb, K_matrix, n=125, rng=None):
    """lib/utils/utlis.py:55-74 (np.random global state replaced by ``rng``; pass the ``np.random`` module
    itself to replay the reference's draws).  One deliberate difference: utlis.py:70 pads BOTH axes of the
    [M,1] inlier array, so the reference returns [1 + n - M, n] identical rows per element when M <= n (and
    np.stack then fails unless all elements share M); here that branch returns the one intended row."""
    rng = rng or np.random.default_rng()
    outs = []
    for j in range(coor_emb.shape[0]):
        ok, _, _, inl = solve_pnp_ransac(coor_emb[j], uv_emb[j], K_matrix[j])
        if inl.shape[0] > n:
            keep = np.zeros(inl.shape[0], dtype=int)
            keep[:n] = 1
            rng.shuffle(keep)
            inl = inl[keep.nonzero()]
        else:
            inl = np.pad(inl, ((0, n - inl.shape[0]), (0, 0)), 'wrap')
        outs.append(inl.T)
    return np.stack(outs, 0)


def eval_frame(metric, frame, diameters, rng, num_points=256):
    """One iteration of the reference's test loop (tools/trainer.py:166-191): get_pose -> cal_dis
    -> success tests.  ``frame`` is a dict from pose_estimation_b200.synth.  Returns a dict."""
    R, t, ok, _ = get_pose(frame['xyz'], frame['extent'], frame['lfborder'], frame['choose'],
                           frame['x_map_choosed'], frame['y_map_choosed'], frame['intrinsic'],
                           num_points=num_points, rng=rng)
    add, r_dis, t_dis = cal_dis(metric, R, t, frame['target_r'], frame['target_t'], frame['target'],
                                frame['model_points'], frame['cls_id'])
    return dict(R=R, t=t, ok=ok, add=add, r_dis=r_dis, t_dis=t_dis,
                succ_rt=add < 0.1 * diameters[int(frame['cls_id'])],   # trainer.py:186
                succ_r=r_dis < 5.0, succ_t=t_dis < 0.05)               # :156-157,189-191
