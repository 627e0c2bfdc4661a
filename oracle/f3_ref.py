"""oracle/f3_ref.py -- CPU restatements for the SURVEY 8(f) row f3 remainder (TEST INFRASTRUCTURE ONLY).

  kabsch                 version/transparent/lib/declarative/leastsquares.py:10-48 (best_batch_fit_transform)
  icp_refine             version/transparent/scripts/train_declarative.py:80-100 (transform -> knn -> gather -> fit -> compose)
  trans_ransac           lib/transform/trans.py:9-123 (the older RANSAC rule + its Umeyama, quirks included)
  backproject / uvd2cloud  dataset/linemod/batchdataset.py:712-721, lib/transform/coordinate.py:91-101

Pinned: tests/golden/gen_golden_f3.py imports leastsquares.py, trans.py and coordinate.py out of /root/reference,
runs them (np.random.randint recorded for trans.py) and stores inputs + outputs in tests/golden/f3_ref.npz;
tests/test_oracle_golden.py checks every function here against those runs.  Unpinned stand-in: the KeOps arg-min
handle inside the ICP loop (float32 brute force, lowest index on ties), as everywhere else.
Only tests/ may import this module.
"""
from __future__ import annotations

import numpy as np


def kabsch(A, B):
    """leastsquares.py:10-48 for one problem: A, B [N,3] -> (R [3,3], t [3], centroid_A, centroid_B), float64."""
    A = np.asarray(A, np.float64)
    B = np.asarray(B, np.float64)
    cA, cB = A.mean(axis=0), B.mean(axis=0)                   # :27-28
    H = (A - cA).T @ (B - cB)                                  # :33
    U, S, Vt = np.linalg.svd(H)                                # :34
    R = Vt.T @ U.T                                             # :36
    if np.linalg.det(R) < 0:                                   # :39-43 reflection case
        Vt[2, :] *= -1
        R = Vt.T @ U.T
    t = cB - R @ cA                                            # :45
    return R, t, cA, cB


def knn_f32(x, y):
    """argmin_j SqDist(x_i, y_j) in float32 with the kernels' operation order, lowest j on ties (KeOps stand-in)."""
    x = np.asarray(x, np.float32)
    y = np.asarray(y, np.float32)
    out = np.empty(x.shape[0], np.int64)
    for s in range(0, x.shape[0], 2048):
        d = x[s:s + 2048, None, :] - y[None, :, :]
        d2 = d[..., 0] * d[..., 0]
        d2 = d[..., 1] * d[..., 1] + d2
        d2 = d[..., 2] * d[..., 2] + d2
        out[s:s + 2048] = d2.argmin(axis=1)
    return out


def icp_refine(model, target, R, t, steps=10, compose="reference"):
    """train_declarative.py:80-100 for one hypothesis: model [N,3], target [M,3], R [3,3], t [3]."""
    model = np.asarray(model, np.float32)
    target = np.asarray(target, np.float32)
    R = np.asarray(R, np.float32).copy()
    t = np.asarray(t, np.float32).copy()
    for _ in range(steps):
        pred = (model @ R.T + t).astype(np.float32)                                  # :81
        idx = knn_f32(pred, target)                                                  # :82
        oR, ot, _, _ = kabsch(pred, target[idx])                                     # :84-86
        if compose == "reference":
            R, t = (R.astype(np.float64) @ oR).astype(np.float32), (t + ot).astype(np.float32)      # :98-99 as written
        else:
            R, t = (oR @ R.astype(np.float64)).astype(np.float32), (oR @ t + ot).astype(np.float32)
    return R, t


# ---- lib/transform/trans.py ---------------------------------------------------------------------------------
def trans_umeyama(SourceHom, TargetHom):
    """trans.py:93-123 (Rotation is the TRANSPOSE of U Vh and goes into OutTransform as it is)."""
    ms, mt = SourceHom[:3].mean(axis=1), TargetHom[:3].mean(axis=1)
    n = SourceHom.shape[1]
    cs, ct = SourceHom[:3] - ms[:, None], TargetHom[:3] - mt[:, None]
    C = ct @ cs.T / n
    U, D, Vh = np.linalg.svd(C, full_matrices=True)
    if np.linalg.det(U) * np.linalg.det(Vh) < 0.0:
        D[-1] = -D[-1]
        U[:, -1] = -U[:, -1]
    Rotation = (U @ Vh).T
    varP = np.var(SourceHom[:3], axis=1).sum()
    s = 1 / varP * np.sum(D)
    Translation = mt - ms.dot(s * Rotation)
    M = np.identity(4)
    M[:3, :3] = np.diag([s, s, s]) @ Rotation
    M[:3, 3] = Translation
    return np.array([s, s, s]), Rotation, Translation, M


def trans_ransac(source, target, rand_idx):
    """trans.py:9-90 with the 5-index draws given ([iterations,5]).  Returns dict: pass_t, residuals, n_inl
    (the reference's nInliers, index 0 never counted), best (iteration), ok, Scales, Rotation, Translation."""
    source = np.asarray(source, np.float64)
    target = np.asarray(target, np.float64)
    n = source.shape[0]
    S = np.vstack([source.T, np.ones(n)])
    T = np.vstack([target.T, np.ones(n)])
    tn, sn = np.mean(np.linalg.norm(target, axis=1)), np.mean(np.linalg.norm(source, axis=1))
    pass_t = max(tn / sn, sn / tn)
    stop_t = pass_t / 100
    best_res, best_ratio, best, best_idx = 1e10, 0.0, -1, np.arange(n)
    residuals, counts = [], []
    for i, ri in enumerate(np.asarray(rand_idx)):
        _, _, _, M = trans_umeyama(S[:, ri], T[:, ri])
        rv = np.linalg.norm((T - M @ S)[:3], axis=0)
        res = np.linalg.norm(rv)
        inl = np.where(rv < pass_t)
        cnt = np.count_nonzero(inl)                              # :88 counts the non-zero INDICES
        residuals.append(res); counts.append(cnt)
        if res < best_res:
            best_res, best_ratio, best, best_idx = res, cnt / n, i, inl[0]
        if best_res < stop_t:
            break
    out = dict(pass_t=pass_t, residuals=np.asarray(residuals), n_inl=np.asarray(counts), best=best,
               ok=best_ratio >= 0.1, n_refit=int(best_idx.shape[0]))
    if out["ok"]:
        out["Scales"], out["Rotation"], out["Translation"], out["OutTransform"] = trans_umeyama(S[:, best_idx], T[:, best_idx])
    return out


# ---- back-projection -----------------------------------------------------------------------------------------
def backproject(depth, choose, xmap, ymap, K4, depth_scale=1.0):
    """batchdataset.py:712-721 for one frame: depth [H,W] f32, choose [NP], x / y lists [NP] f32 -> cloud [NP,3] f32."""
    d = (np.asarray(depth, np.float32).reshape(-1)[np.asarray(choose)] * np.float32(depth_scale)).astype(np.float32)
    fx, fy, cx, cy = [np.float32(v) for v in K4]
    pt0 = (np.asarray(xmap, np.float32) - cx) * d / fx
    pt1 = (np.asarray(ymap, np.float32) - cy) * d / fy
    return np.stack([pt0, pt1, d], axis=1).astype(np.float32)


def uvd2cloud(uvd, K):
    """coordinate.py:91-101 with the docstring's output shape [bs,n,3]."""
    uvd = np.asarray(uvd, np.float32)
    K = np.asarray(K, np.float32)
    x = (uvd[..., 0] - K[0, 2]) * uvd[..., 2] / K[0, 0]
    y = (uvd[..., 1] - K[1, 2]) * uvd[..., 2] / K[1, 1]
    return np.stack([x, y, uvd[..., 2]], axis=-1).astype(np.float32)
