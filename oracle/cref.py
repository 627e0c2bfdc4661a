"""oracle/cref.py -- ctypes binding of the plain-C algorithm oracle (oracle/pose_ref.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
leg; never by pose_estimation_b200/.  See oracle/pose_ref.h for what each function pins.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpose_ref.so")


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in ("pose_ref.c", "sim3_ref.c", "pose_ref.h")]
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "CC=gcc"])
    return _SO


class RansacParams(C.Structure):
    _fields_ = [("n_hyp", C.c_int32), ("thresh_px", C.c_float), ("gn_iters", C.c_int32),
                ("lo_rounds", C.c_int32), ("min_inliers", C.c_int32),
                ("seed_lo", C.c_uint32), ("seed_hi", C.c_uint32)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.ref_ransac_frame.restype = C.c_int
        _lib.ref_compact_map.restype = C.c_int
        _lib.ref_p3p.restype = C.c_int
        _lib.ref_sample4.restype = C.c_int
        _lib.ref_score.restype = C.c_int
        _lib.ref_hypothesis.restype = C.c_int
        _lib.ref_angular_deg.restype = C.c_float
        _lib.ref_sim3_ransac.restype = C.c_int
        _lib.ref_umeyama_fit.restype = C.c_int
        _lib.ref_sim3_inlier_t.restype = C.c_double
        _lib.ref_umeyama_core.restype = C.c_int
    return _lib


def _p(a, t=None):
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def philox(ctr, key):
    out = np.zeros(4, np.uint32)
    lib().ref_philox4x32(_p(np.asarray(ctr, np.uint32)), _p(np.asarray(key, np.uint32)), _p(out))
    return out


def sample4(seed, frame, hyp, P):
    idx = np.zeros(4, np.int32)
    ok = lib().ref_sample4(C.c_uint32(seed & 0xFFFFFFFF), C.c_uint32(seed >> 32), C.c_uint32(frame),
                           C.c_uint32(hyp), C.c_int(P), _p(idx))
    return bool(ok), idx


def p3p(x, y):
    x = _f32(x).reshape(3, 3)
    y = _f32(y).reshape(3, 3)
    R = np.zeros((4, 9), np.float32)
    t = np.zeros((4, 3), np.float32)
    n = lib().ref_p3p(_p(x), _p(y), _p(R), _p(t))
    return R[:n].reshape(n, 3, 3), t[:n]


def make_params(n_hyp=256, thresh_px=1.0, gn_iters=6, lo_rounds=0, min_inliers=5, seed=0x5EED):
    return RansacParams(n_hyp, thresh_px, gn_iters, lo_rounds, min_inliers,
                        seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)


def ransac_frame(pts, K4, frame=0, sample_idx=None, **kw):
    """pts: [5,P] f32 SoA (X,Y,Z,U,V).  Returns dict."""
    pts = _f32(pts)
    P = pts.shape[1]
    prm = make_params(**kw)
    H = prm.n_hyp
    R = np.zeros(9, np.float32); t = np.zeros(3, np.float32)
    R0 = np.zeros(9, np.float32); t0 = np.zeros(3, np.float32)
    n_inl = C.c_int32(); best = C.c_int32(); n_ref = C.c_int32()
    counts = np.zeros(H, np.int32)
    bitmap = np.zeros(max((P + 31) // 32, 1), np.uint32)
    K4 = _f32(K4)
    if sample_idx is not None:
        sample_idx = np.ascontiguousarray(sample_idx, np.int32)
    st = lib().ref_ransac_frame(_p(pts[0]), _p(pts[1]), _p(pts[2]), _p(pts[3]), _p(pts[4]), C.c_int(P),
                                _p(K4), C.byref(prm), C.c_uint32(frame), _p(sample_idx),
                                _p(R), _p(t), C.byref(n_inl), C.byref(best), C.byref(n_ref),
                                _p(counts), _p(bitmap), _p(R0), _p(t0))
    return dict(status=st, R=R.reshape(3, 3), t=t, n_inl=n_inl.value, best_hyp=best.value,
                n_inl_refined=n_ref.value, hyp_counts=counts, bitmap=bitmap,
                R0=R0.reshape(3, 3), t0=t0)


def compact_map(xyz, mask, extent, lfborder, rmin=0, cmin=0, cap=None):
    xyz = _f32(xyz)
    _, Hh, Ww = xyz.shape
    cap = cap or Hh * Ww
    out = np.zeros((5, cap), np.float32)
    total = C.c_int32()
    if mask is not None:
        mask = np.ascontiguousarray(mask, np.uint8)
    n = lib().ref_compact_map(_p(xyz), _p(mask), C.c_int(Hh), C.c_int(Ww), _p(_f32(extent)),
                              _p(_f32(lfborder)), C.c_int(rmin), C.c_int(cmin), C.c_int(cap),
                              _p(out[0]), _p(out[1]), _p(out[2]), _p(out[3]), _p(out[4]), C.byref(total))
    return out[:, :n].copy(), total.value


def knn1(x, y):
    x = _f32(x); y = _f32(y)
    idx = np.zeros(x.shape[0], np.int64)
    d2 = np.zeros(x.shape[0], np.float32)
    lib().ref_knn1(_p(x), _p(y), C.c_int(x.shape[0]), C.c_int(y.shape[0]), _p(idx), _p(d2))
    return idx, d2


def add_adds(pred, tgt, want_adds=True):
    pred = _f32(pred); tgt = _f32(tgt)
    add = C.c_float(); adds = C.c_float()
    lib().ref_add_adds(_p(pred), _p(tgt), C.c_int(pred.shape[0]), C.c_int(tgt.shape[0]),
                       C.byref(add), C.byref(adds) if want_adds else None)
    return add.value, (adds.value if want_adds else add.value)


def transform(model, R, t):
    model = _f32(model)
    out = np.zeros_like(model)
    lib().ref_transform(_p(model), C.c_int(model.shape[0]), _p(_f32(R).reshape(9)), _p(_f32(t)), _p(out))
    return out


def angular_deg(R1, R2):
    return float(lib().ref_angular_deg(_p(_f32(R1).reshape(9)), _p(_f32(R2).reshape(9))))


def vote_translation(cloud, t_res):
    cloud = _f32(cloud); t_res = _f32(t_res)
    t = np.zeros(3, np.float32)
    lib().ref_vote_translation(_p(cloud), _p(t_res), C.c_int(cloud.shape[0]), _p(t))
    return t


def resample_inliers(bitmap, P, n, seed=0x5EED, frame=0):
    bitmap = np.ascontiguousarray(bitmap, np.uint32)
    out = np.zeros(n, np.int32)
    lib().ref_resample_inliers(_p(bitmap), C.c_int(P), C.c_int(n), C.c_uint32(seed & 0xFFFFFFFF),
                               C.c_uint32(seed >> 32), C.c_uint32(frame), _p(out))
    return out


def sim3_ransac(src, dst, n_iter=128, confidence=0.99, min_ratio=0.1, seed=0x5EED, prob=0, sample_idx=None):
    """3D-3D RANSAC + Umeyama of ONE problem (sim3_ref.c) -> dict(status, scale, R, t, n_inl, best_hyp,
    hyp_counts [n_iter], bitmap)."""
    src = _f32(src); dst = _f32(dst)
    n = src.shape[0]
    si = None if sample_idx is None else np.ascontiguousarray(sample_idx, dtype=np.int32).reshape(n_iter, 5)
    scale = C.c_float(); n_inl = C.c_int32(); best = C.c_int32()
    R = np.zeros(9, np.float32); t = np.zeros(3, np.float32)
    counts = np.zeros(n_iter, np.int32)
    bitmap = np.zeros(max((n + 31) // 32, 1), np.uint32)
    st = lib().ref_sim3_ransac(_p(src), _p(dst), C.c_int(n), C.c_int(n_iter), C.c_float(confidence), C.c_float(min_ratio),
                               C.c_uint32(seed & 0xFFFFFFFF), C.c_uint32(seed >> 32), C.c_uint32(prob), _p(si),
                               C.byref(scale), _p(R), _p(t), C.byref(n_inl), C.byref(best), _p(counts), _p(bitmap))
    return dict(status=int(st), scale=float(scale.value), R=R.reshape(3, 3), t=t, n_inl=int(n_inl.value),
                best_hyp=int(best.value), hyp_counts=counts, bitmap=bitmap)


def sim3_sample5(seed, prob, hyp, n):
    idx = np.zeros(5, np.int32)
    lib().ref_sim3_sample5(C.c_uint32(seed & 0xFFFFFFFF), C.c_uint32(seed >> 32), C.c_uint32(prob), C.c_uint32(hyp),
                           C.c_int(n), _p(idx))
    return idx


def umeyama_fit(src, dst):
    src = _f32(src); dst = _f32(dst)
    scale = C.c_float(); R = np.zeros(9, np.float32); t = np.zeros(3, np.float32)
    st = lib().ref_umeyama_fit(_p(src), _p(dst), C.c_int(src.shape[0]), C.byref(scale), _p(R), _p(t))
    return int(st), float(scale.value), R.reshape(3, 3), t


def fps(points, n_samples):
    points = _f32(points)
    out = np.zeros(n_samples, np.int32)
    lib().ref_fps(_p(points), C.c_int(points.shape[0]), C.c_int(n_samples), _p(out))
    return out
