"""3D-3D RANSAC + Umeyama similarity fit on the GPU (csrc/sim3.cu) -- the batched form of
lib/transform/umeyama.py:8-97 (estimateSimilarityTransform / estimateSimilarityUmeyama)."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from . import _lib
from ._lib import call
from .ops import _chk, _ptr, _stream


class Sim3Params(C.Structure):
    _fields_ = [("n_iter", C.c_int32), ("confidence", C.c_float), ("min_inlier_ratio", C.c_float),
                ("seed_lo", C.c_uint32), ("seed_hi", C.c_uint32)]


@dataclass
class Sim3Config:
    """Defaults are the reference's constants: maxIter = 128, confidence = 0.99 (umeyama.py:21-22), and the
    0.1 inlier-ratio floor below which it returns None (:50)."""
    n_iter: int = 128
    confidence: float = 0.99
    min_inlier_ratio: float = 0.1
    seed: int = 0x5EED

    def c_struct(self) -> Sim3Params:
        return Sim3Params(self.n_iter, self.confidence, self.min_inlier_ratio, self.seed & 0xFFFFFFFF,
                          (self.seed >> 32) & 0xFFFFFFFF)


@dataclass
class Sim3Result:
    scale: torch.Tensor        # [B] f32
    R: torch.Tensor            # [B,3,3] f32
    t: torch.Tensor            # [B,3] f32
    n_inl: torch.Tensor        # [B] i32 inliers of the chosen iteration (the refit set)
    best_hyp: torch.Tensor     # [B] i32 chosen iteration (-1: none)
    status: torch.Tensor       # [B] i32 PB_STATUS_*
    hyp_counts: torch.Tensor | None = None   # [B,n_iter] i32
    inl_bitmap: torch.Tensor | None = None   # [B,ceil(cap/32)] u32 (as int32 storage)


def sim3_ransac(src, dst, count=None, cfg: Sim3Config | None = None, sample_idx=None, prob0=0, want_counts=False,
                want_bitmap=False) -> Sim3Result:
    """src, dst [B,cap,3] f32 CUDA (dst ~ scale * R * src + t), count [B] i32 valid pairs (None: all cap).
    sample_idx [B,n_iter,5] i32 overrides the counter RNG (e.g. to replay np.random's draws)."""
    cfg = cfg or Sim3Config()
    src = _chk(src, torch.float32, "src"); dst = _chk(dst, torch.float32, "dst")
    if src.dim() != 3 or src.shape[-1] != 3 or src.shape != dst.shape:
        raise ValueError("src and dst must both be [B,cap,3]")
    B, cap, _ = src.shape
    dev = src.device
    if count is None:
        count = torch.full((B,), cap, dtype=torch.int32, device=dev)
    count = _chk(count, torch.int32, "count")
    sample_idx = _chk(sample_idx, torch.int32, "sample_idx", True)
    if sample_idx is not None and tuple(sample_idx.shape) != (B, cfg.n_iter, 5):
        raise ValueError("sample_idx must be [B,n_iter,5]")
    prm = cfg.c_struct()
    nb = _lib.load().pb_sim3_ransac_workspace_bytes(B, C.byref(prm))
    ws = torch.empty(nb + 16, dtype=torch.uint8, device=dev)
    off = (-ws.data_ptr()) % 16
    res = Sim3Result(scale=torch.empty((B,), dtype=torch.float32, device=dev),
                     R=torch.empty((B, 3, 3), dtype=torch.float32, device=dev),
                     t=torch.empty((B, 3), dtype=torch.float32, device=dev),
                     n_inl=torch.empty((B,), dtype=torch.int32, device=dev),
                     best_hyp=torch.empty((B,), dtype=torch.int32, device=dev),
                     status=torch.empty((B,), dtype=torch.int32, device=dev))
    if want_counts:
        res.hyp_counts = torch.empty((B, cfg.n_iter), dtype=torch.int32, device=dev)
    if want_bitmap:
        res.inl_bitmap = torch.zeros((B, (cap + 31) // 32), dtype=torch.int32, device=dev)
    call("pb_sim3_ransac", _ptr(src), _ptr(dst), _ptr(count), B, cap, int(prob0), C.byref(prm), _ptr(sample_idx),
         _ptr(res.scale), _ptr(res.R), _ptr(res.t), _ptr(res.n_inl), _ptr(res.best_hyp), _ptr(res.status),
         _ptr(res.hyp_counts), _ptr(res.inl_bitmap), C.c_void_p(ws.data_ptr() + off), nb, _stream(res.R))
    return res


def umeyama_fit(src, dst, count=None):
    """Batched estimateSimilarityUmeyama (umeyama.py:65-97) over all pairs -> (scale [B], R [B,3,3], t [B,3], status [B])."""
    src = _chk(src, torch.float32, "src"); dst = _chk(dst, torch.float32, "dst")
    if src.dim() != 3 or src.shape[-1] != 3 or src.shape != dst.shape:
        raise ValueError("src and dst must both be [B,cap,3]")
    B, cap, _ = src.shape
    dev = src.device
    if count is None:
        count = torch.full((B,), cap, dtype=torch.int32, device=dev)
    count = _chk(count, torch.int32, "count")
    scale = torch.empty((B,), dtype=torch.float32, device=dev)
    R = torch.empty((B, 3, 3), dtype=torch.float32, device=dev)
    t = torch.empty((B, 3), dtype=torch.float32, device=dev)
    status = torch.empty((B,), dtype=torch.int32, device=dev)
    call("pb_umeyama_fit", _ptr(src), _ptr(dst), _ptr(count), B, cap, _ptr(scale), _ptr(R), _ptr(t), _ptr(status), _stream(R))
    return scale, R, t, status


def farthest_point_sampling(points, n_samples, count=None):
    """points [B,N,3] or [N,3] f32 CUDA -> int32 indices [B,n_samples] (or [n_samples]); same selection as
    tools/script/sample_model.py:33-46 (start at index 0, float32 distances, first maximum on ties)."""
    points = _chk(points, torch.float32, "points")
    squeeze = points.dim() == 2
    if squeeze:
        points = points.unsqueeze(0)
    if points.dim() != 3 or points.shape[-1] != 3:
        raise ValueError("points must be [B,N,3] or [N,3]")
    B, cap, _ = points.shape
    count = _chk(count, torch.int32, "count", True)
    out = torch.empty((B, n_samples), dtype=torch.int32, device=points.device)
    nb = _lib.load().pb_fps_workspace_bytes(B, cap)
    ws = torch.empty(nb, dtype=torch.uint8, device=points.device) if nb else None
    call("pb_fps", _ptr(points), _ptr(count), B, cap, int(n_samples), _ptr(out), _ptr(ws), nb, _stream(out))
    return out[0] if squeeze else out
