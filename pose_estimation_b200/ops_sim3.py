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
                ("seed_lo", C.c_uint32), ("seed_hi", C.c_uint32), ("rule", C.c_int32)]


@dataclass
class Sim3Config:
    """Defaults are the reference's constants: maxIter = 128, confidence = 0.99 (umeyama.py:21-22), and the
    0.1 inlier-ratio floor below which it returns None (:50)."""
    n_iter: int = 128
    confidence: float = 0.99
    min_inlier_ratio: float = 0.1
    seed: int = 0x5EED
    rule: str = "umeyama"      # "umeyama": lib/transform/umeyama.py; "trans": the older lib/transform/trans.py:47-90
                               # (n_iter = 100 there; smallest total residual wins; confidence unused)

    def c_struct(self) -> Sim3Params:
        if self.rule not in ("umeyama", "trans"):
            raise ValueError("rule must be 'umeyama' or 'trans'")
        return Sim3Params(self.n_iter, self.confidence, self.min_inlier_ratio, self.seed & 0xFFFFFFFF,
                          (self.seed >> 32) & 0xFFFFFFFF, 1 if self.rule == "trans" else 0)


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


def kabsch_fit(A, B, count=None, want_centroids=False):
    """Batched rigid fit B ~ R A + t (version/transparent/lib/declarative/leastsquares.py:10-48).
    A, B [c,N,3] f32 CUDA -> (R [c,3,3], t [c,3], status [c]) (+ centroid_A, centroid_B [c,3])."""
    A = _chk(A, torch.float32, "A"); B = _chk(B, torch.float32, "B")
    if A.dim() != 3 or A.shape[-1] != 3 or A.shape != B.shape:
        raise ValueError("A and B must both be [c,N,3]")
    c, n, _ = A.shape
    dev = A.device
    count = _chk(count, torch.int32, "count", True)
    R = torch.empty((c, 3, 3), dtype=torch.float32, device=dev)
    t = torch.empty((c, 3), dtype=torch.float32, device=dev)
    status = torch.empty((c,), dtype=torch.int32, device=dev)
    cA = torch.empty((c, 3), dtype=torch.float32, device=dev) if want_centroids else None
    cB = torch.empty((c, 3), dtype=torch.float32, device=dev) if want_centroids else None
    call("pb_kabsch_fit", _ptr(A), _ptr(B), _ptr(count), c, n, _ptr(R), _ptr(t), _ptr(cA), _ptr(cB), _ptr(status), _stream(R))
    return (R, t, status, cA, cB) if want_centroids else (R, t, status)


def icp_refine(model, target, R, t, n_steps=10, compose="reference", impl="auto"):
    """The loop of version/transparent/scripts/train_declarative.py:80-100 for B hypotheses at once:
    n_steps x { pred = model R^T + t; nearest target point of every pred point; rigid fit pred -> matches;
    compose }.  model [B,N,3] or [N,3] (shared), target [B,M,3] or [M,3] (shared), R [B,3,3], t [B,3] f32 CUDA.
    compose "reference": R <- R opt_R, t <- t + opt_t (the script's lines 98-99 as written); "rigid": the
    rigid-motion composition R <- opt_R R, t <- opt_R t + opt_t.  Returns new (R, t, status)."""
    if compose not in ("reference", "rigid"):
        raise ValueError("compose must be 'reference' or 'rigid'")
    model = _chk(model, torch.float32, "model"); target = _chk(target, torch.float32, "target")
    R = _chk(R, torch.float32, "R").clone(); t = _chk(t, torch.float32, "t").reshape(-1, 3).clone()
    B = R.shape[0]
    N, M = model.shape[-2], target.shape[-2]
    m_shared, t_shared = model.dim() == 2, target.dim() == 2
    if (not m_shared and model.shape[0] != B) or (not t_shared and target.shape[0] != B) or t.shape[0] != B:
        raise ValueError("batch sizes of model / target / R / t differ")
    dev = R.device
    use_tc = impl == "tc" or (impl == "auto" and M >= 768 and N * M >= (1 << 20))
    nb = _lib.load().pb_icp_refine_workspace_bytes(B, N, M)
    ws = torch.empty(nb + 128, dtype=torch.uint8, device=dev)
    off = (-ws.data_ptr()) % 128
    status = torch.zeros((B,), dtype=torch.int32, device=dev)
    call("pb_icp_refine", _ptr(model), 1 if m_shared else 0, _ptr(target), 1 if t_shared else 0, _ptr(R), _ptr(t), B, N, M,
         int(n_steps), 1 if compose == "rigid" else 0, 1 if use_tc else 0, _ptr(status), C.c_void_p(ws.data_ptr() + off), nb,
         _stream(R))
    return R, t, status


def backproject_cloud(depth, choose, K4, xmap=None, ymap=None, roi=None, depth_scale=1.0):
    """cloud [F,NP,3] from depth [F,H,W] f32 at the choose lists (dataset/linemod/batchdataset.py:712-721):
    ((x - cx) d / fx, (y - cy) d / fy, d), d = depth[choose] * depth_scale.  choose [F,NP] int64 / int32; xmap / ymap
    [F,NP] f32 or None (pixel coordinates from the index and roi [F,2] = (rmin, cmin)); K4 [F,4]."""
    depth = _chk(depth, torch.float32, "depth"); K4 = _chk(K4, torch.float32, "K4")
    if choose.dtype not in (torch.int64, torch.int32) or not choose.is_cuda:
        raise ValueError("choose must be a CUDA int64 / int32 tensor")
    choose = choose.contiguous()
    xmap = _chk(xmap, torch.float32, "xmap", True); ymap = _chk(ymap, torch.float32, "ymap", True)
    roi = _chk(roi, torch.int32, "roi", True)
    F, H, W = depth.shape
    NP = choose.shape[-1]
    cloud = torch.empty((F, NP, 3), dtype=torch.float32, device=depth.device)
    call("pb_backproject_cloud", _ptr(depth), _ptr(choose), 1 if choose.dtype == torch.int32 else 0, _ptr(xmap), _ptr(ymap),
         _ptr(roi), _ptr(K4), float(depth_scale), F, H, W, NP, _ptr(cloud), _stream(cloud))
    return cloud


def uvd2cloud(uvd, K):
    """lib/transform/coordinate.py:91-101: uvd [bs,n,3], K [3,3] -> xyz [bs,n,3] (forward only)."""
    uvd = _chk(uvd, torch.float32, "uvd"); K = _chk(K, torch.float32, "K")
    if uvd.shape[-1] != 3 or K.numel() != 9:
        raise ValueError("uvd must be [...,3] and K [3,3]")
    out = torch.empty_like(uvd)
    n = uvd.shape[-2] if uvd.dim() >= 2 else 1
    B = uvd.numel() // (3 * n)
    call("pb_uvd2cloud", _ptr(uvd), _ptr(K), B, n, _ptr(out), _stream(out))
    return out


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
