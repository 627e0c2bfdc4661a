"""Back-propagatable PnP on the GPU (csrc/bpnp.cu) -- the batched form of lib/network/dnn/BPnP.py:
``pnp_refine`` = the cv2.solvePnP(ITERATIVE, useExtrinsicGuess) step of BPnP.forward (:43-44),
``bpnp_backward`` = BPnP.backward (:54-117)."""
from __future__ import annotations

import torch

from . import _lib, ops
from ._lib import call
from .ops import RansacConfig, _chk, _ptr, _stream

STATUS_CONVERGED, STATUS_ITER_LIMIT, STATUS_DEGENERATE = 0, 1, 2


def _shapes(pts2d, pts3d):
    if pts2d.dim() != 3 or pts2d.shape[-1] != 2:
        raise ValueError("pts2d must be [B,n,2]")
    B, n, _ = pts2d.shape
    if tuple(pts3d.shape) == (n, 3):
        batched = 0
    elif tuple(pts3d.shape) == (B, n, 3):
        batched = 1
    else:
        raise ValueError("pts3d must be [n,3] or [B,n,3] with n matching pts2d")
    return B, n, batched


def pnp_refine(pts2d, pts3d, K, pose6_init=None, R_init=None, t_init=None, max_iter=100, want_rmse=False):
    """pts2d [B,n,2], pts3d [n,3] | [B,n,3], K [3,3] f32 CUDA; start from pose6_init [B,6] (angle-axis, t) or from
    R_init [B,3,3] + t_init [B,3].  Returns (pose6 [B,6] f32, status [B] i32[, rmse [B] f32])."""
    pts2d = _chk(pts2d, torch.float32, "pts2d"); pts3d = _chk(pts3d, torch.float32, "pts3d")
    K = _chk(K, torch.float32, "K")
    pose6_init = _chk(pose6_init, torch.float32, "pose6_init", True)
    R_init = _chk(R_init, torch.float32, "R_init", True); t_init = _chk(t_init, torch.float32, "t_init", True)
    B, n, batched = _shapes(pts2d, pts3d)
    if K.numel() != 9:
        raise ValueError("K must be [3,3]")
    if pose6_init is not None and tuple(pose6_init.shape) != (B, 6):
        raise ValueError("pose6_init must be [B,6]")
    if pose6_init is None and (R_init is None or t_init is None or R_init.numel() != 9 * B or t_init.numel() != 3 * B):
        raise ValueError("need pose6_init [B,6] or R_init [B,3,3] + t_init [B,3]")
    dev = pts2d.device
    pose = torch.empty((B, 6), dtype=torch.float32, device=dev)
    status = torch.empty((B,), dtype=torch.int32, device=dev)
    rmse = torch.empty((B,), dtype=torch.float32, device=dev) if want_rmse else None
    call("pb_pnp_refine", _ptr(pts2d), _ptr(pts3d), batched, _ptr(K), _ptr(pose6_init), _ptr(R_init), _ptr(t_init), B, n,
         int(max_iter), _ptr(pose), _ptr(rmse), _ptr(status), _stream(pose))
    return (pose, status, rmse) if want_rmse else (pose, status)


def bpnp_backward(pts2d, pts3d, K, pose6, grad_out):
    """Returns (grad_pts2d [B,n,2], grad_pts3d (shape of pts3d), grad_K [3,3], status [B] i32)."""
    pts2d = _chk(pts2d, torch.float32, "pts2d"); pts3d = _chk(pts3d, torch.float32, "pts3d")
    K = _chk(K, torch.float32, "K"); pose6 = _chk(pose6, torch.float32, "pose6")
    grad_out = _chk(grad_out, torch.float32, "grad_out")
    B, n, batched = _shapes(pts2d, pts3d)
    if tuple(pose6.shape) != (B, 6) or tuple(grad_out.shape) != (B, 6) or K.numel() != 9:
        raise ValueError("pose6 and grad_out must be [B,6], K [3,3]")
    dev = pts2d.device
    gx = torch.empty_like(pts2d)
    gz = torch.empty_like(pts3d)
    gK = torch.empty((3, 3), dtype=torch.float32, device=dev)
    status = torch.empty((B,), dtype=torch.int32, device=dev)
    nb = _lib.load().pb_bpnp_backward_workspace_bytes(B, n)
    ws = torch.empty(((nb + 7) // 8,), dtype=torch.float64, device=dev)
    call("pb_bpnp_backward", _ptr(pts2d), _ptr(pts3d), batched, _ptr(K), _ptr(pose6), _ptr(grad_out), B, n, _ptr(gx),
         _ptr(gz), _ptr(gK), _ptr(status), _ptr(ws), ws.numel() * 8, _stream(gx))
    return gx, gz, gK, status


def bpnp_forward(pts2d, pts3d, K, ini_pose=None, cfg: RansacConfig | None = None, max_iter=100):
    """BPnP.forward (BPnP.py:24-51): initial pose from RANSAC (3 px, the reference's reprojectionError) unless
    ini_pose [B,6] is given, then the all-point refinement.  Returns (P_6d [B,6], status [B])."""
    if ini_pose is not None:
        return pnp_refine(pts2d, pts3d, K, pose6_init=ini_pose, max_iter=max_iter)
    B, n, batched = _shapes(pts2d, pts3d)
    cfg = cfg or RansacConfig(thresh_px=3.0)
    z = pts3d if batched else pts3d.unsqueeze(0).expand(B, n, 3)
    pts = torch.cat([z.permute(0, 2, 1), pts2d.permute(0, 2, 1)], dim=1).contiguous().float()      # [B,5,n]
    count = torch.full((B,), n, dtype=torch.int32, device=pts.device)
    Kf = K.reshape(9).float()
    K4 = torch.stack([Kf[0], Kf[4], Kf[2], Kf[5]]).unsqueeze(0).expand(B, 4).contiguous()
    out = ops.pnp_ransac_pts(pts, count, K4, cfg)
    return pnp_refine(pts2d, pts3d, K, R_init=out["R"], t_init=out["t"], max_iter=max_iter)
