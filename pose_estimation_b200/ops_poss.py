"""Per-hypothesis pose loss on the GPU (csrc/possloss.cu) -- version/transparent/lib/networks/loss.py:197-267."""
from __future__ import annotations

import torch

from . import _lib
from ._lib import call
from .ops import _chk, _ptr, _stream


def _prep(pred_r, pred_t, pred_c, target, model_points, axis, target_r):
    pred_r = _chk(pred_r, torch.float32, "pred_r"); pred_t = _chk(pred_t, torch.float32, "pred_t")
    pred_c = _chk(pred_c, torch.float32, "pred_c"); target = _chk(target, torch.float32, "target")
    model_points = _chk(model_points, torch.float32, "model_points")
    axis = _chk(axis, torch.float32, "axis"); target_r = _chk(target_r, torch.float32, "target_r")
    if pred_r.dim() != 3 or pred_r.shape[-1] != 4:
        raise ValueError("pred_r must be [bs,num_p,4]")
    bs, num_p, _ = pred_r.shape
    if tuple(pred_t.shape) != (bs, num_p, 3) or pred_c.numel() != bs * num_p:
        raise ValueError("pred_t must be [bs,num_p,3] and pred_c [bs,num_p,1]")
    if model_points.dim() != 3 or model_points.shape[0] != bs or model_points.shape[-1] != 3 or target.shape != model_points.shape:
        raise ValueError("model_points and target must both be [bs,N,3]")
    if axis.numel() != 3 or target_r.numel() != 9:
        raise ValueError("axis must hold 3 and target_r 9 values")
    return pred_r, pred_t, pred_c, target, model_points, axis, target_r, bs, num_p, model_points.shape[1]


def poss_loss_fwd(pred_r, pred_t, pred_c, target, model_points, nn_mode, w, axis, target_r):
    """Returns (vals [3] f32 device = loss, dis[0][argmax c[0]], loss_rotation; ws = the buffer the backward needs)."""
    pred_r, pred_t, pred_c, target, model_points, axis, target_r, bs, num_p, N = _prep(pred_r, pred_t, pred_c, target,
                                                                                     model_points, axis, target_r)
    dev = pred_r.device
    nb = _lib.load().pb_poss_loss_workspace_bytes(bs, num_p, N)
    ws = torch.empty(((nb + 15) // 16 * 4,), dtype=torch.int32, device=dev)
    vals = torch.empty((3,), dtype=torch.float32, device=dev)
    call("pb_poss_loss_fwd", _ptr(pred_r), _ptr(pred_t), _ptr(pred_c), _ptr(model_points), _ptr(target), bs, num_p, N,
         int(bool(nn_mode)), float(w), _ptr(axis), _ptr(target_r), _ptr(vals), _ptr(ws), ws.numel() * 4, _stream(vals))
    return vals, ws


def poss_loss_bwd(pred_r, pred_t, pred_c, target, model_points, nn_mode, w, axis, target_r, grad_vals, ws):
    """grad_vals [3] f32 device (d/d loss, d/d dis_sel, d/d loss_rotation) -> grads shaped like pred_r, pred_t, pred_c."""
    pred_r, pred_t, pred_c, target, model_points, axis, target_r, bs, num_p, N = _prep(pred_r, pred_t, pred_c, target,
                                                                                     model_points, axis, target_r)
    grad_vals = _chk(grad_vals, torch.float32, "grad_vals")
    gr, gt, gc = torch.empty_like(pred_r), torch.empty_like(pred_t), torch.empty_like(pred_c)
    call("pb_poss_loss_bwd", _ptr(pred_r), _ptr(pred_t), _ptr(pred_c), _ptr(model_points), _ptr(target), bs, num_p, N,
         int(bool(nn_mode)), float(w), _ptr(axis), _ptr(target_r), _ptr(grad_vals), _ptr(ws), ws.numel() * 4, _ptr(gr), _ptr(gt),
         _ptr(gc), _stream(gc))
    return gr, gt, gc
