"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's per-hypothesis pose loss
(/root/reference/version/transparent/lib/networks/loss.py:197-267, PossLoss) with the helpers it calls
(lib/transform/allocentric.py:15-45 allo_to_ego_mat_torch, :48-79 quat2mat_torch).  Only tests/,
__graft_entry__.smoke() and bench.py's CPU legs may import this file; the product path never does.

torch (CPU) is used for the arithmetic so that ``dtype`` can be float32 (what the reference computes in) or
float64 (the checker), and so that the gradients come from autograd exactly as in the reference.  The KeOps
handle ``knn`` (third-party, absent here) is restated from its formula string 'SqDist(x, y)', K = 1: argmin_j of
the direct-difference squared distance, lowest index on ties.  Pinned by tests/golden/possloss_ref.npz
(gen_golden_possloss.py runs the reference's own PossLoss)."""
from __future__ import annotations

import numpy as np
import torch


def knn(x, y):
    d = ((x[:, None, :] - y[None, :, :]) ** 2).sum(-1)
    return d.argmin(dim=1, keepdim=True)


def quat2mat(quat, eps=0.0):
    """allocentric.py:48-79."""
    nq = quat / (quat.norm(p=2, dim=1, keepdim=True) + eps)
    qw, qx, qy, qz = nq[:, 0], nq[:, 1], nq[:, 2], nq[:, 3]
    X, Y, Z = qx * 2.0, qy * 2.0, qz * 2.0
    wX, wY, wZ = qw * X, qw * Y, qw * Z
    xX, xY, xZ = qx * X, qx * Y, qx * Z
    yY, yZ, zZ = qy * Y, qy * Z, qz * Z
    return torch.stack([1.0 - (yY + zZ), xY - wZ, xZ + wY, xY + wZ, 1.0 - (xX + zZ), yZ - wX, xZ - wY, yZ + wX,
                        1.0 - (xX + yY)], dim=1).reshape(-1, 3, 3)


def allo_to_ego_mat(translation, rot_allo, eps=1e-4):
    """allocentric.py:15-45."""
    cam_ray = torch.tensor([0, 0, 1.0], dtype=translation.dtype, device=translation.device)
    obj_ray = translation / (torch.norm(translation, dim=1, keepdim=True) + eps)
    angle = obj_ray[:, 2:3].acos()
    axis = torch.cross(cam_ray.expand_as(obj_ray), obj_ray, dim=1)
    axis = axis / (torch.norm(axis, dim=1, keepdim=True) + eps)
    q = torch.cat([torch.cos(angle / 2.0), axis[:, 0:1] * torch.sin(angle / 2.0), axis[:, 1:2] * torch.sin(angle / 2.0),
                   axis[:, 2:3] * torch.sin(angle / 2.0)], dim=1)
    return torch.matmul(quat2mat(q), rot_allo)


def predr2rotation(pred_r, bs, num_p):
    """loss.py:253-267."""
    r = pred_r / (torch.norm(pred_r, dim=2).view(bs, num_p, 1))
    r0, r1, r2, r3 = r[:, :, 0], r[:, :, 1], r[:, :, 2], r[:, :, 3]
    cols = [1.0 - 2.0 * (r2 ** 2 + r3 ** 2), 2.0 * r1 * r2 - 2.0 * r0 * r3, 2.0 * r0 * r2 + 2.0 * r1 * r3,
            2.0 * r1 * r2 + 2.0 * r3 * r0, 1.0 - 2.0 * (r1 ** 2 + r3 ** 2), -2.0 * r0 * r1 + 2.0 * r2 * r3,
            -2.0 * r0 * r2 + 2.0 * r1 * r3, 2.0 * r0 * r1 + 2.0 * r2 * r3, 1.0 - 2.0 * (r1 ** 2 + r2 ** 2)]
    return torch.cat([c.view(bs, num_p, 1) for c in cols], dim=2).contiguous().view(bs * num_p, 3, 3)


def poss_loss(pred_r, pred_t, pred_c, target, model_points, idx, refine, w, axis, target_r, sym_list):
    """loss.py:207-251 on torch tensors (any float dtype); returns (loss, dis[0][argmax c[0]], loss_rotation)."""
    bs, num_p, _ = pred_c.size()
    N = model_points.size(1)
    base = predr2rotation(pred_r, bs, num_p)
    base = allo_to_ego_mat(pred_t.contiguous().view(bs * num_p, 3), base)
    model_points = model_points.view(bs, 1, N, 3).repeat(1, num_p, 1, 1).view(bs * num_p, N, 3)
    target = target.view(bs, 1, N, 3).repeat(1, num_p, 1, 1).view(bs * num_p, N, 3)
    target_r = target_r.repeat(bs * num_p, 1, 1)
    pred_t = pred_t.contiguous().view(bs * num_p, 1, 3)
    pred_c = pred_c.contiguous().view(bs * num_p)
    axis = axis.repeat(bs * num_p, 1)
    pred = torch.add(torch.bmm(model_points, base.transpose(1, 2).contiguous()), pred_t)
    if not refine and idx in sym_list:
        tgt = target[0].transpose(1, 0).contiguous().view(3, -1)
        pr = pred.permute(2, 0, 1).contiguous().view(3, -1)
        inds = knn(pr.T.contiguous(), tgt.T.contiguous())
        tgt = torch.index_select(tgt, 1, inds.view(-1))
        target = tgt.view(3, bs * num_p, N).permute(1, 2, 0).contiguous()
        pred = pr.view(3, bs * num_p, N).permute(1, 2, 0).contiguous()
    dis = torch.mean(torch.norm((pred - target), dim=2), dim=1)
    loss = torch.mean((dis * pred_c - w * torch.log(pred_c + 1e-8)), dim=0)
    cos = torch.nn.CosineSimilarity(dim=1, eps=1e-6)
    bt, tt = base.transpose(1, 2).contiguous(), target_r.transpose(1, 2)
    loss_axis = axis[:, 0] * (1. - cos(bt[:, 0], tt[:, 0])) + axis[:, 1] * (1. - cos(bt[:, 1], tt[:, 1])) \
        + axis[:, 2] * (1. - cos(bt[:, 2], tt[:, 2]))
    loss_rot = torch.mean((pred_c * loss_axis - w * torch.log(pred_c + 1e-8)), dim=0)
    pc = pred_c.view(bs, num_p)
    _, which_max = torch.max(pc, 1)
    return loss, dis.view(bs, num_p)[0][which_max[0]], loss_rot


def poss_loss_np(pred_r, pred_t, pred_c, target, model_points, idx, refine, w, axis, target_r, sym_list, gouts=None,
                 dtype=torch.float64):
    """numpy in / numpy out; with ``gouts`` = (g_loss, g_dis, g_rot) also the gradients wrt pred_r, pred_t, pred_c."""
    T = lambda a: torch.as_tensor(np.asarray(a), dtype=dtype)
    pr, pt, pc = T(pred_r).requires_grad_(), T(pred_t).requires_grad_(), T(pred_c).requires_grad_()
    loss, dsel, lrot = poss_loss(pr, pt, pc, T(target), T(model_points), int(idx), bool(refine), float(w), T(axis), T(target_r),
                                 list(sym_list))
    vals = np.array([loss.item(), dsel.item(), lrot.item()])
    if gouts is None:
        return vals
    (gouts[0] * loss + gouts[1] * dsel + gouts[2] * lrot).backward()
    return vals, pr.grad.numpy(), pt.grad.numpy(), pc.grad.numpy()
