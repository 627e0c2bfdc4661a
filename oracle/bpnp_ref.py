"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's back-propagatable PnP
(/root/reference/lib/network/dnn/BPnP.py).  Only tests/, __graft_entry__.smoke() and bench.py's CPU
legs may import this file; the product path never does.

What is restated, line for line in meaning:
  * ``angle_axis_to_rotation_matrix``  -- kornia 0.6.x (third-party, NOT vendored in the reference and
    NOT installed here: restated from its published formula; this one function is "parity unpinned"
    against kornia itself).  Normal branch  R = cos I + (1-cos) w w^T + sin [w]x  with the axis
    w = aa / (theta + 1e-6), first-order Taylor branch  I + [aa]x  when theta^2 <= 1e-6.
  * ``batch_project``   BPnP.py:141-159
  * ``get_coefs``       BPnP.py:127-138   coefs[i,k,j] = -2 d proj[i,k] / d y_j
  * ``bpnp_forward``    BPnP.py:24-51     cv2.solvePnPRansac(ITERATIVE, 3 px, 0.9999) -> cv2.solvePnP(
                                          ITERATIVE, useExtrinsicGuess) on ALL points
  * ``bpnp_backward``   BPnP.py:54-117    implicit-function gradient: f_j = sum_i coef_ij . r_i,
                                          r_i = x_i S_i - (K P z_i)[0:2];  J_yx = -J_fy^-1 J_fx etc.
The restatement runs the same torch.autograd recipe as the reference (double backward through the
projection); ``dtype`` selects float32 (what the reference computes in) or float64 (the checker the CUDA
kernel, which works in FP64, is compared with).  The file is pinned by tests/golden/bpnp_ref.npz, which
was produced by importing the reference's own BPnP.py in the build container (gen_golden_bpnp.py) with
this module's rotation function standing in for the absent kornia.
"""
from __future__ import annotations

import numpy as np
import torch


def angle_axis_to_rotation_matrix(angle_axis: torch.Tensor) -> torch.Tensor:
    """[B,3] -> [B,3,3]; kornia.geometry.conversions.angle_axis_to_rotation_matrix (0.6.x)."""
    eps = 1e-6
    aa = angle_axis
    theta2 = (aa * aa).sum(dim=1, keepdim=True)                      # [B,1]
    theta = torch.sqrt(theta2)
    w = aa / (theta + eps)
    wx, wy, wz = w[:, 0:1], w[:, 1:2], w[:, 2:3]
    c, s = torch.cos(theta), torch.sin(theta)
    one = 1.0
    r00 = c + wx * wx * (one - c)
    r10 = wz * s + wx * wy * (one - c)
    r20 = -wy * s + wx * wz * (one - c)
    r01 = wx * wy * (one - c) - wz * s
    r11 = c + wy * wy * (one - c)
    r21 = wx * s + wy * wz * (one - c)
    r02 = wy * s + wx * wz * (one - c)
    r12 = -wx * s + wy * wz * (one - c)
    r22 = c + wz * wz * (one - c)
    normal = torch.cat([r00, r01, r02, r10, r11, r12, r20, r21, r22], dim=1).view(-1, 3, 3)
    rx, ry, rz = aa[:, 0:1], aa[:, 1:2], aa[:, 2:3]
    k1 = torch.ones_like(rx)
    taylor = torch.cat([k1, -rz, ry, rz, k1, -rx, -ry, rx, k1], dim=1).view(-1, 3, 3)
    mask = (theta2 > eps).view(-1, 1, 1)
    m1 = mask.type_as(theta2)
    m0 = (~mask).type_as(theta2)
    return m1 * normal + m0 * taylor


def batch_project(P, pts3d, K):
    """BPnP.py:141-159 (angle_axis=True): P [bs,6], pts3d [n,3], K [3,3] -> [bs,n,2]."""
    n = pts3d.size(0)
    bs = P.size(0)
    pts3d_h = torch.cat((pts3d, torch.ones(n, 1, dtype=pts3d.dtype)), dim=-1)
    R_out = angle_axis_to_rotation_matrix(P[:, 0:3].view(bs, 3))
    PM = torch.cat((R_out[:, 0:3, 0:3], P[:, 3:6].view(bs, 3, 1)), dim=-1)
    pts3d_cam = pts3d_h.matmul(PM.transpose(-2, -1))
    pts2d_proj = pts3d_cam.matmul(K.t())
    S = pts2d_proj[:, :, 2].view(bs, n, 1)
    return pts2d_proj[:, :, 0:2].div(S)


def get_coefs(P_6d, pts3d, K, create_graph=True):
    """BPnP.py:127-138."""
    n = pts3d.size(0)
    m = P_6d.size(-1)
    coefs = torch.zeros(n, 2, m, dtype=P_6d.dtype)
    y = P_6d.repeat(n, 1)
    proj = batch_project(y, pts3d, K)                                 # [n,n,2]
    vec = torch.diag(torch.ones(n, dtype=P_6d.dtype))
    for k in range(2):
        y_grad = torch.autograd.grad(proj[:, :, k], y, vec, retain_graph=True, create_graph=create_graph)
        coefs[:, k, :] = -2 * y_grad[0].clone()
    return coefs


def bpnp_forward(pts2d, pts3d, K, ini_pose=None):
    """BPnP.py:24-51 with numpy in/out: pts2d [bs,n,2], pts3d [n,3], K [3,3] -> P_6d [bs,6] float32."""
    import cv2 as cv
    pts2d = np.asarray(pts2d, dtype=np.float32)
    pts3d_np = np.asarray(pts3d, dtype=np.float32)
    K_np = np.asarray(K, dtype=np.float32)
    bs, n = pts2d.shape[0], pts2d.shape[1]
    out = np.zeros((bs, 6), dtype=np.float32)
    for i in range(bs):
        p2 = np.ascontiguousarray(pts2d[i]).reshape((n, 1, 2))
        if ini_pose is None:
            _, rvec0, T0, _ = cv.solvePnPRansac(objectPoints=pts3d_np, imagePoints=p2, cameraMatrix=K_np,
                                                distCoeffs=None, flags=cv.SOLVEPNP_ITERATIVE, confidence=0.9999,
                                                reprojectionError=3)
        else:
            rvec0 = np.array(ini_pose[i, 0:3], dtype=np.float64).reshape(3, 1)
            T0 = np.array(ini_pose[i, 3:6], dtype=np.float64).reshape(3, 1)
        _, rvec, T = cv.solvePnP(objectPoints=pts3d_np, imagePoints=p2, cameraMatrix=K_np, distCoeffs=None,
                                 flags=cv.SOLVEPNP_ITERATIVE, useExtrinsicGuess=True, rvec=rvec0, tvec=T0)
        out[i, 0:3] = np.asarray(rvec, dtype=np.float32).reshape(3)
        out[i, 3:6] = np.asarray(T, dtype=np.float32).reshape(3)
    return out


def bpnp_backward(pts2d, pts3d, K, P_6d, grad_output, dtype=torch.float64, pts3d_batched=False):
    """BPnP.py:54-117.  numpy in/out.  Returns (grad_x [bs,n,2], grad_z [n,3] (or [bs,n,3] when
    pts3d_batched: every sample has its own 3-D points and the sum over the batch is not taken),
    grad_K [3,3]) in ``dtype`` precision."""
    pts2d = torch.as_tensor(np.asarray(pts2d), dtype=dtype)
    pts3d = torch.as_tensor(np.asarray(pts3d), dtype=dtype)
    K = torch.as_tensor(np.asarray(K), dtype=dtype)
    P_6d = torch.as_tensor(np.asarray(P_6d), dtype=dtype)
    g = torch.as_tensor(np.asarray(grad_output), dtype=dtype)
    bs, n, m = pts2d.size(0), pts2d.size(1), 6
    grad_x = torch.zeros_like(pts2d)
    grad_z = torch.zeros_like(pts3d)
    grad_K = torch.zeros_like(K)
    for i in range(bs):
        z_i = pts3d[i] if pts3d_batched else pts3d
        J_fy = torch.zeros(m, m, dtype=dtype)
        J_fx = torch.zeros(m, 2 * n, dtype=dtype)
        J_fz = torch.zeros(m, 3 * n, dtype=dtype)
        J_fK = torch.zeros(m, 9, dtype=dtype)
        with torch.enable_grad():
            x_f = pts2d[i].clone().contiguous().view(-1).detach().requires_grad_()
            y_f = P_6d[i].clone().contiguous().view(-1).detach().requires_grad_()
            z_f = z_i.clone().contiguous().view(-1).detach().requires_grad_()
            K_f = K.clone().contiguous().view(-1).detach().requires_grad_()
            for j in range(m):
                for v in (x_f, y_f, z_f, K_f):
                    if v.grad is not None:
                        v.grad.zero_()
                R = angle_axis_to_rotation_matrix(y_f[0:3].view(1, 3))
                P = torch.cat((R[0, 0:3, 0:3].view(3, 3), y_f[3:6].view(3, 1)), dim=-1)
                KP = torch.mm(K_f.view(3, 3), P)
                x_i = x_f.view(n, 2).transpose(0, 1)
                z_h = torch.cat((z_f.view(n, 3), torch.ones(n, 1, dtype=dtype)), dim=-1).t()
                proj_i = KP.mm(z_h)
                Si = proj_i[2, :].view(1, n)
                r = x_i * Si - proj_i[0:2, :]
                coefs = get_coefs(y_f.view(1, 6), z_f.view(n, 3), K_f.view(3, 3))
                coef = coefs[:, :, j].transpose(0, 1)
                fj = (coef * r).sum()
                fj.backward()
                J_fy[j, :] = y_f.grad.clone()
                J_fx[j, :] = x_f.grad.clone()
                J_fz[j, :] = z_f.grad.clone()
                J_fK[j, :] = K_f.grad.clone()
        inv_J_fy = torch.inverse(J_fy)
        J_yx = (-1) * torch.mm(inv_J_fy, J_fx)
        J_yz = (-1) * torch.mm(inv_J_fy, J_fz)
        J_yK = (-1) * torch.mm(inv_J_fy, J_fK)
        grad_x[i] = g[i].view(1, m).mm(J_yx).view(n, 2)
        if pts3d_batched:
            grad_z[i] = g[i].view(1, m).mm(J_yz).view(n, 3)
        else:
            grad_z += g[i].view(1, m).mm(J_yz).view(n, 3)
        grad_K += g[i].view(1, m).mm(J_yK).view(3, 3)
    return grad_x.numpy(), grad_z.numpy(), grad_K.numpy()
