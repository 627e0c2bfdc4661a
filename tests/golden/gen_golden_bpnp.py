"""Generate tests/golden/bpnp_ref.npz by RUNNING THE REFERENCE: imports
/root/reference/lib/network/dnn/BPnP.py in this container and calls BPnP.apply(...).backward on seeded
synthetic keypoint sets (CPU tensors; the file's cv2 calls run for real).  The reference imports kornia,
which is not installed here: a stub module exposing only
``kornia.geometry.conversions.angle_axis_to_rotation_matrix`` (the restatement in oracle/bpnp_ref.py) is
put in sys.modules first -- everything else that runs is the reference's own code.
Run here only; the GPU box has no /root/reference.

    python tests/golden/gen_golden_bpnp.py
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import bpnp_ref  # noqa: E402

kn = types.ModuleType("kornia")
kn.geometry = types.ModuleType("kornia.geometry")
kn.geometry.conversions = types.ModuleType("kornia.geometry.conversions")
kn.geometry.conversions.angle_axis_to_rotation_matrix = bpnp_ref.angle_axis_to_rotation_matrix
sys.modules["kornia"] = kn
sys.modules["kornia.geometry"] = kn.geometry
sys.modules["kornia.geometry.conversions"] = kn.geometry.conversions

REF = "/root/reference/lib/network/dnn/BPnP.py"
spec = importlib.util.spec_from_file_location("ref_bpnp", REF)
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

K_LM = np.array([[572.4114, 0.0, 325.2611], [0.0, 573.57043, 242.04899], [0.0, 0.0, 1.0]], dtype=np.float32)


def rodrigues(w):
    th = np.linalg.norm(w)
    k = w / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * Kx @ Kx


def make_case(rng, bs, n, noise_px, outliers=0):
    pts3d = ((rng.random((n, 3)) - 0.5) * np.array([0.10, 0.08, 0.14])).astype(np.float32)
    pts2d = np.zeros((bs, n, 2), dtype=np.float32)
    poses = np.zeros((bs, 6))
    for b in range(bs):
        w = rng.normal(size=3); w *= rng.uniform(0.3, 2.5) / np.linalg.norm(w)
        t = np.array([rng.uniform(-0.15, 0.15), rng.uniform(-0.1, 0.1), rng.uniform(0.6, 1.2)])
        q = pts3d.astype(np.float64) @ rodrigues(w).T + t
        p = q @ K_LM.astype(np.float64).T
        uv = p[:, :2] / p[:, 2:3] + rng.normal(size=(n, 2)) * noise_px
        if outliers:
            bad = rng.choice(n, outliers, replace=False)
            uv[bad] += rng.uniform(-40, 40, size=(outliers, 2))
        pts2d[b] = uv
        poses[b] = np.concatenate([w, t])
    return pts2d, pts3d, poses


def run_reference(pts2d, pts3d, K, grad_out, ini_pose=None):
    x = torch.tensor(pts2d, requires_grad=True)
    z = torch.tensor(pts3d, requires_grad=True)
    Kt = torch.tensor(K, requires_grad=True)
    ini = None if ini_pose is None else torch.tensor(ini_pose, dtype=torch.float32)
    P = ref.BPnP.apply(x, z, Kt, ini)
    P.backward(torch.tensor(grad_out))
    return P.detach().numpy(), x.grad.numpy(), z.grad.numpy(), Kt.grad.numpy()


def main():
    rng = np.random.default_rng(20261104)
    out = {}
    cases = [("a", 4, 12, 0.5, 0), ("b", 3, 9, 0.2, 0), ("c", 2, 40, 1.0, 0), ("d", 2, 24, 0.3, 2)]
    for name, bs, n, noise, nout in cases:
        pts2d, pts3d, poses = make_case(rng, bs, n, noise, nout)
        g = rng.normal(size=(bs, 6)).astype(np.float32)
        P, gx, gz, gK = run_reference(pts2d, pts3d, K_LM, g)
        for k, v in (("pts2d", pts2d), ("pts3d", pts3d), ("gt_pose", poses), ("grad_out", g), ("P6d", P), ("grad_x", gx),
                     ("grad_z", gz), ("grad_K", gK)):
            out[f"{name}_{k}"] = v
        print(name, "P", P[0], "|gx|", np.abs(gx).max(), "|gz|", np.abs(gz).max(), "|gK|", np.abs(gK).max())
    out["K"] = K_LM
    out["cases"] = np.array([c[0] for c in cases])
    # one case with ini_pose given (BPnP.py:39-41): the RANSAC call is skipped
    pts2d, pts3d, poses = make_case(rng, 2, 10, 0.4, 0)
    ini = (poses + rng.normal(size=poses.shape) * 0.02).astype(np.float32)
    g = rng.normal(size=(2, 6)).astype(np.float32)
    P, gx, gz, gK = run_reference(pts2d, pts3d, K_LM, g, ini)
    for k, v in (("pts2d", pts2d), ("pts3d", pts3d), ("ini", ini), ("grad_out", g), ("P6d", P), ("grad_x", gx),
                 ("grad_z", gz), ("grad_K", gK)):
        out[f"ini_{k}"] = v
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "bpnp_ref.npz"), **out)


if __name__ == "__main__":
    main()
