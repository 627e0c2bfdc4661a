"""Generate tests/golden/possloss_ref.npz by RUNNING THE REFERENCE: imports
/root/reference/version/transparent/lib/networks/loss.py in this container and runs PossLoss.forward + autograd
backward on seeded synthetic hypothesis sets (CPU float32).  Two third-party imports of that file are absent here
and are stubbed before the import: ``pykeops.torch.generic_argkmin`` (only its K=1 'SqDist' handle is used: replaced
by a float32 brute-force argmin of the direct-difference squared distance, lowest index on ties) and ``transforms3d``
(imported by lib/transform/allocentric.py but unused by allo_to_ego_mat_torch).  Everything else that runs is the
reference's own code.  Run here only; the GPU box has no /root/reference.

    python tests/golden/gen_golden_possloss.py
"""
import os
import sys
import types

import numpy as np
import torch


def _stub(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def knn_bruteforce(x, y):
    d = ((x[:, None, :] - y[None, :, :]) ** 2).sum(-1)
    return d.argmin(dim=1, keepdim=True)


_any = lambda *a, **k: None
_stub("pykeops"); _stub("pykeops.torch", generic_argkmin=lambda *a, **k: knn_bruteforce)
_stub("transforms3d")
_stub("transforms3d.euler", euler2quat=_any, mat2euler=_any, quat2euler=_any, euler2mat=_any)
_stub("transforms3d.quaternions", axangle2quat=_any, mat2quat=_any, qinverse=_any, qmult=_any, quat2mat=_any)
_stub("transforms3d.axangles", axangle2mat=_any)
sys.path.insert(0, "/root/reference/version/transparent")
from lib.networks.loss import PossLoss  # noqa: E402


def rand_rot(rng):
    q = rng.normal(size=4); q /= np.linalg.norm(q)
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]]), q


def make_case(rng, bs, num_p, N):
    model = ((rng.random((bs, N, 3)) - 0.5) * np.array([0.10, 0.08, 0.14])).astype(np.float32)
    Rg, qg = rand_rot(rng)
    tg = np.array([rng.uniform(-0.1, 0.1), rng.uniform(-0.1, 0.1), rng.uniform(0.5, 1.0)])
    target = (model.astype(np.float64) @ Rg.T + tg).astype(np.float32)
    pred_r = (qg[None, None, :] * rng.uniform(0.5, 2.0, size=(bs, num_p, 1)) + rng.normal(size=(bs, num_p, 4)) * 0.15).astype(np.float32)
    pred_t = (tg[None, None, :] + rng.normal(size=(bs, num_p, 3)) * 0.01).astype(np.float32)
    pred_c = rng.uniform(0.05, 0.95, size=(bs, num_p, 1)).astype(np.float32)
    axis = np.array([[1.0, 0.0, 1.0]], dtype=np.float32)
    return dict(model=model, target=target, pred_r=pred_r, pred_t=pred_t, pred_c=pred_c, axis=axis,
                target_r=Rg[None].astype(np.float32))


def run_reference(c, idx, sym_list, refine, w, gouts):
    N = c["model"].shape[1]
    lossfn = PossLoss(N, sym_list, knn_bruteforce)
    pr = torch.tensor(c["pred_r"], requires_grad=True)
    pt = torch.tensor(c["pred_t"], requires_grad=True)
    pc = torch.tensor(c["pred_c"], requires_grad=True)
    loss, dsel, lrot = lossfn(pr, pt, pc, torch.tensor(c["target"]), torch.tensor(c["model"]), idx, refine, w,
                              torch.tensor(c["axis"]), torch.tensor(c["target_r"]))
    (gouts[0] * loss + gouts[1] * dsel + gouts[2] * lrot).backward()
    return (np.array([loss.item(), dsel.item(), lrot.item()], np.float32), pr.grad.numpy(), pt.grad.numpy(), pc.grad.numpy())


def main():
    rng = np.random.default_rng(20261105)
    out = {}
    cases = [("s1", 1, 40, 64, 7, False), ("n1", 1, 40, 64, 2, False), ("s2", 2, 16, 50, 7, False), ("r1", 1, 24, 33, 7, True)]
    for name, bs, num_p, N, idx, refine in cases:
        c = make_case(rng, bs, num_p, N)
        gouts = (1.0, 0.25, 0.5)
        vals, gr, gt, gc = run_reference(c, idx, [7, 8], refine, 0.015, gouts)
        for k, v in c.items():
            out[f"{name}_{k}"] = v
        out[f"{name}_vals"], out[f"{name}_grad_r"], out[f"{name}_grad_t"], out[f"{name}_grad_c"] = vals, gr, gt, gc
        out[f"{name}_meta"] = np.array([idx, int(refine)], np.int32)
        print(name, vals, np.abs(gr).max(), np.abs(gt).max(), np.abs(gc).max())
    out["cases"] = np.array([c[0] for c in cases])
    out["sym_list"] = np.array([7, 8], np.int32)
    out["w"] = np.float32(0.015)
    out["gouts"] = np.array([1.0, 0.25, 0.5], np.float32)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "possloss_ref.npz"), **out)


if __name__ == "__main__":
    main()
