"""GPU parity of the per-hypothesis pose loss (csrc/possloss.cu) against the CPU oracle (oracle/possloss_ref.py,
float64 autograd restatement of version/transparent/lib/networks/loss.py PossLoss) and, through
tests/golden/possloss_ref.npz, against the reference's own run.  Tolerances (floating point, the kernel transforms
and measures distances in float32 like the reference): values 1e-5 relative, gradients 1e-4 of the largest entry."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def _golden():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "possloss_ref.npz"))


def _rel(a, b):
    return float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / np.abs(b).max())


def _run_module(dev, c, g, name, idx, refine, gouts):
    import pose_estimation_b200 as pb
    T = lambda k: torch.from_numpy(np.ascontiguousarray(g[f"{name}_{k}"], dtype=np.float32)).to(dev)
    pr, pt, pc = T("pred_r").requires_grad_(), T("pred_t").requires_grad_(), T("pred_c").requires_grad_()
    lossfn = pb.PossLoss(g[f"{name}_model"].shape[1], [int(s) for s in g["sym_list"]], None)
    loss, dsel, lrot = lossfn(pr, pt, pc, T("target"), T("model"), int(idx), bool(refine), float(g["w"]), T("axis"), T("target_r"))
    (float(gouts[0]) * loss + float(gouts[1]) * dsel + float(gouts[2]) * lrot).backward()
    vals = np.array([loss.item(), dsel.item(), lrot.item()])
    return vals, pr.grad.cpu().numpy(), pt.grad.cpu().numpy(), pc.grad.cpu().numpy()


def test_values_and_gradients_match_oracle_and_reference_run(dev):
    from oracle import possloss_ref as orc
    g = _golden()
    for c in g["cases"]:
        idx, refine = g[f"{c}_meta"]
        v, gr, gt, gc = _run_module(dev, c, g, c, idx, refine, g["gouts"])
        ov, ogr, ogt, ogc = orc.poss_loss_np(g[f"{c}_pred_r"], g[f"{c}_pred_t"], g[f"{c}_pred_c"], g[f"{c}_target"], g[f"{c}_model"],
                                             idx, refine, float(g["w"]), g[f"{c}_axis"], g[f"{c}_target_r"], g["sym_list"],
                                             gouts=g["gouts"], dtype=torch.float64)
        for mine, o, ref, tol in ((v, ov, g[f"{c}_vals"], 1e-5), (gr, ogr, g[f"{c}_grad_r"], 1e-4), (gt, ogt, g[f"{c}_grad_t"], 1e-4),
                                  (gc, ogc, g[f"{c}_grad_c"], 1e-4)):
            assert _rel(mine, o) < tol, (c, _rel(mine, o))
            assert _rel(mine, ref) < tol, (c, _rel(mine, ref))


def test_full_size_symmetric_case_matches_oracle_run_on_device(dev):
    """bs*num_p*N = 250 000 predicted points against 500 target points (loss.py:228 comment) -- the oracle's torch code
    runs on the GPU in float64 here only because its [250000,500,3] temporary is too slow on the host."""
    from oracle import possloss_ref as orc
    import pose_estimation_b200 as pb
    rng = np.random.default_rng(5)
    num_p, N = 500, 500
    model = ((rng.random((1, N, 3)) - 0.5) * np.array([0.10, 0.08, 0.14])).astype(np.float32)
    q = rng.normal(size=4); q /= np.linalg.norm(q)
    Rg = orc.quat2mat(torch.tensor(q[None])).numpy()[0]
    tg = np.array([0.05, -0.03, 0.8])
    target = (model.astype(np.float64) @ Rg.T + tg).astype(np.float32)
    pred_r = (q[None, None] * rng.uniform(0.5, 2.0, size=(1, num_p, 1)) + rng.normal(size=(1, num_p, 4)) * 0.1).astype(np.float32)
    pred_t = (tg[None, None] + rng.normal(size=(1, num_p, 3)) * 0.01).astype(np.float32)
    pred_c = rng.uniform(0.05, 0.95, size=(1, num_p, 1)).astype(np.float32)
    axis = np.array([[0.0, 1.0, 1.0]], np.float32)
    T = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(dev).to(dt)
    pr, pt, pc = T(pred_r).requires_grad_(), T(pred_t).requires_grad_(), T(pred_c).requires_grad_()
    loss, dsel, lrot = pb.PossLoss(N, [7, 8])(pr, pt, pc, T(target), T(model), 8, False, 0.015, T(axis), T(Rg[None]))
    (loss + 0.5 * dsel + 0.25 * lrot).backward()
    d = torch.float64
    opr, opt_, opc = T(pred_r, d).requires_grad_(), T(pred_t, d).requires_grad_(), T(pred_c, d).requires_grad_()
    ol, od, orot = orc.poss_loss(opr, opt_, opc, T(target, d), T(model, d), 8, False, 0.015, T(axis, d), T(Rg[None], d), [7, 8])
    (ol + 0.5 * od + 0.25 * orot).backward()
    assert abs(loss.item() - ol.item()) < 1e-5 * abs(ol.item())
    assert abs(dsel.item() - od.item()) < 1e-5 * abs(od.item())
    assert abs(lrot.item() - orot.item()) < 1e-5 * abs(orot.item())
    assert _rel(pr.grad.cpu().numpy(), opr.grad.cpu().numpy()) < 1e-3
    assert _rel(pt.grad.cpu().numpy(), opt_.grad.cpu().numpy()) < 1e-3
    assert _rel(pc.grad.cpu().numpy(), opc.grad.cpu().numpy()) < 1e-4


def test_argument_errors_and_helper(dev):
    import pose_estimation_b200 as pb
    from pose_estimation_b200 import ops_poss
    z = torch.zeros((1, 4, 4), device=dev)
    with pytest.raises(ValueError):
        ops_poss.poss_loss_fwd(z, torch.zeros((1, 5, 3), device=dev), torch.zeros((1, 4, 1), device=dev), torch.zeros((1, 8, 3), device=dev),
                               torch.zeros((1, 8, 3), device=dev), 0, 0.015, torch.zeros(3, device=dev), torch.eye(3, device=dev))
    q = torch.tensor([[[2.0, 0.0, 0.0, 0.0], [0.0, 0.0, 3.0, 0.0]]], device=dev)
    R = pb.PossLoss.predr2rotation(q, 1, 2).cpu().numpy()
    np.testing.assert_allclose(R[0], np.eye(3), atol=1e-7)
    np.testing.assert_allclose(R[1], np.diag([-1.0, 1.0, -1.0]), atol=1e-7)
