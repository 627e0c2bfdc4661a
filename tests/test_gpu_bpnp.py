"""GPU parity of back-propagatable PnP (csrc/bpnp.cu) against the CPU oracle (oracle/bpnp_ref.py, float64 autograd
restatement of lib/network/dnn/BPnP.py) and, through tests/golden/bpnp_ref.npz, against the reference's own run.
Tolerances (floating point): gradients relative 1e-5 of the largest entry against the float64 oracle (outputs are
rounded to float32), 5e-4 against the reference run (which computes in float32 and inverts J_fy in float32);
poses 1e-4 rad / 1e-4 m against cv2."""
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
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bpnp_ref.npz"))


def _rel(a, b):
    return float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / np.abs(b).max())


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)


def test_backward_matches_oracle_and_reference_run(dev):
    from oracle import bpnp_ref as br
    from pose_estimation_b200 import ops_bpnp
    g = _golden()
    for c in list(g["cases"]) + ["ini"]:
        x, z, K, P, go = (g[f"{c}_pts2d"], g[f"{c}_pts3d"], g["K"], g[f"{c}_P6d"], g[f"{c}_grad_out"])
        gx, gz, gK, st = ops_bpnp.bpnp_backward(_t(x, dev), _t(z, dev), _t(K, dev), _t(P, dev), _t(go, dev))
        assert st.cpu().numpy().tolist() == [0] * x.shape[0]
        ox, oz, oK = br.bpnp_backward(x, z, K, P, go, dtype=torch.float64)
        for mine, orc, ref in ((gx, ox, g[f"{c}_grad_x"]), (gz, oz, g[f"{c}_grad_z"]), (gK, oK, g[f"{c}_grad_K"])):
            mine = mine.cpu().numpy()
            assert _rel(mine, orc) < 1e-5, (c, _rel(mine, orc))
            assert _rel(mine, ref) < 5e-4, (c, _rel(mine, ref))


def test_backward_batched_points_small_angles_and_larger_sets(dev):
    from oracle import bpnp_ref as br
    from pose_estimation_b200 import ops_bpnp
    rng = np.random.default_rng(7)
    K = _golden()["K"]
    for B, n, small in ((3, 150, False), (2, 17, True), (5, 3, False)):
        z = ((rng.random((B, n, 3)) - 0.5) * np.array([0.10, 0.08, 0.14])).astype(np.float32)
        P = np.zeros((B, 6), np.float32)
        x = np.zeros((B, n, 2), np.float32)
        for b in range(B):
            w = rng.normal(size=3); w *= (3e-4 if small else rng.uniform(0.2, 2.8)) / np.linalg.norm(w)
            t = np.array([rng.uniform(-0.1, 0.1), rng.uniform(-0.1, 0.1), rng.uniform(0.6, 1.2)])
            P[b] = np.concatenate([w, t])
            R = br.angle_axis_to_rotation_matrix(torch.tensor(w[None])).numpy()[0]
            p = (z[b].astype(np.float64) @ R.T + t) @ K.astype(np.float64).T
            x[b] = p[:, :2] / p[:, 2:3] + rng.normal(size=(n, 2)) * 0.5
        go = rng.normal(size=(B, 6)).astype(np.float32)
        gx, gz, gK, st = ops_bpnp.bpnp_backward(_t(x, dev), _t(z, dev), _t(K, dev), _t(P, dev), _t(go, dev))
        ox, oz, oK = br.bpnp_backward(x, z, K, P, go, dtype=torch.float64, pts3d_batched=True)
        assert st.cpu().numpy().tolist() == [0] * B
        assert _rel(gx.cpu().numpy(), ox) < 1e-5 and _rel(gz.cpu().numpy(), oz) < 1e-5 and _rel(gK.cpu().numpy(), oK) < 1e-5
        # shared-point form on the first sample's points: the batch sum is taken in sample order
        gx1, gz1, gK1, _ = ops_bpnp.bpnp_backward(_t(x, dev), _t(z[0], dev), _t(K, dev), _t(P, dev), _t(go, dev))
        o1 = br.bpnp_backward(x, z[0], K, P, go, dtype=torch.float64)
        assert _rel(gx1.cpu().numpy(), o1[0]) < 1e-5 and _rel(gz1.cpu().numpy(), o1[1]) < 1e-5 and _rel(gK1.cpu().numpy(), o1[2]) < 1e-5


def test_backward_flags_singular_systems(dev):
    from pose_estimation_b200 import ops_bpnp
    K = _golden()["K"]
    z = np.zeros((6, 3), np.float32)                               # all points coincide: J_fy has no rank
    x = np.full((1, 6, 2), 100.0, np.float32)
    P = np.array([[0.1, 0.2, 0.3, 0.0, 0.0, 1.0]], np.float32)
    gx, gz, gK, st = ops_bpnp.bpnp_backward(_t(x, dev), _t(z, dev), _t(K, dev), _t(P, dev), _t(np.ones((1, 6)), dev))
    assert st.cpu().numpy().tolist() == [1]
    assert float(gx.abs().max()) == 0.0 and float(gz.abs().max()) == 0.0 and float(gK.abs().max()) == 0.0


def _pose_close(P, Q, tol=1e-4):
    from tests.util import rot_angle
    import cv2
    for a, b in zip(np.asarray(P, np.float64), np.asarray(Q, np.float64)):
        Ra, Rb = cv2.Rodrigues(a[:3])[0], cv2.Rodrigues(b[:3])[0]
        assert rot_angle(Ra, Rb) < tol, rot_angle(Ra, Rb)
        assert np.abs(a[3:] - b[3:]).max() < tol, np.abs(a[3:] - b[3:]).max()


def test_forward_matches_cv2_run_of_the_reference(dev):
    from pose_estimation_b200 import ops_bpnp
    g = _golden()
    for c in list(g["cases"]):
        x, z, K = _t(g[f"{c}_pts2d"], dev), _t(g[f"{c}_pts3d"], dev), _t(g["K"], dev)
        P, st = ops_bpnp.bpnp_forward(x, z, K)
        assert int(st.max()) == 0
        if c != "d":                                               # d has gross outliers: both minimise over ALL points
            _pose_close(P.cpu().numpy(), g[f"{c}_P6d"])
        else:
            _pose_close(P.cpu().numpy(), g[f"{c}_P6d"], tol=2e-3)
    x, z, K = _t(g["ini_pts2d"], dev), _t(g["ini_pts3d"], dev), _t(g["K"], dev)
    P, st, rmse = ops_bpnp.pnp_refine(x, z, K, pose6_init=_t(g["ini_ini"], dev), want_rmse=True)
    _pose_close(P.cpu().numpy(), g["ini_P6d"])
    assert float(rmse.max()) < 1.0


def test_autograd_function_mirrors_the_reference_module(dev):
    import pose_estimation_b200 as pb
    g = _golden()
    x = _t(g["a_pts2d"], dev).requires_grad_()
    z = _t(g["a_pts3d"], dev).requires_grad_()
    K = _t(g["K"], dev).requires_grad_()
    P = pb.BPnPModle()(x, z, K)
    P.backward(_t(g["a_grad_out"], dev))
    assert P.shape == (4, 6) and x.grad.shape == x.shape and z.grad.shape == z.shape and K.grad.shape == (3, 3)
    # gradients at the module's own optimum stay close to the reference run's (poses agree to ~1e-5)
    assert _rel(x.grad.cpu().numpy(), g["a_grad_x"]) < 2e-2
    assert _rel(z.grad.cpu().numpy(), g["a_grad_z"]) < 2e-2
    P2 = pb.BPnP.apply(x, z, K, P.detach())
    np.testing.assert_allclose(P2.detach().cpu().numpy(), P.detach().cpu().numpy(), atol=1e-6)
