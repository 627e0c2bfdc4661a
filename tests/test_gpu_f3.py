"""GPU parity of the SURVEY 8(f) row f3 remainder against runs of the reference's own files
(tests/golden/f3_ref.npz, made by tests/golden/gen_golden_f3.py) and the CPU oracle (oracle/f3_ref.py):
batched Kabsch fit, the ICP-style loop, the older RANSAC rule of lib/transform/trans.py, depth back-projection."""
import os

import numpy as np
import pytest
import torch

import pose_estimation_b200 as pb
from oracle import f3_ref
from pose_estimation_b200 import ops_sim3, synth
from tests.test_oracle_golden import GOLD

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def f3():
    return np.load(os.path.join(GOLD, "f3_ref.npz"))


def test_kabsch_fit_vs_reference_run(f3, dev):
    """best_batch_fit_transform (leastsquares.py:10-48): R, t within 5e-5 of the reference's float32-SVD run, 1e-6 of the
    float64 oracle; the mirrored target gives a proper rotation; all problems in ONE batched call too."""
    n = int(f3["kab_n"])
    for k in range(n):
        A, B = torch.from_numpy(f3[f"kab{k}_A"]).to(dev)[None], torch.from_numpy(f3[f"kab{k}_B"]).to(dev)[None]
        R, t, cA, cB = pb.best_batch_fit_transform(A, B)
        Ro, to, cAo, cBo = f3_ref.kabsch(f3[f"kab{k}_A"], f3[f"kab{k}_B"])
        Rn, tn = R[0].cpu().numpy(), t[0].cpu().numpy()
        assert abs(np.linalg.det(Rn.astype(np.float64)) - 1.0) < 1e-5
        if k != 3:
            assert np.abs(Rn - Ro).max() < 2e-6 and np.abs(tn - to).max() < 2e-6, k
            assert np.abs(Rn - f3[f"kab{k}_R"]).max() < 5e-5 and np.abs(tn - f3[f"kab{k}_t"]).max() < 5e-5, k
        else:        # planar cloud: the rotation about the plane normal's sign is fixed by properness; residuals agree
            A64, B64 = f3[f"kab{k}_A"].astype(np.float64), f3[f"kab{k}_B"].astype(np.float64)
            assert np.linalg.norm(A64 @ Rn.T + tn - B64) <= np.linalg.norm(A64 @ Ro.T + to - B64) + 1e-5
        assert np.abs(cA[0].cpu().numpy() - cAo).max() < 1e-6 and np.abs(cB[0].cpu().numpy() - cBo).max() < 1e-6
    # ragged batch through count[]
    cap = max(f3[f"kab{k}_A"].shape[0] for k in range(n))
    A = torch.zeros((n, cap, 3), device=dev); B = torch.zeros((n, cap, 3), device=dev)
    cnt = torch.zeros(n, dtype=torch.int32, device=dev)
    for k in range(n):
        m = f3[f"kab{k}_A"].shape[0]
        A[k, :m], B[k, :m], cnt[k] = torch.from_numpy(f3[f"kab{k}_A"]).to(dev), torch.from_numpy(f3[f"kab{k}_B"]).to(dev), m
    R, t, st = ops_sim3.kabsch_fit(A, B, count=cnt)
    assert int(st.abs().sum()) == 0
    for k in (0, 1, 2, 4):
        assert np.abs(R[k].cpu().numpy() - f3[f"kab{k}_R"]).max() < 5e-5 and np.abs(t[k].cpu().numpy() - f3[f"kab{k}_t"]).max() < 5e-5
    # degenerate: all points identical -> identity + centroid shift, status set
    Z = torch.full((1, 10, 3), 0.25, device=dev)
    R, t, st = ops_sim3.kabsch_fit(Z, Z + 1.0)
    assert int(st[0]) != 0 and torch.equal(R[0].cpu(), torch.eye(3)) and torch.allclose(t[0].cpu(), torch.ones(3))


@pytest.mark.parametrize("mode", ["reference", "rigid"])
def test_icp_refine_vs_reference_loop(f3, dev, mode):
    """train_declarative.py:80-100, 10 steps, three hypotheses in one call: the poses of the loop re-enacted around
    the reference's own best_batch_fit_transform (2e-4: it fits in float32), and of the oracle (2e-5)."""
    model = torch.from_numpy(f3["icp_model"]).to(dev)
    target = torch.from_numpy(f3["icp_target"]).to(dev)
    R0, t0 = torch.from_numpy(f3["icp_R0"]).to(dev), torch.from_numpy(f3["icp_t0"]).to(dev)
    R, t = pb.icp_refine(model, target, R0, t0, steps=10, compose=mode)
    for h in range(3):
        assert np.abs(R[h].cpu().numpy() - f3[f"icp_{mode}_R"][h]).max() < 2e-4, h
        assert np.abs(t[h].cpu().numpy() - f3[f"icp_{mode}_t"][h]).max() < 2e-4, h
        Ro, to = f3_ref.icp_refine(f3["icp_model"], f3["icp_target"], f3["icp_R0"][h], f3["icp_t0"][h], 10, mode)
        assert np.abs(R[h].cpu().numpy() - Ro).max() < 2e-5 and np.abs(t[h].cpu().numpy() - to).max() < 2e-5, h
    # batched clouds and the tensor-core search give the same poses
    Rb, tb, _ = ops_sim3.icp_refine(model[None].expand(3, -1, -1).contiguous(), target[None].expand(3, -1, -1).contiguous(),
                                    R0, t0, n_steps=10, compose=mode, impl="tc")
    assert torch.equal(Rb, R) and torch.equal(tb, t)


def test_trans_rule_vs_reference_run(f3, dev):
    """lib/transform/trans.py with its np.random.randint draws replayed: every iteration's inlier count (the
    reference's nInliers, which skips index 0), the chosen iteration, success, and the returned transform."""
    for k in range(int(f3["tr_n"])):
        src, dst = f3[f"tr{k}_src"], f3[f"tr{k}_dst"]
        n_it = int(f3[f"tr{k}_niter"])
        o = f3_ref.trans_ransac(src, dst, f3[f"tr{k}_idx"][:n_it])
        cfg = ops_sim3.Sim3Config(n_iter=100, rule="trans")
        res = ops_sim3.sim3_ransac(torch.from_numpy(src).to(dev)[None], torch.from_numpy(dst).to(dev)[None], cfg=cfg,
                                   sample_idx=torch.from_numpy(f3[f"tr{k}_idx"]).to(dev)[None], want_counts=True)
        cnt = res.hyp_counts[0].cpu().numpy()
        assert np.array_equal(cnt[:n_it] / src.shape[0], f3[f"tr{k}_ratio"]), k
        assert int(res.best_hyp[0]) == o["best"], k
        assert (int(res.status[0]) == 0) == bool(f3[f"tr{k}_ok"]), k
        if f3[f"tr{k}_ok"]:
            assert int(res.n_inl[0]) == o["n_refit"]
            assert abs(float(res.scale[0]) - f3[f"tr{k}_S"][0]) <= 1e-6 * f3[f"tr{k}_S"][0]
            assert np.abs(res.R[0].cpu().numpy() - f3[f"tr{k}_R"]).max() < 1e-6
            assert np.abs(res.t[0].cpu().numpy() - f3[f"tr{k}_T"]).max() <= 2e-6 * max(1.0, np.abs(f3[f"tr{k}_T"]).max())
    # the reference-shaped call (its own RNG stream: a different sample, the same kind of answer)
    S, R, T, M = pb.trans_estimateSimilarityTransform(f3["tr1_src"], f3["tr1_dst"])
    assert S.shape == (3,) and R.shape == (3, 3) and M.shape == (4, 4) and np.allclose(M[:3, :3], np.diag(S) @ R)
    assert pb.trans_estimateSimilarityTransform(f3["tr3_src"], f3["tr3_dst"])[0] is None


def test_backprojection_vs_reference(f3, dev):
    """uvd2cloud (coordinate.py:91-101) bit for bit against the reference's run; the depth-map form against the
    numpy lines of batchdataset.py:712-721 (oracle), with list coordinates, index-derived coordinates and a crop."""
    got = pb.uvd2cloud(torch.from_numpy(f3["uvd"]), torch.from_numpy(f3["uvd_K"]))
    assert np.array_equal(got.cpu().numpy(), f3["uvd_xyz"])
    rng = np.random.default_rng(9)
    F, H, W, NP = 3, 120, 160, 500
    depth = rng.uniform(0.4, 1.6, (F, H, W)).astype(np.float32)
    choose = np.sort(rng.integers(0, H * W, (F, NP)), axis=1).astype(np.int64)
    roi = rng.integers(0, 200, (F, 2)).astype(np.int32)
    xm = (choose % W + roi[:, 1:2]).astype(np.float32)
    ym = (choose // W + roi[:, 0:1]).astype(np.float32)
    K4 = np.tile(synth.LINEMOD_K4, (F, 1)).astype(np.float32)
    K4[1] = (700.0, 700.0, 320.0, 240.0)
    t = lambda a: torch.from_numpy(a).to(dev)
    want = np.stack([f3_ref.backproject(depth[f], choose[f], xm[f], ym[f], K4[f], 0.001) for f in range(F)])
    c1 = ops_sim3.backproject_cloud(t(depth), t(choose), t(K4), xmap=t(xm), ymap=t(ym), depth_scale=0.001)
    c2 = ops_sim3.backproject_cloud(t(depth), t(choose.astype(np.int32)), t(K4), roi=t(roi), depth_scale=0.001)
    assert np.array_equal(c1.cpu().numpy(), want) and np.array_equal(c2.cpu().numpy(), want)
    # the cloud feeds the T-offset vote: EvalSweep(cloud=...) accepts it as it comes
    t_res = torch.zeros((F, 3, NP), device=dev)
    v = pb.vote_translation(c1, t_res)
    assert torch.allclose(v, c1.mean(dim=1), atol=1e-6)
