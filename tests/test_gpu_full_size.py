"""GPU tests at BASELINE.json's FULL sizes, through properties that do not need the CPU oracle to run the whole
workload: additivity of the accuracy table over a split of the batch (a checksum of checksums), determinism,
batch-size independence, self-match / permutation properties of the nearest-neighbour search -- plus the oracle
on a handful of frames taken out of the full-size launch (same frame index, so the same hypothesis stream)."""
import numpy as np
import pytest
import torch

import bench
import pose_estimation_b200 as pb
from pose_estimation_b200 import ops, ops_metric, synth
from tests.util import rot_angle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2(dev):
    """bench.py's workload: 1024 frames of config C2 resident on the device (4.09 GB)."""
    return bench.build_inputs(torch, dev, 0, bench.FRAMES_PER_GPU)


def _sweep(dev, F, cfg=None):
    return pb.EvalSweep(torch.from_numpy(synth.class_models(bench.N_MODEL)), torch.from_numpy(synth.DIAMETER),
                        torch.from_numpy(synth.IS_SYM), F, synth.IMG_H, synth.IMG_W, bench.CAP,
                        cfg or ops.RansacConfig(n_hyp=bench.N_HYP, seed=bench.SEED), device=dev)


KEYS = ("xyz", "mask", "extent", "lfborder", "K4", "cls", "R_gt", "t_gt")


def test_c2_table_is_additive_deterministic_and_complete(dev, c2):
    F = bench.FRAMES_PER_GPU
    sw = _sweep(dev, F)
    out1 = [x.clone() for x in sw.step(*(c2[k] for k in KEYS))]
    whole = sw.finish().clone()
    sw.reset()
    out2 = sw.step(*(c2[k] for k in KEYS))
    assert torch.equal(sw.finish(), whole) and all(torch.equal(a, b) for a, b in zip(out1, out2))   # deterministic
    # every frame counted once, in its own class column; no failures; the GT poses are recovered
    cls = c2["cls"].cpu().numpy()
    assert np.array_equal(whole[ops_metric.ROW_OBJ_NUM].cpu().numpy(), np.bincount(cls, minlength=13).astype(np.float64))
    s = pb.summarize(whole, synth.OBJLIST)
    assert s["frames"] == F and s["failed"] == 0
    assert s["succ_base_rt"] >= 0.97 and s["succ_base_r"] >= 0.99 and s["succ_base_t"] >= 0.99
    # additivity: two half batches (own launches, own grids; frame0 keeps every frame's hypothesis stream) sum to the whole
    half = _sweep(dev, F // 2)
    parts = []
    for h in range(2):
        lo, hi = h * F // 2, (h + 1) * F // 2
        parts.append([x.clone() for x in half.step(*(c2[k][lo:hi] for k in KEYS), frame0=lo)])
    assert torch.equal(half.finish(), whole)
    for j in range(3):
        assert torch.equal(torch.cat([parts[0][j], parts[1][j]]), out1[j])


def test_c2_full_launch_matches_oracle_on_sampled_frames(cref, dev, c2):
    """Frames taken out of the 1024-frame launch (the balanced scoring grid at its benchmark size) against the C
    oracle: compaction output, every hypothesis' inlier count, chosen hypothesis, bitmap; R, t to 1e-4."""
    F = bench.FRAMES_PER_GPU
    pts, count = ops.compact_map(c2["xyz"], c2["mask"], c2["extent"], c2["lfborder"], bench.CAP)
    cfg = ops.RansacConfig(n_hyp=bench.N_HYP, seed=bench.SEED)
    out = ops.pnp_ransac_pts(pts, count, c2["K4"], cfg, want_counts=True, want_bitmap=True)
    sample = [0, 1, 137, 511, 512, 777, 1022, 1023]
    cnt = count.cpu().numpy()
    for f in sample:
        sp = synth.make_frame_sparse(bench.SEED, bench.CONFIG_ID, f)
        ref_pts = synth.compact_points(sp)
        n = int(cnt[f])
        assert n == ref_pts.shape[1] and np.array_equal(pts[f, :, :n].cpu().numpy(), ref_pts), f
        o = cref.ransac_frame(ref_pts, synth.LINEMOD_K4, frame=f, n_hyp=cfg.n_hyp, thresh_px=cfg.thresh_px,
                              gn_iters=cfg.gn_iters, lo_rounds=cfg.lo_rounds, min_inliers=cfg.min_inliers, seed=cfg.seed)
        assert np.array_equal(out["hyp_counts"][f].cpu().numpy(), o["hyp_counts"]), f
        assert int(out["best_hyp"][f]) == o["best_hyp"] and int(out["n_inl"][f]) == o["n_inl"], f
        words = (n + 31) // 32
        assert np.array_equal(out["inl_bitmap"][f, :words].cpu().numpy().view(np.uint32), o["bitmap"][:words]), f
        assert rot_angle(out["R"][f].cpu().numpy(), o["R"]) <= 1e-4 and np.abs(out["t"][f].cpu().numpy() - o["t"]).max() <= 1e-4


def test_c3_full_batch_nearest_neighbour_properties(dev):
    """C3: 4096 problems of 2620 x 2620 points.  Tensor-core and FP32-scan searches agree on all 10.7 M indices;
    searching a cloud in itself returns the identity; a permuted database returns the permuted indices; sampled rows
    equal a float64 brute force; ADD-S of a cloud against itself is 0 and ADD-S <= ADD."""
    B, N = 4096, 2620
    g = torch.Generator(device=dev).manual_seed(33)
    ext = torch.tensor([0.1502, 0.1071, 0.0692], device=dev)
    y = (torch.rand(B, N, 3, device=dev, generator=g) - 0.5) * ext
    x = y + 0.004 * torch.randn(B, N, 3, device=dev, generator=g)
    i_tc, d_tc = ops_metric.knn1(x, y, want_d2=True, impl="tc")
    i_si, d_si = ops_metric.knn1(x, y, want_d2=True, impl="simt")
    assert torch.equal(i_tc, i_si) and torch.equal(d_tc, d_si)
    ident = torch.arange(N, device=dev).expand(B, N)
    i_self, d_self = ops_metric.knn1(y, y, want_d2=True, impl="tc")
    assert torch.equal(i_self, ident) and float(d_self.abs().max()) == 0.0
    perm = torch.randperm(N, device=dev, generator=g)
    i_perm = ops_metric.knn1(x[:64], y[:64][:, perm].contiguous(), impl="tc")
    assert torch.equal(perm[i_perm], i_tc[:64])
    for b in (0, 2047, 4095):
        rows = torch.randint(0, N, (32,), device=dev, generator=g)
        d = ((x[b, rows].double()[:, None, :] - y[b].double()[None, :, :]) ** 2).sum(-1)
        best = d.gather(1, i_tc[b, rows][:, None])[:, 0]
        assert torch.all(best <= d.min(dim=1).values * (1 + 1e-6))          # float32 rounding can swap near-ties only
    add, adds = ops_metric.add_adds(x, y, None)
    assert torch.all(adds <= add * (1 + 1e-6)) and torch.all(adds > 0)
    _, adds0 = ops_metric.add_adds(y[:256], y[:256], None)
    assert float(adds0.abs().max()) == 0.0


def test_c4_dense_frames_do_not_depend_on_the_batch(cref, dev):
    """C4: P = 10 000 correspondences, 1024 hypotheses, 5 re-score / refit rounds.  A frame's result is the same
    whether it runs in a batch of 64 or alone (frame index kept), and equals the oracle; T-vote on 10 000 points."""
    F, cap, H = 64, 10240, 1024
    sps = synth.make_batch_sparse(0x5EED, 4, 4, p_range=(10000, 10000), outlier_frac=0.3, z_range=(0.30, 0.40), xy_frac=0.25)
    pts = np.zeros((F, 5, cap), np.float32); cnt = np.zeros(F, np.int32)
    for f in range(F):
        p = synth.compact_points(sps[f % 4])
        n = min(p.shape[1], cap); pts[f, :, :n] = p[:, :n]; cnt[f] = n
    cfg = ops.RansacConfig(n_hyp=H, lo_rounds=5)
    K4 = torch.from_numpy(np.tile(synth.LINEMOD_K4, (F, 1))).to(dev)
    pts_d, cnt_d = torch.from_numpy(pts).to(dev), torch.from_numpy(cnt).to(dev)
    full = ops.pnp_ransac_pts(pts_d, cnt_d, K4, cfg, want_counts=True)
    for f in (0, 37, 63):
        one = ops.pnp_ransac_pts(pts_d[f:f + 1], cnt_d[f:f + 1], K4[f:f + 1], cfg, frame0=f, want_counts=True)
        for k in ("R", "t", "n_inl", "best_hyp", "hyp_counts", "status", "n_inl_refined"):
            assert torch.equal(one[k][0], full[k][f]), (f, k)
    f = 37
    o = cref.ransac_frame(pts[f, :, :cnt[f]], synth.LINEMOD_K4, frame=f, n_hyp=H, thresh_px=cfg.thresh_px,
                          gn_iters=cfg.gn_iters, lo_rounds=5, min_inliers=cfg.min_inliers, seed=cfg.seed)
    assert np.array_equal(full["hyp_counts"][f].cpu().numpy(), o["hyp_counts"])
    assert int(full["best_hyp"][f]) == o["best_hyp"] and int(full["n_inl_refined"][f]) == o["n_inl_refined"]
    assert rot_angle(full["R"][f].cpu().numpy(), o["R"]) <= 1e-4 and np.abs(full["t"][f].cpu().numpy() - o["t"]).max() <= 1e-4
    NP = 10000
    g = torch.Generator(device=dev).manual_seed(4)
    cloud, tres = torch.randn(F, NP, 3, device=dev, generator=g), torch.randn(F, 3, NP, device=dev, generator=g)
    v = ops_metric.vote_translation(cloud, tres)
    ref = (cloud.double() + tres.double().permute(0, 2, 1)).mean(dim=1)
    assert torch.allclose(v.double(), ref, atol=2e-6)
