"""K3 with the projection on the tensor cores -- csrc/ransac_score_tc.cu (tcgen05 + TMEM, impl "tc") and
csrc/ransac_score_mma.cu (warp-level mma.sync, impl "mma"), both opt-in through pb_ransac_score_impl -- against the
FP32-pipe kernels and the C oracle: every integer (inlier counts of every hypothesis, chosen hypothesis, bitmap, status) and the
pose must be IDENTICAL -- the tensor-core values only pre-sort the tests, the contract's arithmetic decides inside the
band.  Shapes: two M tiles (256 hypotheses), one (96), a partial second one (200); ragged / empty / too-small /
non-finite frames; another threshold; a narrower band must still be exact on these data, and the kernel must report
no barrier time-out."""
import numpy as np
import pytest
import torch

from pose_estimation_b200 import ops, synth
from tests.util import pack_pts

pytestmark = pytest.mark.gpu

KEYS = ("hyp_counts", "best_hyp", "n_inl", "n_inl_refined", "status", "inl_bitmap", "R", "t")


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def cref():
    from oracle import cref as c
    return c


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _both(pts, cnt, K4, cfg, dev, frame0=0, band_scale=1.0, tc="tc"):
    F, _, cap = pts.shape
    outs = {}
    try:
        for impl in ("fp32", tc):
            ops.set_score_impl(impl, band_scale)
            pr = ops.PtsRansac(F, cap, cfg, device=dev)            # its workspace holds the kernel's statistics
            o = ops.pnp_ransac_pts(_t(pts, dev), _t(cnt, dev), _t(K4, dev), cfg, frame0=frame0, want_counts=True, want_bitmap=True)
            outs[impl] = {k: v.cpu().numpy() for k, v in o.items()}
            if impl == tc:
                pr(_t(pts, dev), _t(cnt, dev), _t(K4, dev), frame0=frame0)
                st = ops.ransac_tc_stats(pr.ws, F, cap, cfg)
                assert st is not None, "the shape did not take the tensor-core kernel"
                assert st["timeout"] == 0
                outs["stats"] = st
    finally:
        ops.set_score_impl("auto", 1.0)
    outs["tc"] = outs[tc]
    for k in KEYS:
        assert np.array_equal(outs["fp32"][k], outs["tc"][k], equal_nan=True), (tc, k)
    return outs


def _frames(F):
    base = [synth.compact_points(s) for s in synth.make_batch_sparse(0x5EED, 2, 16)]
    base[3] = base[3][:, :3]; base[9] = base[9][:, :0]; base[12] = base[12][:, :200]; base[5] = base[5][:, :64]; base[6] = base[6][:, :65]
    return (base * ((F + 15) // 16))[:F]


@pytest.mark.parametrize("tc", ["tc", "mma"])
@pytest.mark.parametrize("n_hyp,F,thr", [(256, 320, 1.0), (96, 300, 1.0), (200, 304, 3.0), (40, 300, 1.0)])
def test_tc_scoring_equals_fp32_and_oracle(cref, dev, n_hyp, F, thr, tc):
    cap = 1280
    pts, cnt = pack_pts(_frames(F), cap)
    K4 = np.tile(synth.LINEMOD_K4, (F, 1))
    cfg = ops.RansacConfig(n_hyp=n_hyp, thresh_px=thr, lo_rounds=1)
    outs = _both(pts, cnt, K4, cfg, dev, frame0=5, tc=tc)
    assert outs["stats"]["reevaluated_groups"] > 0                 # the band is exercised
    for f in (0, 3, 5, 6, 9, 12, 17, F - 1):
        if cnt[f] < 4:
            continue
        o = cref.ransac_frame(pts[f, :, :cnt[f]], K4[f], frame=5 + f, n_hyp=n_hyp, thresh_px=thr, gn_iters=cfg.gn_iters,
                              lo_rounds=1, min_inliers=cfg.min_inliers, seed=cfg.seed)
        assert np.array_equal(outs["tc"]["hyp_counts"][f], o["hyp_counts"]), f
        assert outs["tc"]["best_hyp"][f] == o["best_hyp"] and outs["tc"]["n_inl"][f] == o["n_inl"], f


@pytest.mark.parametrize("tc", ["tc", "mma"])
def test_tc_scoring_ragged_and_degenerate_frames(dev, tc):
    cap = 1280
    base = [synth.compact_points(s) for s in synth.make_batch_sparse(0x5EED, 2, 6)]
    frames = [base[0][:, :n].copy() for n in (0, 1, 3, 4, 5, 31, 32, 33, 63, 64, 65, 127, 128, 129)]
    frames.append(np.concatenate([base[1], base[2]], axis=1)[:, :cap])          # exactly cap correspondences
    frames.append(np.repeat(base[3][:, :1], 200, axis=1))                       # 200 copies of one correspondence
    col = base[4][:, :300].copy()
    tl = np.linspace(-1, 1, 300, dtype=np.float32)
    col[0], col[1], col[2] = 0.01 * tl, 0.02 * tl, -0.015 * tl                  # collinear model points
    frames.append(col)
    behind = base[5][:, :400].copy(); behind[2] -= 5.0; frames.append(behind)
    bad = base[0][:, :500].copy()
    bad[0, ::7] = np.nan; bad[1, 3::11] = np.inf; bad[3, 5::13] = -np.inf       # non-finite entries among good points
    frames.append(bad)
    huge = base[1][:, :300].copy(); huge[:3] *= 3.0e5; frames.append(huge)      # coordinates beyond the FP16 scaling range
    tiny = base[2][:, :300].copy(); tiny[:3] *= 1.0e-7; frames.append(tiny)
    frames += [base[1], base[2]]
    frames = (frames * 14)[:308]
    pts, cnt = pack_pts(frames, cap)
    K4 = np.tile(synth.LINEMOD_K4, (len(frames), 1))
    for n_hyp in (256, 64):
        _both(pts, cnt, K4, ops.RansacConfig(n_hyp=n_hyp), dev, frame0=100, tc=tc)


@pytest.mark.parametrize("tc", ["tc", "mma"])
def test_tc_scoring_narrower_band_still_exact_and_cheaper(dev, tc):
    """Half the derived band is still exact on these data (the calibration pass measures the largest deviation at about
    an eighth of the bound) and re-evaluates fewer groups."""
    cap = 1280
    F = 320
    pts, cnt = pack_pts(_frames(F), cap)
    K4 = np.tile(synth.LINEMOD_K4, (F, 1))
    cfg = ops.RansacConfig()
    full = _both(pts, cnt, K4, cfg, dev, tc=tc)["stats"]["reevaluated_groups"]
    half = _both(pts, cnt, K4, cfg, dev, band_scale=0.5, tc=tc)["stats"]["reevaluated_groups"]
    assert half < full
    try:
        ops.set_score_impl(tc + "_calib", 1.0)
        pr = ops.PtsRansac(F, cap, cfg, device=dev)
        pr(_t(pts, dev), _t(cnt, dev), _t(K4, dev))
        st = ops.ransac_tc_stats(pr.ws, F, cap, cfg)
    finally:
        ops.set_score_impl("auto", 1.0)
    assert st["timeout"] == 0 and st["calib_mismatch"] == 0
    assert 0.0 < st["calib_max_ratio"] < 0.5, st                   # the bound holds with a factor of two to spare
