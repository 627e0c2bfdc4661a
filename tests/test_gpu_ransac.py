"""GPU parity: K0 compaction and K1-K4 PnP-RANSAC against the plain-C oracle (bit-exact for every
integer/index result; R,t within 1e-4 rad / 1e-4 m as BASELINE.json's north_star states)."""
import numpy as np
import pytest
import torch

from pose_estimation_b200 import ops, synth
from tests.util import batch_arrays, pack_pts, rot_angle

pytestmark = pytest.mark.gpu
R_TOL_RAD, T_TOL_M = 1e-4, 1e-4


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


@pytest.mark.parametrize("F,use_mask", [(3, True), (3, False), (40, True)])
def test_compact_matches_oracle(cref, dev, F, use_mask):
    sps = synth.make_batch_sparse(0x5EED, 11, F)
    b = batch_arrays(sps)
    cap = 1280
    pts, count, total = ops.compact_map(_t(b["xyz"], dev), _t(b["mask"], dev) if use_mask else None,
                                        _t(b["extent"], dev), _t(b["lfborder"], dev), cap, want_total=True)
    pts, count, total = pts.cpu().numpy(), count.cpu().numpy(), total.cpu().numpy()
    for f in range(F):
        ref, tot = cref.compact_map(b["xyz"][f], b["mask"][f] if use_mask else None, b["extent"][f],
                                    b["lfborder"][f], cap=cap)
        assert count[f] == ref.shape[1] and total[f] == tot
        assert np.array_equal(pts[f, :, :count[f]], ref)


def test_compact_truncation_roi_and_odd_shape(cref, dev):
    rng = np.random.default_rng(3)
    # odd H*W (scalar path), dense mask, small cap, ROI offsets
    F, H, W, cap = 5, 37, 53, 200
    xyz = rng.random((F, 3, H, W)).astype(np.float32)
    mask = (rng.random((F, H, W)) < 0.4).astype(np.uint8)
    mask[1] = 0                      # empty frame
    mask[2] = 1                      # full frame -> truncated at cap
    ext = rng.random((F, 3)).astype(np.float32) + 0.1
    lfb = -rng.random((F, 3)).astype(np.float32)
    roi = rng.integers(0, 100, (F, 2)).astype(np.int32)
    for m in (mask, None):
        pts, count, total = ops.compact_map(_t(xyz, dev), None if m is None else _t(m, dev), _t(ext, dev),
                                            _t(lfb, dev), cap, roi=_t(roi, dev), want_total=True)
        pts, count, total = pts.cpu().numpy(), count.cpu().numpy(), total.cpu().numpy()
        for f in range(F):
            ref, tot = cref.compact_map(xyz[f], None if m is None else m[f], ext[f], lfb[f],
                                        rmin=int(roi[f, 0]), cmin=int(roi[f, 1]), cap=cap)
            assert count[f] == ref.shape[1] and total[f] == tot
            assert np.array_equal(pts[f, :, :count[f]], ref)
    # vector path with H*W % 4 == 0 and several row bands per frame (few frames)
    F, H, W, cap = 2, 480, 640, 5000
    xyz = np.zeros((F, 3, H, W), np.float32)
    mask = (rng.random((F, H, W)) < 0.02).astype(np.uint8)
    xyz[:, :, :, :] = rng.random((F, 3, H, W)).astype(np.float32) * mask[:, None]
    ext = np.ones((F, 3), np.float32); lfb = np.zeros((F, 3), np.float32)
    pts, count, total = ops.compact_map(_t(xyz, dev), _t(mask, dev), _t(ext, dev), _t(lfb, dev), cap, want_total=True)
    for f in range(F):
        ref, tot = cref.compact_map(xyz[f], mask[f], ext[f], lfb[f], cap=cap)
        assert int(count[f]) == ref.shape[1] and int(total[f]) == tot
        assert np.array_equal(pts[f, :, :ref.shape[1]].cpu().numpy(), ref)


def test_compact_mask_first_kernel_edges_and_pinned_host(cref, dev):
    """Mask-first (sparse) K0: any non-zero mask byte is foreground, dense masks truncate at cap in
    pixel order, ROI offsets, empty frames; the same kernel reading PINNED HOST maps (zero-copy
    over PCIe) is bit-identical."""
    rng = np.random.default_rng(7)
    F, H, W, cap = 6, 64, 96, 300
    xyz = rng.random((F, 3, H, W)).astype(np.float32)
    mask = (rng.random((F, H, W)) < 0.03).astype(np.uint8) * rng.integers(1, 256, (F, H, W)).astype(np.uint8)
    mask[1] = 0                       # empty
    mask[2] = 255                     # full frame -> truncated at cap
    mask[3] = 0; mask[3, -1, -1] = 7  # only the very last pixel
    mask[4] = 0; mask[4, 0, 0] = 128; mask[4, 0, 31] = 1; mask[4, 0, 32] = 1
    ext = rng.random((F, 3)).astype(np.float32) + 0.1
    lfb = -rng.random((F, 3)).astype(np.float32)
    roi = rng.integers(0, 100, (F, 2)).astype(np.int32)
    refs = [cref.compact_map(xyz[f], mask[f], ext[f], lfb[f], rmin=int(roi[f, 0]), cmin=int(roi[f, 1]), cap=cap)
            for f in range(F)]
    xyz_p, mask_p = torch.from_numpy(xyz).pin_memory(), torch.from_numpy(mask).pin_memory()
    comp = ops.MapCompactor(F, H, W, cap, device=dev)
    assert comp.launches(True) == 1
    for src_xyz, src_mask in ((_t(xyz, dev), _t(mask, dev)), (xyz_p, mask_p)):
        pts, count = comp(src_xyz, src_mask, _t(ext, dev), _t(lfb, dev), roi=_t(roi, dev))
        pts, count = pts.cpu().numpy(), count.cpu().numpy()
        for f in range(F):
            ref, tot = refs[f]
            assert count[f] == ref.shape[1] == min(tot, cap)
            assert np.array_equal(pts[f, :, :count[f]], ref)
    with pytest.raises(Exception):    # pageable host memory is rejected (no CPU path)
        comp(torch.from_numpy(xyz), _t(mask, dev), _t(ext, dev), _t(lfb, dev))


def _check_ransac(cref, dev, sps, cfg, cap, frame0=0, sample_idx=None):
    pts_list = [synth.compact_points(s) for s in sps]
    pts, cnt = pack_pts(pts_list, cap)
    F = len(sps)
    K4 = np.tile(synth.LINEMOD_K4, (F, 1))
    out = ops.pnp_ransac_pts(_t(pts, dev), _t(cnt, dev), _t(K4, dev), cfg, frame0=frame0,
                             sample_idx=None if sample_idx is None else _t(sample_idx, dev),
                             want_counts=True, want_bitmap=True, want_minimal=True)
    out = {k: v.cpu().numpy() for k, v in out.items()}
    for f in range(F):
        o = cref.ransac_frame(pts[f, :, :cnt[f]], K4[f], frame=frame0 + f,
                              sample_idx=None if sample_idx is None else sample_idx[f],
                              n_hyp=cfg.n_hyp, thresh_px=cfg.thresh_px, gn_iters=cfg.gn_iters,
                              lo_rounds=cfg.lo_rounds, min_inliers=cfg.min_inliers, seed=cfg.seed)
        assert out["status"][f] == o["status"], f
        assert np.array_equal(out["hyp_counts"][f], o["hyp_counts"]), f      # bit-exact inlier counts
        assert out["best_hyp"][f] == o["best_hyp"] and out["n_inl"][f] == o["n_inl"], f
        words = (cnt[f] + 31) // 32
        assert np.array_equal(out["inl_bitmap"][f, :words].view(np.uint32), o["bitmap"][:words]), f
        assert np.array_equal(out["R0"][f], o["R0"]) and np.array_equal(out["t0"][f], o["t0"]), f
        assert rot_angle(out["R"][f], o["R"]) <= R_TOL_RAD, f
        assert np.abs(out["t"][f] - o["t"]).max() <= T_TOL_M, f
        assert out["n_inl_refined"][f] == o["n_inl_refined"], f
    return out


@pytest.mark.parametrize("n_hyp,F,lo", [(256, 6, 0), (256, 6, 2), (64, 3, 0), (100, 2, 1), (1024, 2, 0),
                                        (512, 300, 0)])
def test_ransac_bit_exact_vs_oracle(cref, dev, n_hyp, F, lo):
    nf = min(F, 12)
    sps = synth.make_batch_sparse(0x5EED, 2, nf)
    sps = (sps * ((F + nf - 1) // nf))[:F]
    cfg = ops.RansacConfig(n_hyp=n_hyp, lo_rounds=lo)
    _check_ransac(cref, dev, sps[:F] if F <= 12 else sps, cfg, cap=1280, frame0=7)


@pytest.mark.parametrize("lo", [0, 2])
def test_ransac_warp_per_frame_finalize_matches_oracle_and_cta_path(cref, dev, lo):
    """F >= 592 frames take the warp-per-frame K4 kernel; fewer take the CTA-per-frame one.  Both
    must agree with the oracle (integers bit-exact) and with EACH OTHER bit for bit (R, t included):
    results may not depend on launch geometry."""
    F = 640
    base = synth.make_batch_sparse(0x5EED, 2, 16)
    base[3] = dict(base[3]); base[3]["rows"] = base[3]["rows"][:3]; base[3]["cols"] = base[3]["cols"][:3]
    base[3]["xyzn"] = base[3]["xyzn"][:3]                          # a too-few-points frame in the mix
    sps = (base * (F // 16))[:F]
    cfg = ops.RansacConfig(n_hyp=64, lo_rounds=lo)
    big = _check_ransac(cref, dev, sps, cfg, cap=1280, frame0=11)
    small = _check_ransac(cref, dev, sps[:40], cfg, cap=1280, frame0=11)
    for k in ("R", "t", "n_inl_refined", "status", "best_hyp", "inl_bitmap"):
        assert np.array_equal(big[k][:40], small[k]), k


@pytest.mark.parametrize("n_hyp,F", [(256, 320), (128, 300)])
def test_ransac_balanced_scoring_matches_oracle_and_frame_per_cta_path(cref, dev, n_hyp, F):
    """296 <= F <= 2368 with H <= 256 takes the balanced scoring kernel (CTAs own equal ranges of the batch's
    point list, counts by integer atomics, arg-max in K4).  Ragged frames, an empty one and a too-few-points one;
    the oracle and the frame-per-CTA path (same frames in a small batch) must be reproduced bit for bit."""
    base = synth.make_batch_sparse(0x5EED, 2, 16)
    for j, keep in ((3, 3), (9, 0), (12, 200)):
        base[j] = dict(base[j])
        for k in ("rows", "cols", "xyzn"):
            base[j][k] = base[j][k][:keep]
    sps = (base * ((F + 15) // 16))[:F]
    cfg = ops.RansacConfig(n_hyp=n_hyp, lo_rounds=1)
    big = _check_ransac(cref, dev, sps, cfg, cap=1280, frame0=5)
    small = _check_ransac(cref, dev, sps[:48], cfg, cap=1280, frame0=5)
    for k in ("R", "t", "n_inl", "n_inl_refined", "status", "best_hyp", "inl_bitmap", "hyp_counts"):
        assert np.array_equal(big[k][:48], small[k]), k


def test_ransac_forced_minimal_sets_and_large_P(cref, dev):
    sps = synth.make_batch_sparse(0x5EED, 4, 2, p_range=(5000, 6000), outlier_frac=0.3)
    cap = 6144
    rng = np.random.default_rng(5)
    H = 128
    sidx = np.stack([rng.integers(0, s["rows"].shape[0], (H, 4)) for s in sps]).astype(np.int32)
    sidx[0, 3] = [-1, 2, 3, 4]                   # invalid index -> hypothesis dropped
    sidx[1, 5] = [9, 9, 10, 11]                  # duplicate point -> degenerate
    _check_ransac(cref, dev, sps, ops.RansacConfig(n_hyp=H, lo_rounds=1), cap=cap, sample_idx=sidx)


def test_ransac_failure_status(dev, cref):
    F, cap = 3, 64
    pts = np.zeros((F, 5, cap), np.float32)
    cnt = np.array([0, 3, 40], np.int32)
    rng = np.random.default_rng(0)
    pts[2, :, :40] = rng.random((5, 40))         # garbage correspondences: no consensus
    pts[2, 3:5, :40] *= 600.0                    # image points spread over the frame (not all within 1 px)
    K4 = np.tile(synth.LINEMOD_K4, (F, 1))
    out = ops.pnp_ransac_pts(_t(pts, dev), _t(cnt, dev), _t(K4, dev), ops.RansacConfig(n_hyp=32, min_inliers=30))
    st = out["status"].cpu().numpy()
    assert st[0] == 1 and st[1] == 1 and st[2] == 2
    o = cref.ransac_frame(pts[2, :, :40], K4[2], frame=2, n_hyp=32, min_inliers=30)
    assert o["status"] == 2
    assert np.array_equal(out["R"][0].cpu().numpy(), np.eye(3, dtype=np.float32))
    assert np.all(out["t"].cpu().numpy()[:2] == 0)


def test_map_ransac_equals_two_step(dev):
    sps = synth.make_batch_sparse(0x5EED, 2, 5)
    b = batch_arrays(sps)
    cap = 1280
    cfg = ops.RansacConfig()
    mr = ops.MapRansac(5, synth.IMG_H, synth.IMG_W, cap, cfg, device=dev)
    mr(_t(b["xyz"], dev), _t(b["mask"], dev), _t(b["extent"], dev), _t(b["lfborder"], dev), _t(b["K4"], dev))
    pts, count = ops.compact_map(_t(b["xyz"], dev), _t(b["mask"], dev), _t(b["extent"], dev), _t(b["lfborder"], dev), cap)
    out = ops.pnp_ransac_pts(pts, count, _t(b["K4"], dev), cfg)
    assert torch.equal(mr.R, out["R"]) and torch.equal(mr.t, out["t"]) and torch.equal(mr.n_inl, out["n_inl"])
    # pose quality against ground truth (1 mm noise, 25 % outliers)
    for f in range(5):
        assert rot_angle(mr.R[f].cpu().numpy(), b["R_gt"][f]) < np.radians(5.0)
        assert np.linalg.norm(mr.t[f].cpu().numpy() - b["t_gt"][f]) < 0.05


def test_ransac_ragged_and_degenerate_batches_match_oracle(cref, dev):
    """One launch holding every awkward frame at once: 0 / 1 / 3 / 4 / 5 / 31 / 32 / 33 correspondences, exactly cap,
    all points identical, collinear model points, points behind the camera, NaN / Inf coordinates, and normal frames
    between them.  Status, every hypothesis' inlier count, chosen hypothesis, bitmap and pose equal the C oracle's."""
    cap = 1280
    base = [synth.compact_points(s) for s in synth.make_batch_sparse(0x5EED, 2, 6)]
    rng = np.random.default_rng(12)
    frames = []
    for n in (0, 1, 3, 4, 5, 31, 32, 33):
        frames.append(base[0][:, :n].copy())
    full = np.concatenate([base[1], base[2]], axis=1)[:, :cap]                 # exactly cap correspondences (two objects mixed)
    frames.append(full)
    same = np.repeat(base[3][:, :1], 200, axis=1); frames.append(same)          # 200 copies of one correspondence
    col = base[4][:, :300].copy()
    tline = np.linspace(-1, 1, 300, dtype=np.float32)
    col[0], col[1], col[2] = 0.01 * tline, 0.02 * tline, -0.015 * tline          # collinear model points
    frames.append(col)
    behind = base[5][:, :400].copy(); behind[2] -= 5.0; frames.append(behind)   # far outside the object: no consensus
    bad = base[0][:, :500].copy()
    bad[0, ::7] = np.nan; bad[1, 3::11] = np.inf; bad[3, 5::13] = -np.inf       # non-finite entries among good points
    frames.append(bad)
    frames += [base[1], base[2]]                                                # ordinary frames after the odd ones
    pts, cnt = pack_pts(frames, cap)
    F = len(frames)
    K4 = np.tile(synth.LINEMOD_K4, (F, 1))
    for n_hyp, lo in ((256, 0), (96, 2)):
        cfg = ops.RansacConfig(n_hyp=n_hyp, lo_rounds=lo)
        out = ops.pnp_ransac_pts(_t(pts, dev), _t(cnt, dev), _t(K4, dev), cfg, frame0=100, want_counts=True, want_bitmap=True)
        out = {k: v.cpu().numpy() for k, v in out.items()}
        for f in range(F):
            o = cref.ransac_frame(pts[f, :, :cnt[f]], K4[f], frame=100 + f, n_hyp=cfg.n_hyp, thresh_px=cfg.thresh_px,
                                  gn_iters=cfg.gn_iters, lo_rounds=cfg.lo_rounds, min_inliers=cfg.min_inliers, seed=cfg.seed)
            assert out["status"][f] == o["status"], (f, cnt[f])
            if cnt[f] >= 4:
                assert np.array_equal(out["hyp_counts"][f], o["hyp_counts"]), f
            if o["status"] == 0:
                assert out["best_hyp"][f] == o["best_hyp"] and out["n_inl"][f] == o["n_inl"], f
                words = (cnt[f] + 31) // 32
                assert np.array_equal(out["inl_bitmap"][f, :words].view(np.uint32), o["bitmap"][:words]), f
                assert rot_angle(out["R"][f], o["R"]) <= R_TOL_RAD and np.abs(out["t"][f] - o["t"]).max() <= T_TOL_M, f
            else:
                assert np.array_equal(out["R"][f], np.eye(3, dtype=np.float32)) and np.all(out["t"][f] == 0), f
        assert np.all(np.isfinite(out["R"])) and np.all(np.isfinite(out["t"]))
    assert list(out["status"][:3]) == [1, 1, 1]                                 # fewer than 4 correspondences


@pytest.mark.parametrize("cap", [40000, 6000])
def test_compact_dense_silhouettes_across_tiles(cref, dev, cap):
    """Dense, compact masks shaped like real LineMOD silhouettes (every pixel of the ellipsoid, 2 k .. 30 k px)
    through the streaming mask-first kernel: blobs span several 16 K-pixel tiles of a frame, so the per-tile
    ranks, the look-back over the earlier tiles and the cut at ``cap`` in the middle of a tile are all exercised;
    many frames so that the persistent grid takes several tickets per CTA."""
    F = 24
    sps = [synth.make_frame_sparse(0x5EED, 12, f, dense=True, z_range=(0.45, 1.0)) for f in range(F)]
    assert max(s["rows"].shape[0] for s in sps) > 16384 > min(s["rows"].shape[0] for s in sps)
    b = batch_arrays(sps)
    comp = ops.MapCompactor(F, synth.IMG_H, synth.IMG_W, cap, device=dev)
    for rep in range(2):                                   # the workspace (ticket, tile states) is reset by every call
        pts, count = comp(_t(b["xyz"], dev), _t(b["mask"], dev), _t(b["extent"], dev), _t(b["lfborder"], dev))
        pts_h, count_h = pts.cpu().numpy(), count.cpu().numpy()
        for f in range(F):
            ref, tot = cref.compact_map(b["xyz"][f], b["mask"][f], b["extent"][f], b["lfborder"][f], cap=cap)
            assert tot == sps[f]["rows"].shape[0] and count_h[f] == ref.shape[1] == min(tot, cap), f
            assert np.array_equal(pts_h[f, :, :count_h[f]], ref), f
    pts2, count2, total2 = ops.compact_map(_t(b["xyz"], dev), _t(b["mask"], dev), _t(b["extent"], dev),
                                           _t(b["lfborder"], dev), cap, want_total=True)
    assert torch.equal(count2, count) and np.array_equal(total2.cpu().numpy(), [s["rows"].shape[0] for s in sps])
