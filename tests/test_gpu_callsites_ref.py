"""GPU parity through the reference-shaped API against outputs of the REFERENCE'S OWN FUNCTIONS
(tests/golden/callsites_ref.npz: Trainer.get_pose / cal_dis, lib/utils/metric.py, lib/utils/utlis.py and
PoseLoss from /root/reference, run in the build container by tests/golden/gen_golden_callsites.py)."""
import os

import numpy as np
import pytest
import torch

import pose_estimation_b200 as pb
from pose_estimation_b200 import synth
from tests.test_gpu_api import _datas
from tests.test_oracle_golden import GOLD, load_frames
from tests.util import rot_angle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLD, "callsites_ref.npz"))


def test_get_pose_vs_reference_get_pose(g):
    """tools/trainer.py:383-438.  The reference solves on 256 random pixels with cv2; here every masked pixel is
    scored.  Clear-consensus frames: R within 1e-4 rad, t within 1e-4 m (north_star's bar).  Noisy frames:
    ADD(-S) / 5 degree / 5 cm verdicts identical (tools/trainer.py:186-191)."""
    models = synth.class_models(500)
    metric = pb.Metric(synth.SYM_CLS)
    for tag in ("clean", "noisy"):
        for f, sp in enumerate(load_frames(g, tag)):
            pred, data = _datas(sp, models)
            R, t = pb.get_pose(pred, data)
            if tag == "clean":
                assert rot_angle(R[0].numpy(), g["clean_R"][f]) <= 1e-4
                assert np.abs(t[0].numpy() - g["clean_t"][f]).max() <= 1e-4
            add, r_dis, t_dis = pb.cal_dis(metric, R, t, data)
            ref_add, ref_r, ref_t = g[f"{tag}_dis"][f]
            d = synth.DIAMETER[sp["cls"]]
            assert [add < 0.1 * d, r_dis < 5.0, t_dis < 0.05] == [ref_add < 0.1 * d, ref_r < 5.0, ref_t < 0.05]


def test_cal_dis_vs_reference_cal_dis(g):
    """tools/trainer.py:370-381 on the reference's own (R, t): same three numbers."""
    models = synth.class_models(500)
    metric = pb.Metric(synth.SYM_CLS)
    for tag in ("clean", "noisy"):
        for f, sp in enumerate(load_frames(g, tag)):
            _, data = _datas(sp, models)
            R, t = torch.from_numpy(g[f"{tag}_R"][f])[None], torch.from_numpy(g[f"{tag}_t"][f])[None]
            add, r_dis, t_dis = pb.cal_dis(metric, R, t, data)
            ref_add, ref_r, ref_t = g[f"{tag}_dis"][f]
            assert abs(add - ref_add) <= 1e-5 * ref_add + 1e-7
            assert abs(t_dis - ref_t) <= 1e-5 * ref_t + 1e-7
            # float32 acos near 1 in the reference: quantised in ~0.03 degree steps below 0.1 degree
            assert abs(r_dis - ref_r) < 0.06


def test_metric_vs_reference_metric(g, dev):
    """lib/utils/metric.py:17-36,67-84."""
    for cls in (7, 8, 3):
        model, R, t, Rg, tg = (torch.from_numpy(g[f"m{cls}_{k}"]).to(dev) for k in ("model", "R", "t", "Rg", "tg"))
        pred, tgt = model @ R.T + t, model @ Rg.T + tg
        vals = g[f"m{cls}_vals"]
        add, adds = pb.Metric([7, 8]).cal_adds_cuda(pred, tgt, cls)
        assert add == adds
        assert abs(add - vals[0]) <= 1e-5 * vals[0] and abs(adds - vals[1]) <= 1e-5 * vals[1]
        assert abs(pb.Metric.angular_distance(R[None], Rg[None]).item() - vals[2]) < 0.02
        assert abs(pb.Metric.translation_distance(t[None], tg[None]).item() - vals[3]) <= 1e-6
    assert abs(pb.Metric([7, 8]).cal_auc(list(g["auc_in"])) - float(g["auc_out"])) < 1e-9


def test_utlis_vs_reference_utlis(g):
    """lib/utils/utlis.py:12-34,55-74.  choose_ransac_batch: the reference's index list is a random subset of
    cv2's inlier set; ours is a random subset of OUR inlier set (every point scored)."""
    K33 = pb.batch_intrinsic_transform(torch.from_numpy(g["K4"]))
    assert np.array_equal(K33.numpy(), g["K33"])
    assert np.array_equal(pb.batch_intrinsic_transform(K33).numpy(), g["K4_back"])
    coor, uv, Ks = g["crb_coor"], g["crb_uv"], g["crb_K"]
    out = pb.choose_ransac_batch(coor, uv, Ks, n=125)
    assert out.shape == g["crb_out125"].shape and out.dtype == g["crb_out125"].dtype
    wrap = pb.choose_ransac_batch(coor, uv, Ks, n=400)       # fewer than n inliers everywhere: the wrap-pad branch
    assert wrap.shape == (4, 1, 400)                         # (the reference cannot stack these, see the generator)
    for j in range(4):
        cv2_inliers = np.unique(g[f"crb_wrap{j}"])           # the wrap-padded list holds cv2's whole inlier set
        ours = np.unique(wrap[j])                             # our whole inlier set (hypotheses are keyed by batch index)
        assert np.array_equal(wrap[j, 0], np.resize(ours, 400))                     # np.pad(..., 'wrap')
        if j < 2:       # noise-free elements (outliers >= 5 px off): the two inlier sets are the same set
            assert np.array_equal(ours, cv2_inliers)
        else:           # 1 mm noise is ~1 px here, the threshold: both keep the inliers of their best minimal model,
            #             65-80 % of the true ones, different models -> both inside the truth, of comparable size
            truth = g[f"crb_truth{j}"]
            assert np.isin(ours, truth).all() and np.isin(cv2_inliers, truth).all()
            assert ours.size >= 0.85 * cv2_inliers.size
        sel = out[j, 0]
        assert np.all(np.diff(sel) > 0) and np.isin(sel, ours).all()                # order-preserving subset


def test_pose_loss_vs_reference_pose_loss(g, dev):
    """lib/network/loss.py:18-41 forward and autograd."""
    R = torch.from_numpy(g["pl_R"]).to(dev).requires_grad_(True)
    t = torch.from_numpy(g["pl_t"]).to(dev).requires_grad_(True)
    loss = pb.PoseLoss([7, 8])(R, t, torch.from_numpy(g["pl_target"]).to(dev), torch.from_numpy(g["pl_model"]).to(dev),
                               torch.from_numpy(g["pl_idxs"]))
    loss.backward()
    assert abs(loss.item() - float(g["pl_loss"])) <= 1e-5 * float(g["pl_loss"])
    assert np.abs(R.grad.cpu().numpy() - g["pl_grad_R"]).max() <= 1e-4 * np.abs(g["pl_grad_R"]).max()
    assert np.abs(t.grad.cpu().numpy() - g["pl_grad_t"]).max() <= 1e-4 * np.abs(g["pl_grad_t"]).max()


def test_transparent_metric_vs_reference(g, dev):
    """version/transparent/lib/utils.py:17-46 (the older branch's Metric)."""
    M = pb.TransparentMetric([7, 8])
    for cls in (7, 8, 3):
        model, R, t, Rg, tg = (torch.from_numpy(g[f"m{cls}_{k}"]).to(dev) for k in ("model", "R", "t", "Rg", "tg"))
        pred, tgt = model @ R.T + t, model @ Rg.T + tg
        got = [*M.cal_dis(pred, tgt, cls), *M.cal_adds_cuda(pred, tgt, cls)]
        assert np.allclose(got, g[f"m{cls}_tvals"], rtol=1e-5, atol=0)


def test_model_ingestion_sampling(cref, dev, tmp_path):
    """tools/script/sample_model.py:78-85 through pose_estimation_b200.models: ply -> farthest point samples ->
    pickle -> ModelBank, against the C restatement of the reference's selection loop."""
    import pickle
    from pose_estimation_b200 import models
    from tests.test_host_cpu import _write_ply
    rng = np.random.default_rng(5)
    root = tmp_path
    (root / "models").mkdir()
    clouds = {}
    for o in (1, 10):
        clouds[o] = (rng.normal(size=(3000, 3)) * 40).astype(np.float32)          # millimetres
        _write_ply(root / "models" / f"obj_{o:02d}.ply", clouds[o], binary=(o == 10))
    models.sample_models(str(root), objlist=(1, 10), n_samples=700)
    for o in (1, 10):
        with open(root / "models" / f"obj_{o:02d}.pkl", "rb") as f:
            got = pickle.load(f)
        assert np.array_equal(got, clouds[o][cref.fps(clouds[o], 700)])
    info = {o: dict(diameter=0.1, lfborder=np.zeros(3), extent=np.ones(3)) for o in (1, 10)}
    yml = root / "info.yml"
    yml.write_text("".join(f"{o}: {{diameter: 100.0, min_x: 0, min_y: 0, min_z: 0, size_x: 1000, size_y: 1000, size_z: 1000}}\n"
                           for o in (1, 10)))
    bank = models.ModelBank.from_directory(str(root), str(yml), objlist=(1, 10), sym_cls=(1,), num_points=500)
    assert bank.model.shape == (2, 500, 3) and np.array_equal(bank.is_sym, [0, 1])
    assert np.allclose(bank.diameter, 0.1) and np.allclose(bank.extent, 1.0)
    assert np.abs(bank.model).max() < 0.3                                          # metres


def test_region_coded_maps_vs_reference(g, dev):
    """Dense decode against the reference's own runs: arg-max form bit for bit (tools/script/eval.py:93-103), softmax
    form to 1e-6 (lib/network/loss_utils.py:18-28; torch's softmax sums in another order)."""
    from pose_estimation_b200 import ops
    off, lg, rp = (torch.from_numpy(g[k]).to(dev) for k in ("reg_xyz_off", "reg_logits", "reg_points"))
    hard = ops.region_decode_map(off, lg, rp, mode="hard")
    assert np.array_equal(hard.cpu().numpy(), g["reg_hard"])
    soft = pb.get_xyz(off, lg, rp)
    assert np.abs(soft.cpu().numpy() - g["reg_soft"]).max() <= 1e-6
    # compacted form: only the listed pixels are decoded; same numbers as the dense map followed by the plain gather
    bs, _, h, w = off.shape
    gen = torch.Generator(device=dev).manual_seed(3)
    choose = torch.stack([torch.randperm(h * w, device=dev, generator=gen)[:200] for _ in range(bs)])
    xm = (choose % w).float() + 300.0
    ym = (choose // w).float() + 200.0
    ext = torch.tensor([[0.0759, 0.0776, 0.0918], [0.2157, 0.1219, 0.2194]], device=dev)
    lfb = -0.5 * ext
    for mode, dense in (("hard", hard), ("soft", soft)):
        pts, cnt = ops.gather_correspondences(off, choose, xm, ym, ext, lfb)
        ops.region_decode_points(pts, cnt, off, lg, rp, ext, lfb, choose=choose, mode=mode)
        want, _ = ops.gather_correspondences(dense, choose, xm, ym, ext, lfb)
        assert torch.equal(pts, want), mode
    # compact_map output (pixel recovered from U, V and the crop offset)
    mask = torch.zeros((bs, h, w), dtype=torch.uint8, device=dev)
    mask.view(bs, -1).scatter_(1, choose, 1)
    roi = torch.tensor([[200, 300], [17, 5]], dtype=torch.int32, device=dev)
    pts, cnt = ops.compact_map(off, mask, ext, lfb, 256, roi=roi)
    ops.region_decode_points(pts, cnt, off, lg, rp, ext, lfb, roi=roi, mode="hard")
    want, cnt2 = ops.compact_map(hard, mask, ext, lfb, 256, roi=roi)
    assert torch.equal(cnt, cnt2) and torch.equal(pts[:, :, :200], want[:, :, :200])


def test_get_pose_with_region_coded_map(g):
    """get_pose(region_mode='hard'): the offset map + region logits of a frame give the pose of the plain map."""
    models = synth.class_models(500)
    sp = load_frames(g, "clean")[2]
    pred, data = _datas(sp, models)
    R0, t0 = pb.get_pose(pred, data)
    n = 65
    gen = torch.Generator().manual_seed(9)
    rp = torch.rand(1, n, 3, generator=gen)
    rp[:, 0] = 0
    _, _, h, w = pred["xyz"].shape
    cls = torch.randint(0, n, (1, h, w), generator=gen)
    logits = torch.randn(1, n, h, w, generator=gen)
    logits.scatter_(1, cls[:, None], 10.0)                                  # the drawn class wins at every pixel
    base = rp[0][cls[0]].permute(2, 0, 1)[None]                             # [1,3,h,w]
    pred2 = {"xyz": pred["xyz"] - base, "region": logits}
    data2 = dict(data, region_point=rp)
    R1, t1 = pb.get_pose(pred2, data2, region_mode="hard")
    assert rot_angle(R1[0].numpy(), R0[0].numpy()) <= 1e-4 and (t1 - t0).abs().max() <= 1e-4
    Rp, tp = pb.get_pose(pred2, data2)                                      # offsets alone are not a pose
    assert rot_angle(Rp[0].numpy(), R0[0].numpy()) > 1e-2 or (tp - t0).abs().max() > 1e-2
