"""CPU tests: the oracles against the committed golden vectors (tests/golden/gen_golden.py made them
with the real cv2 in-container) and against each other."""
import os

import numpy as np
import pytest

from oracle import ref_callsites as rc
from pose_estimation_b200 import synth
from tests.util import rot_angle

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def load_frames(g, tag):
    out = []
    for f in range(int(g[f"{tag}_n"])):
        sp = {k: g[f"{tag}_f{f}_{k}"] for k in ("rows", "cols", "xyzn", "R_gt", "t_gt")}
        sp["cls"] = int(g[f"{tag}_f{f}_cls"])
        out.append(sp)
    return out


def test_philox_known_answers(cref):
    # Random123 kat_vectors for philox4x32-10
    assert list(cref.philox([0, 0, 0, 0], [0, 0])) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert list(cref.philox([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2)) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert list(cref.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0])) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    g = np.load(os.path.join(GOLD, "ransac_c.npz"))
    assert np.array_equal(g["philox"][1], cref.philox([0, 0, 0, 0], [0, 0]))


def test_sampler_distinct_and_uniform(cref):
    seen = np.zeros(50, int)
    for h in range(2000):
        ok, idx = cref.sample4(0x5EED, 3, h, 50)
        assert ok and len(set(idx.tolist())) == 4 and idx.min() >= 0 and idx.max() < 50
        seen[idx] += 1
    assert seen.min() > 100 and seen.max() < 230          # 160 expected per bin
    assert not cref.sample4(1, 0, 0, 3)[0]
    ok, idx = cref.sample4(1, 0, 0, 4)
    assert ok and sorted(idx.tolist()) == [0, 1, 2, 3]


def test_p3p_recovers_exact_pose(cref):
    rng = np.random.default_rng(1)
    errs = []
    for _ in range(300):
        R = synth.random_rotation(rng)
        t = np.array([rng.uniform(-.15, .15), rng.uniform(-.1, .1), rng.uniform(.6, 1.2)])
        X = (rng.random((3, 3)) - .5) * 0.1
        Xc = X @ R.T + t
        Rs, ts = cref.p3p(X, Xc / np.linalg.norm(Xc, axis=1, keepdims=True))
        assert len(Rs) <= 4
        if len(Rs):
            errs.append(min(np.abs(Rs[k] - R).max() + np.abs(ts[k] - t).max() for k in range(len(Rs))))
    errs = np.array(errs)
    assert len(errs) >= 295 and np.median(errs) < 5e-4 and np.quantile(errs, 0.9) < 5e-3


def test_c_oracle_matches_its_golden(cref):
    g = np.load(os.path.join(GOLD, "ransac_c.npz"))
    sps = synth.make_batch_sparse(0x5EED, 2, 6)
    for lo in (0, 2):
        for f, sp in enumerate(sps):
            r = cref.ransac_frame(synth.compact_points(sp), synth.LINEMOD_K4, frame=f, lo_rounds=lo)
            ints = g[f"lo{lo}_f{f}_ints"]
            assert [r["status"], r["n_inl"], r["best_hyp"], r["n_inl_refined"]] == ints.tolist()
            assert np.array_equal(r["hyp_counts"], g[f"lo{lo}_f{f}_hyp_counts"])
            assert np.array_equal(r["bitmap"], g[f"lo{lo}_f{f}_bitmap"])
            assert np.array_equal(r["R"], g[f"lo{lo}_f{f}_R"]) and np.array_equal(r["t"], g[f"lo{lo}_f{f}_t"])


def test_c_oracle_vs_cv2_golden(cref):
    """Algorithm oracle vs the reference's cv2 path: clear-consensus frames agree to 1e-4 rad /
    1e-4 m; noisy frames give identical ADD(-S)/r/t pass-fail (tools/trainer.py:186-191)."""
    g = np.load(os.path.join(GOLD, "pnp_cv2.npz"))
    models = synth.class_models(500)
    metric = rc.Metric(synth.SYM_CLS)
    for tag in ("clean", "noisy"):
        assert g[f"{tag}_ok"].all()
        for f, sp in enumerate(load_frames(g, tag)):
            o = cref.ransac_frame(synth.compact_points(sp), synth.LINEMOD_K4, frame=f)
            assert o["status"] == 0
            if tag == "clean":
                assert rot_angle(o["R"], g["clean_R"][f]) <= 1e-4
                assert np.abs(o["t"] - g["clean_t"][f]).max() <= 1e-4
            fr = synth.reference_frame_dict(sp, models[sp["cls"]])
            a, rd, td = rc.cal_dis(metric, o["R"], o["t"], fr["target_r"], fr["target_t"], fr["target"],
                                   fr["model_points"], fr["cls_id"])
            succ = [a < 0.1 * synth.DIAMETER[sp["cls"]], rd < 5.0, td < 0.05]
            assert succ == g[f"{tag}_succ"][f].tolist()


@pytest.mark.skipif(rc.cv2 is None, reason="cv2 missing")
def test_ref_callsites_reproduce_golden():
    """The golden file is what ref_callsites (real cv2) produces today: guards against drift."""
    g = np.load(os.path.join(GOLD, "pnp_cv2.npz"))
    models = synth.class_models(500)
    metric = rc.Metric(synth.SYM_CLS)
    for f, sp in enumerate(load_frames(g, "clean")[:4]):
        fr = synth.reference_frame_dict(sp, models[sp["cls"]])
        e = rc.eval_frame(metric, fr, synth.DIAMETER, np.random.default_rng([7, 101, f]))
        assert np.allclose(e["R"], g["clean_R"][f], atol=1e-6) and np.allclose(e["t"], g["clean_t"][f], atol=1e-6)
        assert abs(e["add"] - g["clean_add"][f]) < 1e-6


def test_metric_oracles_agree(cref):
    """C oracle (what the GPU is compared with) vs ref_callsites' restatement of lib/utils/metric.py."""
    g = np.load(os.path.join(GOLD, "metric.npz"))
    for cls in (7, 8, 0):
        model, R, t, Rg, tg = (g[f"c{cls}_{k}"] for k in ("model", "R", "t", "Rg", "tg"))
        pred, tgt = cref.transform(model, R, t), cref.transform(model, Rg, tg)
        add, adds = cref.add_adds(pred, tgt)
        vals = g[f"c{cls}_vals"]
        assert abs(add - vals[0]) <= 2e-6 * vals[0] + 1e-9 and abs(adds - vals[1]) <= 2e-6 * vals[1] + 1e-9
        assert abs(cref.angular_deg(R, Rg) - vals[2]) < 2e-3
        idx, _ = cref.knn1(pred, tgt)
        ref_idx = g[f"c{cls}_knn"]
        # FP32 formulas differ in fma contraction: indices may differ only at numerical near-ties
        diff = np.nonzero(idx != ref_idx)[0]
        assert len(diff) <= 2
        for i in diff:
            da = np.sum((pred[i].astype(np.float64) - tgt[idx[i]]) ** 2)
            db = np.sum((pred[i].astype(np.float64) - tgt[ref_idx[i]]) ** 2)
            assert abs(da - db) <= 1e-6 * max(da, db)


def test_compaction_oracle_matches_reference_gather(cref):
    """ref_compact_map == the reference's de-normalise + gather at `choose` (trainer.py:415-421)."""
    sp = synth.make_frame_sparse(0x5EED, 1, 0)
    xyz, mask = synth.dense_from_sparse(sp)
    cls = sp["cls"]
    pts, total = cref.compact_map(xyz, mask, synth.EXTENT[cls], synth.LFBORDER[cls])
    assert total == sp["rows"].shape[0] == pts.shape[1]
    coord = (xyz * synth.EXTENT[cls].reshape(3, 1, 1) + synth.LFBORDER[cls].reshape(3, 1, 1))   # float64
    choose = sp["rows"].astype(np.int64) * synth.IMG_W + sp["cols"]
    ref = coord.reshape(3, -1)[:, choose]
    assert np.abs(pts[:3] - ref).max() < 2e-8            # FP32 fma vs the reference's float64
    assert np.array_equal(pts[3], sp["cols"]) and np.array_equal(pts[4], sp["rows"])
    pts2, _ = cref.compact_map(xyz, None, synth.EXTENT[cls], synth.LFBORDER[cls])
    assert np.array_equal(pts, pts2)                     # xyz != 0 convention gives the same set
    e, _ = cref.compact_map(np.zeros((3, 8, 8), np.float32), None, np.ones(3), np.zeros(3))
    assert e.shape == (5, 0)


# ---- 3D-3D RANSAC + Umeyama: the oracle against vectors produced by RUNNING the reference
# (tests/golden/gen_golden_sim3.py imports /root/reference/lib/transform/umeyama.py) ---------------------
def _sim3_golden():
    import os
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sim3_ref.npz"))


def test_sim3_oracle_matches_reference_run_iteration_by_iteration():
    """With the reference's own np.random index sets: every iteration's inlier count, the chosen iteration (so
    the confidence break fired at the same place), the refit set size and the final transform."""
    from oracle import cref
    g = _sim3_golden()
    for k in range(5):
        src, dst, idx, counts = g[f"src{k}"], g[f"dst{k}"], g[f"idx{k}"], g[f"counts{k}"]
        it = idx.shape[0]                                   # iterations the reference ran before its break
        si = np.concatenate([idx, np.repeat(idx[-1:], 128 - it, 0)]) if it < 128 else idx
        r = cref.sim3_ransac(src, dst, sample_idx=si)
        assert r["status"] == 0
        assert np.array_equal(r["hyp_counts"][:it], counts)                 # integer work: exact
        assert r["best_hyp"] == int(np.argmax(counts))                      # first maximum = strict '>' update
        assert r["n_inl"] == int(counts.max())
        assert abs(r["scale"] - float(g[f"scale{k}"])) <= 2e-7 * max(1.0, float(g[f"scale{k}"]))
        assert np.abs(r["R"] - g[f"R{k}"]).max() <= 1e-6                    # FP32 outputs of an FP64 fit
        assert np.abs(r["t"] - g[f"t{k}"]).max() <= 1e-6


def test_umeyama_oracle_matches_reference_fit_including_reflection():
    from oracle import cref
    g = _sim3_golden()
    for k in range(4):
        st, sc, R, t = cref.umeyama_fit(g[f"u_src{k}"], g[f"u_dst{k}"])
        assert st == 0
        assert abs(sc - float(g[f"u_scale{k}"])) <= 2e-7 * max(1.0, sc)
        assert np.abs(R - g[f"u_R{k}"]).max() <= 1e-6 and np.abs(t - g[f"u_t{k}"]).max() <= 1e-6
        assert abs(np.linalg.det(R.astype(np.float64)) - 1.0) < 1e-5       # proper rotation even for the mirrored case


def test_sim3_oracle_edge_cases():
    from oracle import cref
    rng = np.random.default_rng(3)
    src = rng.random((40, 3)).astype(np.float32)
    # no consistent transform: unrelated clouds -> best ratio stays below 0.1 -> the reference's None return
    r = cref.sim3_ransac(src, rng.random((40, 3)).astype(np.float32) * 50, n_iter=16)
    assert r["status"] in (0, 2)
    # all draws identical (np.random can return five equal indices): zero variance -> no inliers for that iteration
    si = np.zeros((16, 5), np.int32)
    r = cref.sim3_ransac(src, src * 2 + 1, n_iter=16, sample_idx=si)
    assert r["status"] == 2 and r["best_hyp"] == -1 and not r["hyp_counts"].any()
    # counter RNG draws are in range and reproducible
    a = cref.sim3_sample5(0x5EED, 3, 7, 40)
    assert np.array_equal(a, cref.sim3_sample5(0x5EED, 3, 7, 40)) and a.min() >= 0 and a.max() < 40


def test_fps_oracle_matches_reference_selection():
    """tools/script/sample_model.py:33-46 run here (tests/golden/gen_golden_fps.py): identical index sequences,
    including duplicated points and n_samples > N."""
    import os
    from oracle import cref
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fps_ref.npz"))
    for k in range(5):
        want = g[f"idx{k}"]
        assert np.array_equal(cref.fps(g[f"pts{k}"], want.shape[0]), want), k


def _rel(a, b):
    return float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / np.abs(b).max())


def test_bpnp_oracle_matches_reference_run():
    """oracle/bpnp_ref.py against tests/golden/bpnp_ref.npz, which gen_golden_bpnp.py made by importing the
    reference's own BPnP.py (forward through cv2, backward through its autograd recipe, float32)."""
    import torch
    from oracle import bpnp_ref as br
    pytest.importorskip("cv2")
    g = np.load(os.path.join(GOLD, "bpnp_ref.npz"))
    for c in list(g["cases"]) + ["ini"]:
        ini = g["ini_ini"] if c == "ini" else None
        P = br.bpnp_forward(g[f"{c}_pts2d"], g[f"{c}_pts3d"], g["K"], ini)
        np.testing.assert_allclose(P, g[f"{c}_P6d"], rtol=0, atol=1e-6)                 # same cv2 calls
        for dt, tol in ((torch.float32, 2e-5), (torch.float64, 5e-4)):                  # f64: the reference's f32 rounding
            gx, gz, gK = br.bpnp_backward(g[f"{c}_pts2d"], g[f"{c}_pts3d"], g["K"], g[f"{c}_P6d"], g[f"{c}_grad_out"], dtype=dt)
            assert _rel(gx, g[f"{c}_grad_x"]) < tol and _rel(gz, g[f"{c}_grad_z"]) < tol and _rel(gK, g[f"{c}_grad_K"]) < tol


def test_bpnp_oracle_rotation_formula_is_a_rotation_and_switches_branch():
    import torch
    from oracle import bpnp_ref as br
    aa = torch.tensor([[0.3, -1.2, 0.5], [1e-4, 2e-4, -1e-4], [0.0, 0.0, 3.0]], dtype=torch.float64)
    R = br.angle_axis_to_rotation_matrix(aa).numpy()
    for k in (0, 2):
        np.testing.assert_allclose(R[k] @ R[k].T, np.eye(3), atol=1e-5)                 # eps in the axis: 1e-6 off unit
    np.testing.assert_array_equal(R[1], np.array([[1, 1e-4, 2e-4], [-1e-4, 1, -1e-4], [-2e-4, 1e-4, 1]]))


def test_possloss_oracle_matches_reference_run():
    """oracle/possloss_ref.py against tests/golden/possloss_ref.npz (gen_golden_possloss.py ran the reference's own
    PossLoss.forward + autograd backward in float32)."""
    import torch
    from oracle import possloss_ref as pr
    g = np.load(os.path.join(GOLD, "possloss_ref.npz"))
    for c in g["cases"]:
        idx, refine = g[f"{c}_meta"]
        for dt, tol in ((torch.float32, 2e-5), (torch.float64, 2e-5)):
            v, gr, gt, gc = pr.poss_loss_np(g[f"{c}_pred_r"], g[f"{c}_pred_t"], g[f"{c}_pred_c"], g[f"{c}_target"], g[f"{c}_model"],
                                            idx, refine, float(g["w"]), g[f"{c}_axis"], g[f"{c}_target_r"], g["sym_list"],
                                            gouts=g["gouts"], dtype=dt)
            assert _rel(v, g[f"{c}_vals"]) < tol and _rel(gr, g[f"{c}_grad_r"]) < tol
            assert _rel(gt, g[f"{c}_grad_t"]) < tol and _rel(gc, g[f"{c}_grad_c"]) < tol


# ---------------------------------------------------------------------------------------------------
# callsites_ref.npz: outputs of the reference's OWN get_pose / cal_dis / Metric / utlis / PoseLoss run in the
# build container (tests/golden/gen_golden_callsites.py) -- this is what pins oracle/ref_callsites.py.
def _callsites():
    return np.load(os.path.join(GOLD, "callsites_ref.npz"))


@pytest.mark.skipif(rc.cv2 is None, reason="cv2 missing")
def test_ref_callsites_get_pose_cal_dis_match_reference_run():
    g = _callsites()
    models = synth.class_models(500)
    metric = rc.Metric(synth.SYM_CLS)
    for tag in ("clean", "noisy"):
        for f, sp in enumerate(load_frames(g, tag)):
            fr = synth.reference_frame_dict(sp, models[sp["cls"]])
            R, t, ok, _ = rc.get_pose(fr["xyz"], fr["extent"], fr["lfborder"], fr["choose"], fr["x_map_choosed"],
                                      fr["y_map_choosed"], fr["intrinsic"], pick=g[f"{tag}_pick"][f])
            assert ok
            # same cv2 call on the same 256 correspondences: rvec / tvec are identical, R differs only by
            # Rodrigues (cv2, float64) vs the kornia formula in float32
            assert np.abs(R - g[f"{tag}_R"][f]).max() < 2e-6
            assert np.array_equal(t, g[f"{tag}_t"][f])
            add, r_dis, t_dis = rc.cal_dis(metric, g[f"{tag}_R"][f], g[f"{tag}_t"][f], fr["target_r"], fr["target_t"],
                                           fr["target"], fr["model_points"], fr["cls_id"])
            ref_add, ref_r, ref_t = g[f"{tag}_dis"][f]
            assert abs(add - ref_add) <= 1e-5 * ref_add + 1e-7
            assert abs(t_dis - ref_t) <= 1e-5 * ref_t + 1e-7
            # the reference takes acos of a float32 dot product clamped to 1 - 1e-7: below ~0.1 degree the
            # angle is quantised in steps of ~0.03 degree (an exact match reads 0.056), so compare loosely
            assert abs(r_dis - ref_r) < 0.06
            d = synth.DIAMETER[sp["cls"]]
            assert [add < 0.1 * d, r_dis < 5.0, t_dis < 0.05] == [ref_add < 0.1 * d, ref_r < 5.0, ref_t < 0.05]


def test_ref_callsites_metric_matches_reference_run():
    g = _callsites()
    for cls in (7, 8, 3):
        model, R, t, Rg, tg = (g[f"m{cls}_{k}"] for k in ("model", "R", "t", "Rg", "tg"))
        pred, tgt = (model @ R.T + t).astype(np.float32), (model @ Rg.T + tg).astype(np.float32)
        add, adds = rc.Metric([7, 8]).cal_adds_cuda(pred, tgt, cls)
        vals = g[f"m{cls}_vals"]
        assert abs(add - vals[0]) <= 1e-5 * vals[0] and abs(adds - vals[1]) <= 1e-5 * vals[1]
        assert (add == adds) and ((cls in (7, 8)) or adds == add)
        assert abs(rc.Metric.angular_distance(R, Rg) - vals[2]) < 0.02
        assert abs(rc.Metric.translation_distance(t, tg) - vals[3]) <= 1e-6
    M = rc.Metric([7, 8])
    assert abs(M.cal_auc(list(g["auc_in"])) - float(g["auc_out"])) < 1e-9
    assert M.cal_auc([0.5, 0.7]) == float(g["auc_none"]) == 0.0


@pytest.mark.skipif(rc.cv2 is None, reason="cv2 missing")
def test_ref_callsites_utlis_matches_reference_run():
    g = _callsites()
    assert np.array_equal(rc.batch_intrinsic_transform(g["K4"]), g["K33"])
    assert np.array_equal(rc.batch_intrinsic_transform(g["K33"]), g["K4_back"])
    coor, uv, Ks = g["crb_coor"], g["crb_uv"], g["crb_K"]
    np.random.seed(77)
    assert np.array_equal(rc.choose_ransac_batch(coor, uv, Ks, n=125, rng=np.random), g["crb_out125"])
    for j in range(4):          # wrap-pad branch: the reference returns 1 + n - M identical rows per element
        ref = g[f"crb_wrap{j}"]
        assert ref.shape[0] == 1 and ref.shape[2] == 400 and (ref == ref[:, :1, :]).all()
        np.random.seed(78)
        ours = rc.choose_ransac_batch(coor[j:j + 1], uv[j:j + 1], Ks[j:j + 1], n=400, rng=np.random)
        assert ours.shape == (1, 1, 400) and np.array_equal(ours, ref[:, :1, :])


def test_ref_callsites_pose_loss_matches_reference_run():
    g = _callsites()
    val = rc.pose_loss([7, 8], g["pl_R"], g["pl_t"], g["pl_target"], g["pl_model"], g["pl_idxs"])
    assert abs(val - float(g["pl_loss"])) <= 2e-6 * float(g["pl_loss"])


def test_ref_callsites_transparent_metric_matches_reference_run():
    """version/transparent/lib/utils.py:17-46 run here: cal_dis takes the minimum over the TARGET points for every
    predicted point, cal_adds_cuda over the predicted points for every target point; a class in ``sys`` returns
    ADD-S twice."""
    g = _callsites()
    for cls in (7, 8, 3):
        model, R, t, Rg, tg = (g[f"m{cls}_{k}"] for k in ("model", "R", "t", "Rg", "tg"))
        pred, tgt = (model @ R.T + t).astype(np.float32), (model @ Rg.T + tg).astype(np.float32)
        add = float(np.linalg.norm(pred - tgt, axis=1).astype(np.float32).mean(dtype=np.float32))
        _, adds_tgt = rc.Metric([cls]).cal_adds_cuda(pred, tgt, cls)        # min over pred for every target
        _, adds_prd = rc.Metric([cls]).cal_adds_cuda(tgt, pred, cls)        # roles swapped: min over target
        sym = cls in (7, 8)
        want = [adds_prd if sym else add, adds_prd, adds_tgt if sym else add, adds_tgt]
        assert np.allclose(want, g[f"m{cls}_tvals"], rtol=1e-5, atol=0)


def test_ref_callsites_region_maps_match_reference_run():
    """Region-coded maps: tools/script/eval.py:93-103 (arg-max class, run as the lines stand) and
    lib/network/loss_utils.py:18-28 get_xyz (softmax-weighted mean), including an exact tie of two logits."""
    g = _callsites()
    for f in range(g["reg_xyz_off"].shape[0]):
        hard = rc.region_xyz_hard(g["reg_xyz_off"][f], g["reg_logits"][f], g["reg_points"][f])
        assert np.array_equal(hard, g["reg_hard"][f])
        soft = rc.get_xyz(g["reg_xyz_off"][f], g["reg_logits"][f], g["reg_points"][f])
        assert np.abs(soft - g["reg_soft"][f]).max() <= 1e-7
    assert np.array_equal(g["reg_hard"][0][:, 3, 4], g["reg_xyz_off"][0][:, 3, 4] + g["reg_points"][0][7])   # tie -> class 7


def test_auc_of_oracle_and_product_match_reference_run():
    """lib/utils/metric.py:38-65 run here on ties, border values, a single sample, mostly-out-of-range and unsorted
    lists: the restatement and the product's host-side Metric.cal_auc (pure numpy, no GPU involved) give the same AUC."""
    from pose_estimation_b200.api import Metric as ProductMetric
    g = _callsites()
    for k in range(int(g["auc_ncases"])):
        d, want = list(g[f"auc_case{k}_in"]), float(g[f"auc_case{k}_out"])
        assert abs(rc.Metric([]).cal_auc(d) - want) < 1e-9, k
        assert abs(ProductMetric([]).cal_auc(d) - want) < 1e-9, k
    assert abs(ProductMetric([]).cal_auc(list(g["auc_in"])) - float(g["auc_out"])) < 1e-9
    assert ProductMetric([]).cal_auc([0.5, 0.7]) == 0


# ---- SURVEY 8(f) row f3 remainder: oracle/f3_ref.py against runs of the reference's own files -----------------
@pytest.fixture(scope="module")
def f3():
    return np.load(os.path.join(GOLD, "f3_ref.npz"))


def test_f3_kabsch_oracle_matches_reference_run(f3):
    """leastsquares.py:10-48 run by gen_golden_f3.py (float32 SVD) vs the float64 restatement."""
    from oracle import f3_ref
    for k in range(int(f3["kab_n"])):
        R, t, cA, cB = f3_ref.kabsch(f3[f"kab{k}_A"], f3[f"kab{k}_B"])
        assert abs(np.linalg.det(R) - 1.0) < 1e-9
        if k != 3:                                  # planar cloud: the reference's float32 SVD is loose there
            assert np.abs(R - f3[f"kab{k}_R"]).max() < 5e-5, k
            assert np.abs(t - f3[f"kab{k}_t"]).max() < 5e-5, k
        assert np.abs(cA - f3[f"kab{k}_cA"]).max() < 1e-6 and np.abs(cB - f3[f"kab{k}_cB"]).max() < 1e-6
        # both are least-squares optimal: same residual
        A, B = f3[f"kab{k}_A"].astype(np.float64), f3[f"kab{k}_B"].astype(np.float64)
        r_ours = np.linalg.norm(A @ R.T + t - B)
        r_ref = np.linalg.norm(A @ f3[f"kab{k}_R"].astype(np.float64).T + f3[f"kab{k}_t"] - B)
        assert r_ours <= r_ref * (1 + 1e-5) + 1e-7, k


def test_f3_icp_oracle_matches_reference_loop(f3):
    from oracle import f3_ref
    for mode in ("reference", "rigid"):
        for h in range(3):
            R, t = f3_ref.icp_refine(f3["icp_model"], f3["icp_target"], f3["icp_R0"][h], f3["icp_t0"][h], 10, mode)
            assert np.abs(R - f3[f"icp_{mode}_R"][h]).max() < 2e-4, (mode, h)
            assert np.abs(t - f3[f"icp_{mode}_t"][h]).max() < 2e-4, (mode, h)
    # the rigid composition converges onto the ground-truth pose; the script's own update rule need not
    for h in range(3):
        assert np.abs(f3["icp_rigid_R"][h] - f3["icp_Rg"]).max() < 2e-3 and np.abs(f3["icp_rigid_t"][h] - f3["icp_tg"]).max() < 1e-3


def test_f3_trans_oracle_matches_reference_run(f3):
    """lib/transform/trans.py run with np.random.randint recorded: every iteration's residual and inlier ratio,
    the threshold, success and the returned transform."""
    from oracle import f3_ref
    for k in range(int(f3["tr_n"])):
        n_it = int(f3[f"tr{k}_niter"])
        o = f3_ref.trans_ransac(f3[f"tr{k}_src"], f3[f"tr{k}_dst"], f3[f"tr{k}_idx"][:n_it])
        n = f3[f"tr{k}_src"].shape[0]
        assert abs(o["pass_t"] - float(f3[f"tr{k}_pass_t"])) < 1e-12
        assert np.allclose(o["residuals"], f3[f"tr{k}_residual"], rtol=1e-10, atol=0)
        assert np.array_equal(o["n_inl"] / n, f3[f"tr{k}_ratio"])
        assert bool(o["ok"]) == bool(f3[f"tr{k}_ok"])
        if o["ok"]:
            assert np.allclose(o["Scales"], f3[f"tr{k}_S"], rtol=1e-10) and np.allclose(o["Rotation"], f3[f"tr{k}_R"], atol=1e-10)
            assert np.allclose(o["Translation"], f3[f"tr{k}_T"], atol=1e-9) and np.allclose(o["OutTransform"], f3[f"tr{k}_M"], atol=1e-9)


def test_f3_uvd2cloud_oracle_matches_reference_run(f3):
    from oracle import f3_ref
    assert np.array_equal(f3_ref.uvd2cloud(f3["uvd"], f3["uvd_K"]), f3["uvd_xyz"])


def test_report_format_matches_reference_lines():
    """sweep.format_report against the statements of tools/trainer.py:281-330 executed as they stand."""
    import json
    import torch
    from pose_estimation_b200 import sweep
    rep = json.load(open(os.path.join(GOLD, "report_ref.json")))
    table = torch.tensor(rep["table"], dtype=torch.float64)
    for key, opt_pose in (("base", False), ("opt_pose", True)):
        lines, all_content = sweep.format_report(table, rep["objlist"], opt_pose=opt_pose)
        ref = rep[key]
        assert ref[-1] == all_content and ref[-2] == '\n'
        assert ref[:len(lines)] == lines
