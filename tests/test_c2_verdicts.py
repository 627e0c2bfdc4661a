"""Pass/fail parity on ALL 1024 frames of the benchmark workload (C2) against the reference's own
``Trainer.get_pose`` + ``Trainer.cal_dis`` (tools/trainer.py:370-438), run in the build container by
tests/golden/gen_golden_c2_verdicts.py -> tests/golden/c2_ref_verdicts.npz.

The reference is a randomised estimator: it solves on 256 pixels drawn by an unseeded ``torch.randperm``
(trainer.py:407) with cv2's 100-iteration EPnP RANSAC, so two runs of the REFERENCE ITSELF give different verdicts
on frames whose error sits near a threshold (1 mm coordinate noise + 25 % outliers put the small objects' ADD at
0.3 .. 1.5 of ``0.1 * diameter``).  The golden file therefore holds two runs of it (seeds A and B).  The bar:

  * our verdicts (deterministic: every masked pixel, 256 P3P hypotheses, Philox stream keyed by the frame index)
    disagree with a reference run on NO MORE frames than the reference disagrees with itself;
  * every disagreeing frame is an ADD(-S) / rotation verdict of a frame whose errors are both small against the
    object (listed by the test with both sides' numbers) -- never a translation verdict, never a failed solve;
  * the pose deviation from cv2's on these noisy frames is stated (median / 90 % / max), not asserted to 1e-4: that
    bar applies to clear-consensus frames and is tested in test_gpu_callsites_ref.py.
"""
import os

import numpy as np
import pytest

from oracle import ref_callsites as rc
from pose_estimation_b200 import synth
from tests.util import rot_angle

GOLD = os.path.join(os.path.dirname(__file__), "golden")
SEED, CONFIG_ID, N_MODEL, F = 0x5EED, 2, 500, 1024


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLD, "c2_ref_verdicts.npz"))


def _verdicts(dis):
    cls = np.arange(dis.shape[0]) % synth.N_CLS
    return np.stack([dis[:, 0] < 0.1 * synth.DIAMETER[cls], dis[:, 1] < 5.0, dis[:, 2] < 0.05], axis=1)


def compare_with_reference(g, ours_dis, R=None, t=None):
    """ours_dis [F,3] (add, r_deg, t_m) -> report dict; asserts the bar of the module docstring."""
    ref_a, ref_b = g["verdict"].astype(bool), g["verdict_b"].astype(bool)
    assert np.array_equal(ref_a, _verdicts(g["dis"])) and g["ok"].all()
    ours = _verdicts(ours_dis)
    self_dis = int((ref_a != ref_b).any(axis=1).sum())
    dis_a = np.nonzero((ours != ref_a).any(axis=1))[0]
    dis_b = np.nonzero((ours != ref_b).any(axis=1))[0]
    rep = {"frames": int(ours.shape[0]), "reference_vs_itself": self_dis, "ours_vs_run_a": int(dis_a.size),
           "ours_vs_run_b": int(dis_b.size), "succ_ours": ours.mean(axis=0).round(4).tolist(),
           "succ_ref_a": ref_a.mean(axis=0).round(4).tolist(), "succ_ref_b": ref_b.mean(axis=0).round(4).tolist()}
    lines = []
    for f in dis_a:
        thr = 0.1 * float(synth.DIAMETER[f % synth.N_CLS])
        lines.append(f"frame {f} cls {f % synth.N_CLS} thr {thr:.5f}: ours add {ours_dis[f, 0]:.5f} r {ours_dis[f, 1]:.3f} "
                     f"{ours[f].astype(int).tolist()} | ref add {g['dis'][f, 0]:.5f} r {g['dis'][f, 1]:.3f} "
                     f"{ref_a[f].astype(int).tolist()} | ref run B add {g['dis_b'][f, 0]:.5f} {ref_b[f].astype(int).tolist()}")
    rep["listing"] = lines
    print("\n".join([str({k: v for k, v in rep.items() if k != "listing"})] + lines))
    assert dis_a.size <= self_dis and dis_b.size <= self_dis, rep
    assert np.array_equal(ours[:, 2], ref_a[:, 2])                       # the 5 cm verdict never differs
    for f in dis_a:                                                      # both sides' errors stay small: < 2.5x the threshold
        thr = 0.1 * float(synth.DIAMETER[f % synth.N_CLS])
        assert ours_dis[f, 0] < 2.5 * thr and g["dis"][f, 0] < 2.5 * thr and max(ours_dis[f, 1], g["dis"][f, 1]) < 12.5, lines
    # accuracy against the ground truth is not worse than the reference's
    assert np.median(ours_dis[:, 0]) <= np.median(g["dis"][:, 0]) and ours[:, 0].mean() >= ref_a[:, 0].mean() - 0.005
    if R is not None:
        dR = np.degrees([rot_angle(R[f], g["R"][f]) for f in range(ours.shape[0])])
        dt = np.abs(t - g["t"]).max(axis=1)
        rep["dev_vs_cv2"] = {"rot_deg": [float(np.median(dR)), float(np.quantile(dR, 0.9)), float(dR.max())],
                             "t_m": [float(np.median(dt)), float(np.quantile(dt, 0.9)), float(dt.max())]}
        print("deviation from the reference's pose (median / 90 % / max):", rep["dev_vs_cv2"])
        # the two estimators see the same 1 mm noise through different samples: a few degrees / about a centimetre
        assert np.median(dR) < 2.0 and np.median(dt) < 0.006 and dt.max() < 0.05
    return rep


def test_oracle_replays_reference_on_c2_frames(g):
    """oracle/ref_callsites.get_pose / cal_dis given the recorded randperm draws reproduce the reference's own
    run on the first 128 frames of the workload: R, t bit for bit, the three distances to 1e-5."""
    models = synth.class_models(N_MODEL)
    metric = rc.Metric(synth.SYM_CLS)
    for f in range(g["picks"].shape[0]):
        sp = synth.make_frame_sparse(SEED, CONFIG_ID, f)
        fr = synth.reference_frame_dict(sp, models[sp["cls"]])
        R, t, ok, _ = rc.get_pose(fr["xyz"], fr["extent"], fr["lfborder"], fr["choose"], fr["x_map_choosed"],
                                  fr["y_map_choosed"], fr["intrinsic"], pick=g["picks"][f].astype(np.int64))
        assert ok and np.array_equal(t, g["t"][f]) and np.abs(R - g["R"][f]).max() < 2e-6, f
        a, rd, td = rc.cal_dis(metric, g["R"][f], g["t"][f], fr["target_r"], fr["target_t"], fr["target"],
                               fr["model_points"], fr["cls_id"])
        ra, rr, rt = g["dis"][f]
        assert abs(a - ra) <= 1e-5 * ra + 1e-7 and abs(td - rt) <= 1e-5 * rt + 1e-7 and abs(rd - rr) < 0.06, f


def test_algorithm_oracle_verdicts_vs_reference_on_all_c2_frames(cref, g):
    """oracle/pose_ref.c (bit-identical to the CUDA kernels' integer results, tests/test_gpu_ransac.py) over all
    1024 frames against the two runs of the reference."""
    models = synth.class_models(N_MODEL)
    metric = rc.Metric(synth.SYM_CLS)
    dis, R, t = np.zeros((F, 3)), np.zeros((F, 3, 3), np.float32), np.zeros((F, 3), np.float32)
    for f in range(F):
        sp = synth.make_frame_sparse(SEED, CONFIG_ID, f)
        o = cref.ransac_frame(synth.compact_points(sp), synth.LINEMOD_K4, frame=f)
        assert o["status"] == 0
        fr = synth.reference_frame_dict(sp, models[sp["cls"]])
        dis[f] = rc.cal_dis(metric, o["R"], o["t"], fr["target_r"], fr["target_t"], fr["target"], fr["model_points"],
                            fr["cls_id"])
        R[f], t[f] = o["R"], o["t"]
    compare_with_reference(g, dis, R, t)


@pytest.mark.gpu
def test_eval_sweep_verdicts_vs_reference_on_all_c2_frames(cref, dev, g):
    """The 1024-frame EvalSweep launch bench.py times, frame by frame against the reference's run; and against the
    C oracle on every 8th frame (same numbers, so the CPU test above speaks for the GPU path)."""
    import torch
    import bench
    import pose_estimation_b200 as pb
    from pose_estimation_b200 import ops
    inp = bench.build_inputs(torch, dev, 0, F)
    sw = pb.EvalSweep(torch.from_numpy(synth.class_models(N_MODEL)), torch.from_numpy(synth.DIAMETER),
                      torch.from_numpy(synth.IS_SYM), F, synth.IMG_H, synth.IMG_W, bench.CAP,
                      ops.RansacConfig(n_hyp=bench.N_HYP, seed=SEED), device=dev)
    add, r_deg, t_m = sw.step(*(inp[k] for k in ("xyz", "mask", "extent", "lfborder", "K4", "cls", "R_gt", "t_gt")))
    dis = torch.stack([add, r_deg, t_m], dim=1).double().cpu().numpy()
    R, t = sw.ransac.R.cpu().numpy(), sw.ransac.t.cpu().numpy()
    assert int(sw.ransac.status.abs().sum()) == 0
    rep = compare_with_reference(g, dis, R, t)
    # the table's success counts are these verdicts
    tb = sw.finish().cpu().numpy()
    v = _verdicts(dis)
    assert tb[pb.sweep.ops_metric.ROW_VARIANT0].sum() == v[:, 0].sum() and tb[pb.sweep.ops_metric.ROW_VARIANT0 + 2].sum() == v[:, 1].sum()
    models = synth.class_models(N_MODEL)
    metric = rc.Metric(synth.SYM_CLS)
    for f in range(0, F, 8):
        sp = synth.make_frame_sparse(SEED, CONFIG_ID, f)
        o = cref.ransac_frame(synth.compact_points(sp), synth.LINEMOD_K4, frame=f)
        assert rot_angle(R[f], o["R"]) <= 1e-4 and np.abs(t[f] - o["t"]).max() <= 1e-4, f
        fr = synth.reference_frame_dict(sp, models[sp["cls"]])
        a, rd, td = rc.cal_dis(metric, o["R"], o["t"], fr["target_r"], fr["target_t"], fr["target"], fr["model_points"],
                               fr["cls_id"])
        assert abs(dis[f, 0] - a) <= 2e-5 * a + 1e-7 and abs(dis[f, 2] - td) <= 1e-5, f
    assert rep["ours_vs_run_a"] <= rep["reference_vs_itself"]
