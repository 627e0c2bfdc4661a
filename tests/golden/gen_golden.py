"""Generate tests/golden/*.npz in THIS container (real cv2 4.13.0 + the C oracle).

  python tests/golden/gen_golden.py

pnp_cv2.npz     reference-behaviour pins: the reference's exact per-frame call sequence
                (oracle/ref_callsites.py: get_pose -> cal_dis, i.e. tools/trainer.py:383-438 and
                :370-381 around the real cv2.solvePnPRansac) on seeded synthetic frames.
                'clean' frames are noise-free with outliers >= 5 px off (clear consensus: R,t must
                agree to 1e-4 rad / 1e-4 m); 'noisy' frames carry 1 mm noise + 25 % outliers
                (pass/fail of ADD(-S) < 0.1 d, r < 5 deg, t < 5 cm must be identical).
ransac_c.npz    algorithm pins: bit-level outputs of oracle/pose_ref.c for fixed seeds.
metric.npz      ADD / ADD-S / KNN / angular pins from ref_callsites on seeded clouds.
Frames are stored sparsely (rows, cols, xyzn) -- synth.dense_from_sparse rebuilds the maps.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cref, ref_callsites as rc  # noqa: E402
from pose_estimation_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
N_MODEL = 500


def frames(config_id, n, **kw):
    return [synth.make_frame_sparse(0x5EED, config_id, f, **kw) for f in range(n)]


def pack(sps):
    d = {}
    for i, s in enumerate(sps):
        for k in ("rows", "cols", "xyzn", "R_gt", "t_gt"):
            d[f"f{i}_{k}"] = s[k]
        d[f"f{i}_cls"] = np.int32(s["cls"])
    return d


def main():
    models = synth.class_models(N_MODEL)
    metric = rc.Metric(synth.SYM_CLS)
    out = {}
    for tag, cfgid, n, kw in (("clean", 101, 13, dict(noise_m=0.0, outlier_min_px=5.0)),
                              ("noisy", 102, 26, dict())):
        sps = frames(cfgid, n, **kw)
        out.update({f"{tag}_{k}": v for k, v in pack(sps).items()})
        res = []
        for f, sp in enumerate(sps):
            fr = synth.reference_frame_dict(sp, models[sp["cls"]])
            e = rc.eval_frame(metric, fr, synth.DIAMETER, np.random.default_rng([7, cfgid, f]))
            res.append(e)
        out[f"{tag}_n"] = np.int32(n)
        out[f"{tag}_R"] = np.stack([e["R"] for e in res])
        out[f"{tag}_t"] = np.stack([e["t"] for e in res])
        out[f"{tag}_ok"] = np.array([e["ok"] for e in res])
        out[f"{tag}_add"] = np.array([e["add"] for e in res], np.float64)
        out[f"{tag}_r_dis"] = np.array([e["r_dis"] for e in res], np.float64)
        out[f"{tag}_t_dis"] = np.array([e["t_dis"] for e in res], np.float64)
        out[f"{tag}_succ"] = np.array([[e["succ_rt"], e["succ_r"], e["succ_t"]] for e in res])
    np.savez_compressed(os.path.join(OUT, "pnp_cv2.npz"), **out)

    # ---- C-oracle pins
    sps = frames(2, 6)
    o = {}
    for lo in (0, 2):
        for f, sp in enumerate(sps):
            r = cref.ransac_frame(synth.compact_points(sp), synth.LINEMOD_K4, frame=f, lo_rounds=lo)
            for k in ("R", "t", "hyp_counts", "bitmap", "R0", "t0"):
                o[f"lo{lo}_f{f}_{k}"] = r[k]
            o[f"lo{lo}_f{f}_ints"] = np.array([r["status"], r["n_inl"], r["best_hyp"], r["n_inl_refined"]], np.int32)
    o["philox"] = np.stack([cref.philox([1, 2, 3, 4], [5, 6]), cref.philox([0, 0, 0, 0], [0, 0]),
                            cref.philox([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2)])
    np.savez_compressed(os.path.join(OUT, "ransac_c.npz"), **o)

    # ---- metric pins
    rng = np.random.default_rng(11)
    m = {}
    for cls, N in ((7, 2620), (8, 2620), (0, 500)):
        model = (synth.LFBORDER[cls] + rng.random((N, 3)) * synth.EXTENT[cls]).astype(np.float32)
        Rg = synth.random_rotation(rng).astype(np.float32)
        tg = np.array([0.02, -0.03, 0.8], np.float32)
        ang = np.radians(4.0)
        dR = np.array([[np.cos(ang), -np.sin(ang), 0], [np.sin(ang), np.cos(ang), 0], [0, 0, 1]], np.float32)
        R = (dR @ Rg).astype(np.float32)
        t = tg + np.array([0.003, -0.002, 0.004], np.float32)
        pred = (model @ R.T + t).astype(np.float32)
        tgt = (model @ Rg.T + tg).astype(np.float32)
        add, adds = rc.Metric([cls]).cal_adds_cuda(pred, tgt, cls)
        add_ns, _ = rc.Metric([]).cal_adds_cuda(pred, tgt, cls)
        m[f"c{cls}_model"] = model; m[f"c{cls}_R"] = R; m[f"c{cls}_t"] = t; m[f"c{cls}_Rg"] = Rg; m[f"c{cls}_tg"] = tg
        m[f"c{cls}_vals"] = np.array([add_ns, adds, rc.Metric.angular_distance(R, Rg),
                                      rc.Metric.translation_distance(t, tg)], np.float64)
        m[f"c{cls}_knn"] = rc.knn(pred, tgt).reshape(-1)
        m[f"c{cls}_loss"] = np.float64(rc.pose_loss([7, 8], R[None], t[None], tgt[None], model[None], [cls]))
    np.savez_compressed(os.path.join(OUT, "metric.npz"), **m)
    print("wrote", [f for f in os.listdir(OUT) if f.endswith(".npz")])


if __name__ == "__main__":
    main()
