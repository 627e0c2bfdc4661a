"""Generate tests/golden/f3_ref.npz (+ report_ref.json) by RUNNING THE REFERENCE'S OWN CODE:

  best_batch_fit_transform    /root/reference/version/transparent/lib/declarative/leastsquares.py:10-48 (imports cleanly)
  the optimisation loop       /root/reference/version/transparent/scripts/train_declarative.py:80-100 -- its lines need the
                              whole network to run, so the loop is re-enacted here AROUND the reference's own
                              best_batch_fit_transform (KeOps arg-min stood in for by a float32 brute force)
  estimateSimilarityTransform /root/reference/lib/transform/trans.py (imports cleanly; np.random.randint recorded,
                              evaluateModel wrapped to record every iteration's residual / inlier count)
  uvd2cloud                   /root/reference/lib/transform/coordinate.py:91-101
  the per-object report       /root/reference/tools/trainer.py:281-330 -- the statements are taken out of the syntax tree
                              of Trainer.test_epoch and executed as they stand on a filled ``result`` dict

Run here only (the GPU box has no /root/reference):   python tests/golden/gen_golden_f3.py
"""
import ast
import importlib.util
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import f3_ref  # noqa: E402
from pose_estimation_b200 import synth  # noqa: E402

REF = "/root/reference"


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


ls = load(f"{REF}/version/transparent/lib/declarative/leastsquares.py", "ref_leastsquares")
trans = load(f"{REF}/lib/transform/trans.py", "ref_trans")
sys.path.insert(0, REF)
coord = load(f"{REF}/lib/transform/coordinate.py", "ref_coordinate")


def rot(rng, deg):
    w = rng.normal(size=3)
    w *= np.radians(deg) / np.linalg.norm(w)
    K = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    th = np.linalg.norm(w)
    return np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / th ** 2 * K @ K


def main():
    rng = np.random.default_rng(2024)
    out = {}
    # ---- Kabsch: c = 1 problems run as they stand (noisy, clean, reflected = the det < 0 branch, planar)
    cases = []
    for k in range(5):
        N = [500, 64, 1000, 300, 4][k]
        A = (rng.random((N, 3)) - 0.5) * [0.2, 0.1, 0.3]
        R0 = synth.random_rotation(rng)
        B = A @ R0.T + rng.normal(0, 0.3, 3)
        if k == 0:
            B += rng.normal(0, 0.002, B.shape)
        if k == 2:
            B = B * [1, 1, -1]                      # mirrored target: the fit must still return a proper rotation
        if k == 3:
            A[:, 2] = 0.0; B = A @ R0.T + 0.1       # planar cloud: smallest singular value 0
        cases.append((A.astype(np.float32), B.astype(np.float32)))
    for k, (A, B) in enumerate(cases):
        R, t, cA, cB = ls.best_batch_fit_transform(torch.from_numpy(A)[None], torch.from_numpy(B)[None])
        out[f"kab{k}_A"], out[f"kab{k}_B"] = A, B
        out[f"kab{k}_R"], out[f"kab{k}_t"] = R[0].numpy(), t.reshape(3).numpy()
        out[f"kab{k}_cA"], out[f"kab{k}_cB"] = cA.reshape(3).numpy(), cB.reshape(3).numpy()
    out["kab_n"] = np.int32(len(cases))

    # ---- the ICP-style loop, 10 steps, three hypotheses against one target cloud
    N = 400
    model = ((rng.random((N, 3)) - 0.5) * [0.15, 0.1, 0.07]).astype(np.float32)
    Rg = synth.random_rotation(rng)
    tg = np.array([0.02, -0.01, 0.7])
    target = (model @ Rg.T + tg + rng.normal(0, 0.0005, (N, 3))).astype(np.float32)
    Rs, ts, outs = [], [], {"reference": [], "rigid": []}
    for h in range(3):
        R0 = (rot(rng, 4.0 + 3 * h) @ Rg).astype(np.float32)
        t0 = (tg + rng.normal(0, 0.004, 3)).astype(np.float32)
        Rs.append(R0); ts.append(t0)
        for mode in ("reference", "rigid"):
            fr, ft = torch.from_numpy(R0)[None], torch.from_numpy(t0)[None]
            mp, tt = torch.from_numpy(model)[None], torch.from_numpy(target)[None]
            for step in range(10):                                                     # train_declarative.py:80
                pred = mp @ fr.permute(0, 2, 1) + ft.view(1, 1, 3)                     # :81
                inds = torch.from_numpy(f3_ref.knn_f32(pred[0].numpy(), target))       # :82 (KeOps stand-in)
                corr = torch.index_select(tt, 1, inds)                                 # :84
                oR, ot, _, _ = ls.best_batch_fit_transform(pred, corr)                 # :86 (diff_fit's forward)
                if mode == "reference":
                    fr, ft = fr @ oR, ft + ot.view(1, 3)                               # :98-99 as written
                else:
                    fr, ft = oR @ fr, (oR @ ft.view(1, 3, 1)).view(1, 3) + ot.view(1, 3)
            outs[mode].append((fr[0].numpy(), ft[0].numpy()))
    out["icp_model"], out["icp_target"] = model, target
    out["icp_R0"], out["icp_t0"] = np.stack(Rs), np.stack(ts)
    for mode in outs:
        out[f"icp_{mode}_R"] = np.stack([o[0] for o in outs[mode]])
        out[f"icp_{mode}_t"] = np.stack([o[1] for o in outs[mode]])
    out["icp_Rg"], out["icp_tg"] = Rg, tg

    # ---- trans.py: record np.random.randint and every iteration's residual / inlier count
    n_tr = 0
    for k in range(4):
        n = [600, 300, 1000, 200][k]
        src = ((rng.random((n, 3)) - 0.5) * [0.2, 0.15, 0.1] + [0.0, 0.0, 0.0]).astype(np.float32)
        s0, R0, T0 = [1.7, 0.6, 1.0, 2.2][k], synth.random_rotation(rng), rng.normal(0, 0.2, 3)
        # symmetric rotation (180 degrees about an axis) for case 0: R = R^T, where trans.py's transposed rotation is right
        if k == 0:
            ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
            R0 = 2 * np.outer(ax, ax) - np.eye(3)
        dst = (s0 * src @ R0.T + T0 + rng.normal(0, 0.0008, (n, 3))).astype(np.float32)
        bad = rng.random(n) < [0.2, 0.3, 0.1, 0.6][k]
        dst[bad] = (rng.random((int(bad.sum()), 3)) - 0.5).astype(np.float32)
        if k in (0, 3):                         # coordinates of tens of units: PassThreshold (~1-2) now separates inliers
            src, dst = src * np.float32(200.0), dst * np.float32(200.0)
        draws, evals = [], []
        orig_randint, orig_eval = np.random.randint, trans.evaluateModel

        def rec_randint(high, size=None):
            r = orig_randint(high, size=size)
            draws.append(np.asarray(r).copy())
            return r

        def rec_eval(M, S, T, P):
            res = orig_eval(M, S, T, P)
            evals.append((res[0], res[1], P))
            return res
        np.random.seed(500 + k)
        np.random.randint = rec_randint
        trans.evaluateModel = rec_eval
        try:
            Sc, Rot, Tr, M = trans.estimateSimilarityTransform(src.astype(np.float64), dst.astype(np.float64))
        finally:
            np.random.randint, trans.evaluateModel = orig_randint, orig_eval
        idx = np.zeros((100, 5), np.int32)
        idx[:len(draws)] = np.stack(draws)
        out[f"tr{k}_src"], out[f"tr{k}_dst"], out[f"tr{k}_idx"] = src, dst, idx
        out[f"tr{k}_niter"] = np.int32(len(draws))
        out[f"tr{k}_residual"] = np.array([e[0] for e in evals])
        out[f"tr{k}_ratio"] = np.array([e[1] for e in evals])
        out[f"tr{k}_pass_t"] = np.float64(evals[0][2])
        out[f"tr{k}_ok"] = np.int32(Sc is not None)
        if Sc is not None:
            out[f"tr{k}_S"], out[f"tr{k}_R"], out[f"tr{k}_T"], out[f"tr{k}_M"] = np.asarray(Sc), Rot, Tr, M
        n_tr += 1
        print(f"trans case {k}: iterations {len(draws)}, ok {Sc is not None}, best ratio {max(e[1] for e in evals):.3f}")
    out["tr_n"] = np.int32(n_tr)

    # ---- uvd2cloud (torch.cat along the last axis returns [bs,3n]: x block, y block, z block)
    uvd = np.stack([rng.uniform(0, 640, (2, 50)), rng.uniform(0, 480, (2, 50)), rng.uniform(0.3, 1.5, (2, 50))], axis=-1).astype(np.float32)
    K = np.array([[572.4114, 0, 325.2611], [0, 573.57043, 242.04899], [0, 0, 1]], np.float32)
    xyz = coord.uvd2cloud(torch.from_numpy(uvd), torch.from_numpy(K)).numpy()
    out["uvd"], out["uvd_K"] = uvd, K
    out["uvd_xyz"] = xyz.reshape(2, 3, 50).transpose(0, 2, 1).copy()

    np.savez_compressed(os.path.join(HERE, "f3_ref.npz"), **out)
    print("wrote f3_ref.npz", os.path.getsize(os.path.join(HERE, "f3_ref.npz")), "bytes")

    # ---- the report block of Trainer.test_epoch, executed as it stands
    tree = ast.parse(open(f"{REF}/tools/trainer.py").read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "Trainer")
    fn = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "test_epoch")
    loop = next(n for n in fn.body if isinstance(n, ast.For) and isinstance(n.iter, ast.Attribute) and n.iter.attr == "objlist"
                and "Object" in ast.unparse(n))
    k0 = fn.body.index(loop)
    stmts = [loop]
    for n in fn.body[k0 + 1:]:
        stmts.append(n)
        if "all_content" in ast.unparse(n) and isinstance(n, ast.Expr):
            break
    mod = ast.Module(body=stmts, type_ignores=[])
    ast.fix_missing_locations(mod)
    objlist = list(synth.OBJLIST)
    rr = np.random.default_rng(5)
    table = np.zeros((24, len(objlist)))
    table[0] = table[1] = rr.integers(20, 90, len(objlist))
    for v in range(3):
        for c in range(6):
            table[2 + 6 * v + c] = table[1] * (rr.random(len(objlist)) * ([1, 0.02, 1, 4.0, 1, 0.03][c]))
    table[20:23] = table[1] * rr.random((3, len(objlist))) * 0.05
    keys = ["all_num", "obj_num"] + [f"{a}_{v}_{b}" for v in ("base", "reg", "final") for (a, b) in
                                    (("succ", "rt"), ("dis", "rt"), ("succ", "r"), ("dis", "r"), ("succ", "t"), ("dis", "t"))] \
        + ["dis_mask", "dis_normal", "dis_xyz"]
    result = {k: {o: float(table[r, c]) for c, o in enumerate(objlist)} for r, k in enumerate(keys)}
    rep = {"table": table.tolist(), "objlist": objlist}
    for opt_pose in (False, True):
        lines = []

        class L:
            def info(self, s):
                lines.append(s)
        ns = dict(self=type("S", (), {"objlist": objlist})(), result=result, opt_pose=opt_pose, test_logger=L())
        exec(compile(mod, f"{REF}/tools/trainer.py", "exec"), ns)
        rep["opt_pose" if opt_pose else "base"] = lines
    json.dump(rep, open(os.path.join(HERE, "report_ref.json"), "w"), ensure_ascii=False, indent=0)
    print("wrote report_ref.json;", rep["base"][0])


if __name__ == "__main__":
    main()
