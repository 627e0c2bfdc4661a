"""Generate tests/golden/callsites_ref.npz by RUNNING THE REFERENCE'S OWN CODE for the main path:

  Trainer.get_pose / Trainer.cal_dis   /root/reference/tools/trainer.py:370-438
  Metric (all methods)                  /root/reference/lib/utils/metric.py
  batch_intrinsic_transform,
  choose_ransac_batch                   /root/reference/lib/utils/utlis.py
  PoseLoss.forward (+ autograd)         /root/reference/lib/network/loss.py:18-41
  Metric.cal_dis / cal_adds_cuda        /root/reference/version/transparent/lib/utils.py:12-46 (the older branch)
  get_xyz (region-coded map, soft)      /root/reference/lib/network/loss_utils.py:18-28
  region arg-max decode (hard)          /root/reference/tools/script/eval.py:93-103 (the lines as they stand)

``utlis.py`` imports cleanly.  ``metric.py`` / ``loss.py`` import kornia / mmcv (absent here) at module level,
so stub modules are put into ``sys.modules`` first; ``tools/trainer.py`` cannot be imported at all (it pulls in
the network, whose files are missing from the reference), so the two ``Trainer`` methods are taken out of the
file's syntax tree and compiled as they stand, with the reference's own ``batch_intrinsic_transform``, the real
``cv2`` and ``torch`` in their namespace.  What is NOT the reference's code in these runs (and therefore stays
unpinned): kornia's ``angle_axis_to_rotation_matrix`` (oracle/bpnp_ref.py restates the published formula) and
``rotation_matrix_to_quaternion`` (trace-based, float32), the KeOps arg-min handle (float32 brute force, lowest
index on ties), and ``Tensor.cuda()`` (identity: there is no GPU here; the arithmetic after it is torch's).
Frames are stored sparsely.  Run here only; the GPU box has no /root/reference.

    python tests/golden/gen_golden_callsites.py
"""
import ast
import importlib.util
import os
import sys
import types

import cv2
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import bpnp_ref  # noqa: E402
from pose_estimation_b200 import synth  # noqa: E402

REFROOT = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))
N_MODEL = 500


# ------------------------------------------------------------------ stand-ins for the absent packages
def rotation_matrix_to_quaternion(R, order=None, eps=1e-8):
    """kornia's trace-based conversion (w, x, y, z), float32 like the reference's tensors."""
    R = R.reshape(-1, 3, 3)
    out = []
    for m in R:
        m00, m01, m02, m10, m11, m12, m20, m21, m22 = [m[i, j] for i in range(3) for j in range(3)]
        tr = m00 + m11 + m22
        if tr > 0:
            s = torch.sqrt(tr + 1.0 + eps) * 2
            q = [0.25 * s, (m21 - m12) / s, (m02 - m20) / s, (m10 - m01) / s]
        elif m00 > m11 and m00 > m22:
            s = torch.sqrt(1.0 + m00 - m11 - m22 + eps) * 2
            q = [(m21 - m12) / s, 0.25 * s, (m01 + m10) / s, (m02 + m20) / s]
        elif m11 > m22:
            s = torch.sqrt(1.0 + m11 - m00 - m22 + eps) * 2
            q = [(m02 - m20) / s, (m01 + m10) / s, 0.25 * s, (m12 + m21) / s]
        else:
            s = torch.sqrt(1.0 + m22 - m00 - m11 + eps) * 2
            q = [(m10 - m01) / s, (m02 + m20) / s, (m12 + m21) / s, 0.25 * s]
        out.append(torch.stack(q))
    return torch.stack(out)


def knn_handle(x, y):
    """generic_argkmin('SqDist(x, y)', 'a = Vi(1)', 'x = Vi(3)', 'y = Vj(3)') (train.py:126): int64 [N,1]."""
    d = x[:, None, :] - y[None, :, :]
    d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
    return d2.argmin(dim=1, keepdim=True)


kn = types.ModuleType("kornia")
kn.geometry = types.ModuleType("kornia.geometry")
kn.geometry.conversions = types.ModuleType("kornia.geometry.conversions")
kn.geometry.conversions.angle_axis_to_rotation_matrix = bpnp_ref.angle_axis_to_rotation_matrix
kn.geometry.conversions.rotation_matrix_to_quaternion = rotation_matrix_to_quaternion
kn.geometry.conversions.QuaternionCoeffOrder = types.SimpleNamespace(WXYZ="wxyz", XYZW="xyzw")
sys.modules["kornia"] = kn
sys.modules["kornia.geometry"] = kn.geometry
sys.modules["kornia.geometry.conversions"] = kn.geometry.conversions
mmcv = types.ModuleType("mmcv")
mmcv.Config = types.SimpleNamespace(fromfile=lambda *_a, **_k: None)   # loss.py:15 loads a config file that does not exist
sys.modules["mmcv"] = mmcv
pk = types.ModuleType("pykeops")
pk.torch = types.ModuleType("pykeops.torch")
pk.torch.generic_argkmin = lambda *_a, **_k: knn_handle                  # version/transparent/lib/utils.py:15
sys.modules["pykeops"] = pk
sys.modules["pykeops.torch"] = pk.torch
sys.modules["cv2.cv2"] = cv2                                             # utils.py:5 (`import cv2.cv2 as cv2`, old wheels)
torch.Tensor.cuda = lambda self, *a, **k: self                           # metric.py:28-29 on a box without a GPU


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


sys.path.insert(0, REFROOT)
ref_utlis = load(f"{REFROOT}/lib/utils/utlis.py", "ref_utlis")
ref_metric = load(f"{REFROOT}/lib/utils/metric.py", "ref_metric")
ref_loss = load(f"{REFROOT}/lib/network/loss.py", "ref_loss")
ref_tutils = load(f"{REFROOT}/version/transparent/lib/utils.py", "ref_tutils")


def trainer_methods():
    """get_pose and cal_dis exactly as they stand in tools/trainer.py, compiled from the file's syntax tree."""
    src = open(f"{REFROOT}/tools/trainer.py").read()
    tree = ast.parse(src)
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "Trainer")
    keep = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in ("get_pose", "cal_dis")]
    assert len(keep) == 2
    mod = ast.Module(body=[ast.ClassDef(name="Trainer", bases=[], keywords=[], body=keep, decorator_list=[])],
                     type_ignores=[])
    ast.fix_missing_locations(mod)
    ns = dict(torch=torch, cv2=cv2, np=np, kn=kn, batch_intrinsic_transform=ref_utlis.batch_intrinsic_transform)
    exec(compile(mod, f"{REFROOT}/tools/trainer.py", "exec"), ns)
    return ns["Trainer"]


Trainer = trainer_methods()


def torch_frame(fr):
    """the ``pred`` / ``datas`` dicts of tools/trainer.py (shapes and dtypes of dataset/linemod/batchdataset.py)."""
    NP = fr["choose"].shape[0]
    pred = dict(xyz=torch.from_numpy(fr["xyz"])[None], region=torch.zeros(1))
    data = dict(extent=torch.from_numpy(np.asarray(fr["extent"], np.float64))[None],
                lfborder=torch.from_numpy(np.asarray(fr["lfborder"], np.float64))[None],
                choose=torch.from_numpy(fr["choose"]).view(1, 1, NP),
                x_map_choosed=torch.from_numpy(fr["x_map_choosed"]).view(1, NP, 1),
                y_map_choosed=torch.from_numpy(fr["y_map_choosed"]).view(1, NP, 1),
                intrinsic=torch.from_numpy(fr["intrinsic"]).view(1, 4),
                target_r=torch.from_numpy(np.asarray(fr["target_r"], np.float32))[None],
                target_t=torch.from_numpy(np.asarray(fr["target_t"], np.float32))[None],
                target=torch.from_numpy(fr["target"])[None],
                cls_id=torch.tensor([int(fr["cls_id"])]),
                model_points=torch.from_numpy(fr["model_points"])[None])
    return pred, data


def main():
    models = synth.class_models(N_MODEL)
    out = {}
    me = types.SimpleNamespace(metric=ref_metric.Metric(list(synth.SYM_CLS)))
    for tag, cfgid, n, kw in (("clean", 201, 13, dict(noise_m=0.0, outlier_min_px=5.0)), ("noisy", 202, 13, dict())):
        Rs, ts, picks, dis = [], [], [], []
        for f in range(n):
            sp = synth.make_frame_sparse(0x5EED, cfgid, f, **kw)
            for k in ("rows", "cols", "xyzn", "R_gt", "t_gt"):
                out[f"{tag}_f{f}_{k}"] = sp[k]
            out[f"{tag}_f{f}_cls"] = np.int32(sp["cls"])
            fr = synth.reference_frame_dict(sp, models[sp["cls"]])
            pred, data = torch_frame(fr)
            seed = 1000 * cfgid + f
            torch.manual_seed(seed)
            picks.append(torch.randperm(fr["choose"].shape[0])[:256].numpy())      # what trainer.py:407 will draw
            torch.manual_seed(seed)
            R, t = Trainer.get_pose(pred, data)                                     # the reference's method
            add, r_dis, t_dis = Trainer.cal_dis(me, R, t, data)                     # the reference's method
            Rs.append(R[0].numpy()); ts.append(t[0].numpy()); dis.append([add, r_dis, t_dis])
        out[f"{tag}_n"] = np.int32(n)
        out[f"{tag}_pick"] = np.stack(picks)
        out[f"{tag}_R"] = np.stack(Rs)
        out[f"{tag}_t"] = np.stack(ts)
        out[f"{tag}_dis"] = np.asarray(dis, np.float64)

    # ---- Metric on its own (symmetric and not), AUC
    rng = np.random.default_rng(21)
    M = ref_metric.Metric([7, 8])
    for cls, N in ((7, 600), (8, 2620), (3, 500)):
        model = (synth.LFBORDER[cls] + rng.random((N, 3)) * synth.EXTENT[cls]).astype(np.float32)
        Rg, dR = synth.random_rotation(rng).astype(np.float32), synth.random_rotation(rng)
        ang = np.radians(rng.uniform(1.0, 8.0))
        ax = dR[:, 0]
        Kx = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
        R = ((np.eye(3) + np.sin(ang) * Kx + (1 - np.cos(ang)) * Kx @ Kx) @ Rg).astype(np.float32)
        tg = np.array([0.02, -0.03, 0.8], np.float32)
        t = tg + rng.normal(0, 0.004, 3).astype(np.float32)
        pred = torch.from_numpy(model @ R.T + t)
        tgt = torch.from_numpy(model @ Rg.T + tg)
        add, adds = M.cal_adds_cuda(pred, tgt, cls)
        r = M.angular_distance(torch.from_numpy(R)[None], torch.from_numpy(Rg)[None]).squeeze().item()
        td = M.translation_distance(torch.from_numpy(t)[None], torch.from_numpy(tg)[None]).squeeze().item()
        out[f"m{cls}_model"] = model; out[f"m{cls}_R"] = R; out[f"m{cls}_t"] = t
        out[f"m{cls}_Rg"] = Rg; out[f"m{cls}_tg"] = tg
        out[f"m{cls}_vals"] = np.array([add, adds, r, td], np.float64)
        TM = ref_tutils.Metric([7, 8])                       # the older branch's Metric on the same clouds
        out[f"m{cls}_tvals"] = np.array([*TM.cal_dis(pred, tgt, cls), *TM.cal_adds_cuda(pred, tgt, cls)], np.float64)
    d = np.abs(rng.normal(0.03, 0.04, 200))
    out["auc_in"] = d
    out["auc_out"] = np.float64(M.cal_auc(list(d)))
    out["auc_none"] = np.float64(M.cal_auc([0.5, 0.7]))
    # more AUC cases: ties, values on the 0.1 border, a single sample, mostly-out-of-range lists
    rng_auc = np.random.default_rng(77)                      # its own stream: the vectors made after this keep their inputs
    cases = [np.array([0.01, 0.01, 0.02, 0.02, 0.02, 0.09]), np.array([0.1, 0.1, 0.05]), np.array([0.03]),
             np.concatenate([np.full(50, 0.5), np.abs(rng_auc.normal(0.02, 0.01, 10))]),
             np.sort(rng_auc.random(500) * 0.2)[::-1].copy()]
    for k, c in enumerate(cases):
        out[f"auc_case{k}_in"] = c
        out[f"auc_case{k}_out"] = np.float64(M.cal_auc(list(c)))
    out["auc_ncases"] = np.int32(len(cases))

    # ---- utlis.py
    K4 = torch.tensor([[572.4114, 573.57043, 325.2611, 242.04899], [700.0, 700.0, 320.0, 240.0]])
    K33 = ref_utlis.batch_intrinsic_transform(K4)
    out["K4"] = K4.numpy(); out["K33"] = K33.numpy()
    out["K4_back"] = ref_utlis.batch_intrinsic_transform(K33).numpy()
    coor, uv, Ks = [], [], []
    for f in range(4):                                      # 300 correspondences per element, two of them clean
        sp = synth.make_frame_sparse(0x5EED, 203, f, **(dict(noise_m=0.0, outlier_min_px=5.0) if f < 2 else {}))
        pts = synth.compact_points(sp)[:, :300]
        coor.append(pts[:3].T.astype(np.float64)); uv.append(pts[3:].T.astype(np.float32))
        Ks.append(K33[0].numpy())
        out[f"crb_truth{f}"] = np.nonzero(sp["inlier"][:300])[0].astype(np.int32)    # the generator's true inliers
    coor, uv, Ks = np.stack(coor), np.stack(uv), np.stack(Ks)
    out["crb_coor"] = coor; out["crb_uv"] = uv; out["crb_K"] = Ks
    np.random.seed(77)                                      # subset branch (every element has > 125 inliers)
    out["crb_out125"] = ref_utlis.choose_ransac_batch(coor, uv, Ks, n=125)
    # wrap-pad branch: utlis.py:70 pads BOTH axes of the [M,1] inlier array, so an element comes out as
    # [1 + n - M, n] (identical rows) and np.stack fails unless all elements share M -- run per element.
    for j in range(4):
        np.random.seed(78)
        out[f"crb_wrap{j}"] = ref_utlis.choose_ransac_batch(coor[j:j + 1], uv[j:j + 1], Ks[j:j + 1], n=400)

    # ---- PoseLoss forward + autograd (bs = 3: symmetric, not symmetric, symmetric)
    PL = ref_loss.PoseLoss([7, 8], knn_handle)
    N = 700
    idxs = torch.tensor([[7], [2], [8]])
    mp_, tg_, Rp, tp = [], [], [], []
    for b, cls in enumerate([7, 2, 8]):
        model = (synth.LFBORDER[cls] + rng.random((N, 3)) * synth.EXTENT[cls]).astype(np.float32)
        Rg = synth.random_rotation(rng).astype(np.float32)
        tgv = np.array([0.01 * b, -0.02, 0.7 + 0.1 * b], np.float32)
        w = rng.normal(size=3); w *= np.radians(6.0) / np.linalg.norm(w)
        dR, _ = cv2.Rodrigues(w)
        mp_.append(model); tg_.append((model @ Rg.T + tgv).astype(np.float32))
        Rp.append((dR @ Rg).astype(np.float32)); tp.append((tgv + rng.normal(0, 0.005, 3)).astype(np.float32))
    pred_r = torch.from_numpy(np.stack(Rp)).requires_grad_(True)
    pred_t = torch.from_numpy(np.stack(tp))[:, None, :].clone().requires_grad_(True)
    loss = PL(pred_r, pred_t, torch.from_numpy(np.stack(tg_)), torch.from_numpy(np.stack(mp_)), idxs)
    loss.backward()
    out["pl_model"] = np.stack(mp_); out["pl_target"] = np.stack(tg_)
    out["pl_R"] = pred_r.detach().numpy(); out["pl_t"] = pred_t.detach().numpy(); out["pl_idxs"] = idxs.numpy()
    out["pl_loss"] = np.float64(loss.item())
    out["pl_grad_R"] = pred_r.grad.numpy(); out["pl_grad_t"] = pred_t.grad.numpy()

    # ---- region-coded maps: loss_utils.get_xyz (soft) and the block of tools/script/eval.py:93-103 (hard) as it stands
    bs, n, h, w = 2, 65, 24, 24
    xyz_off = (rng.normal(0, 0.05, (bs, 3, h, w))).astype(np.float32)
    logits = (rng.normal(0, 3.0, (bs, n, h, w))).astype(np.float32)
    logits[0, 7, 3, 4] = logits[0, 9, 3, 4] = 20.0                             # an exact tie: the lower class wins
    rpts = rng.random((bs, n, 3)).astype(np.float32)
    rpts[:, 0] = 0.0                                                            # batchdataset.py:728: row 0 is the background
    out["reg_xyz_off"] = xyz_off; out["reg_logits"] = logits; out["reg_points"] = rpts
    get_xyz = sys.modules["lib.network.loss_utils"].get_xyz
    out["reg_soft"] = get_xyz(torch.from_numpy(xyz_off), torch.from_numpy(logits), torch.from_numpy(rpts)).numpy()
    src = open(f"{REFROOT}/tools/script/eval.py").read().split("\n")
    a = next(i for i, ln in enumerate(src) if ln.strip() == "region = pred['region'].cpu()")
    b = next(i for i, ln in enumerate(src) if ln.strip() == "xyz = xyz_off + xyz_base.squeeze(1)")
    assert 0 < b - a < 15
    import textwrap
    ns = dict(torch=torch, pred=dict(region=torch.from_numpy(logits), xyz=torch.from_numpy(xyz_off)),
              datas=dict(region_point=torch.from_numpy(rpts)))
    exec(compile(textwrap.dedent("\n".join(src[a:b + 1])), f"{REFROOT}/tools/script/eval.py", "exec"), ns)
    out["reg_hard"] = ns["xyz"].numpy()

    np.savez_compressed(os.path.join(OUT, "callsites_ref.npz"), **out)
    print("wrote callsites_ref.npz:", os.path.getsize(os.path.join(OUT, "callsites_ref.npz")), "bytes")
    for tag in ("clean", "noisy"):
        print(tag, "add / r / t of the first frames:", out[f"{tag}_dis"][:3].round(5).tolist())


if __name__ == "__main__":
    main()
