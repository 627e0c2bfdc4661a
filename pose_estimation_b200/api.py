"""Host-side mirror of the reference's call sites for the pose-recovery + evaluation path.

Same names, argument meaning and return conventions as the reference (file:line under
/root/reference), each routed to the C ABI (include/poseb200.h):

  batch_intrinsic_transform   lib/utils/utlis.py:12-34
  get_pose(pred, data)        tools/trainer.py:383-438   (Trainer.get_pose, @staticmethod)
  cal_dis(metric, r, t, d)    tools/trainer.py:370-381   (Trainer.cal_dis)
  Metric                      lib/utils/metric.py:13-84
  knn(x, y)                   train.py:126               (KeOps generic_argkmin handle)
  PoseLoss                    lib/network/loss.py:18-41
  choose_ransac_batch         lib/utils/utlis.py:55-74
  vote_translation            lib/network/krrn.py:153
  get_xyz                     lib/network/loss_utils.py:18-28 (region-coded map, forward)
  BPnP / BPnPModle            lib/network/dnn/BPnP.py:8-124
  PossLoss                    version/transparent/lib/networks/loss.py:197-267

Differences that are deliberate: every point in ``choose`` is scored instead of a
random 256 subset; the map is de-normalised in FP32 (the reference promotes to float64); RANSAC
failure yields identity/zero plus a status instead of uninitialised memory.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops, ops_bpnp, ops_metric, ops_poss, ops_sim3
from .ops import RansacConfig

DEFAULT_RANSAC = RansacConfig()


def _dev(t):
    if isinstance(t, np.ndarray):
        t = torch.from_numpy(t)
    return t.cuda(non_blocking=True) if not t.is_cuda else t


def batch_intrinsic_transform(K):
    """[bs,4] (fx,fy,cx,cy) -> [bs,3,3] and back (lib/utils/utlis.py:12-34)."""
    if K.dim() == 2 and K.size(-1) == 4:
        out = torch.zeros((K.size(0), 3, 3), dtype=K.dtype, device=K.device)
        out[:, 0, 0], out[:, 1, 1], out[:, 0, 2], out[:, 1, 2] = K[:, 0], K[:, 1], K[:, 2], K[:, 3]
        out[:, 2, 2] = 1
        return out
    if K.dim() == 3 and K.size(-1) == 3:
        return torch.stack([K[:, 0, 0], K[:, 1, 1], K[:, 0, 2], K[:, 1, 2]], dim=1)
    raise Exception('Wrong Shape!')


def get_xyz(xyz_off, region, region_point):
    """lib/network/loss_utils.py:18-28: xyz_off [bs,3,h,w] + mean_k softmax(region)[k] * region_point[k]
    (region [bs,N,h,w], region_point [bs,N,3]) -> [bs,3,h,w].  Forward only (no autograd)."""
    return ops.region_decode_map(_dev(xyz_off).float(), _dev(region).float(), _dev(region_point).float(), mode="soft")


def get_pose(pred, data, cfg: RansacConfig | None = None, return_info=False, region_mode=None):
    """Trainer.get_pose (tools/trainer.py:383-438).  pred['xyz'] [bs,3,h,w]; data['extent'],
    data['lfborder'] [bs,3]; data['choose'] [bs,1,NP] int64; data['x_map_choosed'],
    data['y_map_choosed'] [bs,NP,1]; data['intrinsic'] [bs,4].  Returns (R [bs,3,3], t [bs,3]) as
    CPU float32 tensors like the reference (which asserts bs == 1; any bs works here).
    region_mode: None = the map is pred['xyz'] as it stands (what trainer.py runs today: :400 ``xyz = xyz_off``);
    'hard' = tools/script/eval.py:93-103 / the block commented out at trainer.py:391-398 -- pred['xyz'] is an offset
    map and the class arg-max of pred['region'] [bs,N,h,w] picks the base point out of data['region_point'] [bs,N,3];
    'soft' = loss_utils.get_xyz's softmax-weighted base.  Only the chosen pixels touch the logits."""
    xyz = _dev(pred['xyz']).float()
    bs = xyz.size(0)
    choose = _dev(data['choose']).long().reshape(bs, -1)
    xm = _dev(data['x_map_choosed']).float().reshape(bs, -1)
    ym = _dev(data['y_map_choosed']).float().reshape(bs, -1)
    ext = _dev(data['extent']).float().reshape(bs, 3)
    lfb = _dev(data['lfborder']).float().reshape(bs, 3)
    K4 = _dev(data['intrinsic']).float().reshape(bs, 4)
    pts, count = ops.gather_correspondences(xyz, choose, xm, ym, ext, lfb)
    if region_mode is not None:
        ops.region_decode_points(pts, count, xyz, _dev(pred['region']).float(), _dev(data['region_point']).float(),
                                 ext, lfb, choose=choose, mode=region_mode)
    out = ops.pnp_ransac_pts(pts, count, K4, cfg or DEFAULT_RANSAC)
    R, t = out['R'].cpu(), out['t'].cpu()   # the reference returns CPU tensors (host sync here)
    if return_info:
        return R, t, {k: v.cpu() for k, v in out.items() if k not in ('R', 't')}
    return R, t


class Metric:
    """lib/utils/metric.py:13-84."""

    def __init__(self, sys):
        self.sys = list(sys)

    def cal_adds_cuda(self, pred, target, idx):
        """metric.py:17-36: returns (add_dis, adds_dis) Python floats; for a class in ``sys`` both
        are ADD-S = mean_i min_j |pred_j - target_i|, otherwise both are ADD."""
        assert pred.dim() == target.dim() == 2
        p = _dev(pred).float().unsqueeze(0)
        g = _dev(target).float().unsqueeze(0)
        sym = int(idx) in self.sys
        want = torch.full((1,), 1 if sym else 0, dtype=torch.uint8, device=p.device)
        add, adds = ops_metric.add_adds(p, g, want)
        add_v, adds_v = float(add.item()), float(adds.item())
        return (adds_v, adds_v) if sym else (add_v, add_v)

    def cal_auc(self, add_dis, max_dis=0.1):
        """metric.py:38-50 (host side: one call per object per sweep)."""
        d = np.sort(np.where(np.asarray(add_dis, dtype=np.float64) > max_dis, np.inf, add_dis))
        n = len(add_dis)
        acc = np.cumsum(np.ones((1, n)), dtype=np.float32) / n
        return self.voc_ap(d, acc) * 100.0

    @staticmethod
    def voc_ap(rec, prec):
        """metric.py:52-65: area under the monotone envelope of (rec, prec) on [0, 0.1], times 10."""
        ok = np.isfinite(rec)
        if not ok.any():
            return 0
        rec, prec = rec[ok], prec[ok]
        mrec = np.concatenate([[0.0], rec, [0.1]])
        mpre = np.concatenate([[0.0], prec, [prec[-1]]])
        # running maximum over the prefix the reference's loop covers (indices 1..len(prec)-1)
        m = prec.shape[0]
        if m > 1:
            mpre[:m] = np.maximum.accumulate(mpre[:m])
        step = np.nonzero(mrec[1:] != mrec[:-1])[0] + 1
        return np.sum((mrec[step] - mrec[step - 1]) * mpre[step]) * 10

    @staticmethod
    def angular_distance(q1, q2, eps: float = 1e-7):
        """metric.py:67-80 for rotation matrices [B,3,3] -> degrees [B,1] (pairwise diagonal)."""
        R1, R2 = _dev(q1).float().reshape(-1, 3, 3), _dev(q2).float().reshape(-1, 3, 3)
        B = R1.size(0)
        z = torch.zeros((B, 3), device=R1.device)
        model = torch.zeros((1, 1, 3), device=R1.device)
        cls = torch.zeros((B,), dtype=torch.int32, device=R1.device)
        _, r_deg, _ = ops_metric.pose_metrics(R1, z, R2, z, model, cls, None)
        return r_deg.reshape(B, 1)

    @staticmethod
    def translation_distance(t1, t2):
        """metric.py:82-84."""
        a, b = _dev(t1).float().reshape(-1, 3), _dev(t2).float().reshape(-1, 3)
        B = a.size(0)
        eye = torch.eye(3, device=a.device).expand(B, 3, 3).contiguous()
        model = torch.zeros((1, 1, 3), device=a.device)
        cls = torch.zeros((B,), dtype=torch.int32, device=a.device)
        _, _, t_m = ops_metric.pose_metrics(eye, a, eye, b, model, cls, None)
        return t_m


class TransparentMetric(Metric):
    """version/transparent/lib/utils.py:12-46 -- the older branch's ``Metric``: ``cal_dis`` matches every PREDICTED
    point to its nearest TARGET point (KeOps arg-min, then gather), i.e. ADD-S taken in the direction opposite to
    ``cal_adds_cuda``; both methods always compute both distances and return ADD-S twice for a class in ``sys``."""

    def cal_dis(self, pred, target, idx):
        """utils.py:17-30: (add_dis, adds_dis) with adds = mean_i min_j |pred_i - target_j|."""
        p = _dev(pred).float().unsqueeze(0).contiguous()
        g = _dev(target).float().unsqueeze(0).contiguous()
        one = torch.ones((1,), dtype=torch.uint8, device=p.device)
        add, _ = ops_metric.add_adds(p, g, torch.zeros_like(one))
        _, adds = ops_metric.add_adds(g, p, one)          # roles swapped: the minimum runs over the target points
        add_v, adds_v = float(add.item()), float(adds.item())
        return (adds_v if int(idx) in self.sys else add_v), adds_v

    def cal_adds_cuda(self, pred, target, idx):
        """utils.py:32-46: like lib/utils/metric.py's, but ADD-S is computed for every class."""
        p = _dev(pred).float().unsqueeze(0).contiguous()
        g = _dev(target).float().unsqueeze(0).contiguous()
        one = torch.ones((1,), dtype=torch.uint8, device=p.device)
        add, adds = ops_metric.add_adds(p, g, one)
        add_v, adds_v = float(add.item()), float(adds.item())
        return (adds_v if int(idx) in self.sys else add_v), adds_v


def cal_dis(metric: Metric, pred_r, pred_t, datas):
    """Trainer.cal_dis (tools/trainer.py:370-381) -> (add_dis, r_dis, t_dis) Python floats."""
    R = _dev(pred_r).float().reshape(1, 3, 3)
    t = _dev(pred_t).float().reshape(1, 3)
    Rg = _dev(datas['target_r']).float().reshape(1, 3, 3)
    tg = _dev(datas['target_t']).float().reshape(1, 3)
    model = _dev(datas['model_points']).float().reshape(1, -1, 3)
    target = _dev(datas['target']).float().reshape(1, -1, 3)
    cls_id = int(torch.as_tensor(datas['cls_id']).reshape(-1)[0].item())
    is_sym = torch.full((1,), 1 if cls_id in metric.sys else 0, dtype=torch.uint8, device=R.device)
    cls = torch.zeros((1,), dtype=torch.int32, device=R.device)
    add, r_deg, t_m = ops_metric.pose_metrics(R, t, Rg, tg, model, cls, is_sym, target)
    vals = torch.stack([add[0], r_deg[0], t_m[0]]).cpu()    # one host sync instead of three
    return float(vals[0]), float(vals[1]), float(vals[2])


def knn(x, y):
    """The KeOps handle of train.py:126: x [N,3], y [M,3] contiguous f32 CUDA -> int64 [N,1]."""
    if x.dim() != 2 or y.dim() != 2 or x.size(-1) != 3 or y.size(-1) != 3:
        raise ValueError("knn expects x [N,3] and y [M,3]")
    if x.device != y.device:
        raise ValueError("knn: x and y must be on the same device")
    return ops_metric.knn1(x, y).view(-1, 1)


class _PoseLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred_r, pred_t, targets, model_points, sym, sym_host=None):
        loss_b, pred, nn_idx = ops_metric.pose_loss_fwd(pred_r.contiguous(), pred_t.reshape(-1, 3).contiguous(),
                                                        model_points.contiguous(), targets.contiguous(), sym,
                                                        sym_host=sym_host)
        ctx.save_for_backward(pred, model_points, targets, nn_idx)
        ctx.t_shape = pred_t.shape
        return loss_b.mean()

    @staticmethod
    def backward(ctx, gout):
        pred, model_points, targets, nn_idx = ctx.saved_tensors
        gR, gt = ops_metric.pose_loss_bwd(pred, model_points.contiguous(), targets.contiguous(), nn_idx, 1.0)
        return gR * gout, (gt * gout).reshape(ctx.t_shape), None, None, None, None


class PoseLoss(torch.nn.Module):
    """lib/network/loss.py:18-41.  ``knn`` is accepted for signature compatibility; the fused
    kernel performs the same argmin."""

    def __init__(self, sym_list, knn=None, num_point=1000):
        super().__init__()
        self.num_point = num_point
        self.sym_list = list(sym_list)
        self.knn = knn

    def forward(self, pred_r, pred_t, targets, model_points, idxs):
        ids = torch.as_tensor(idxs).reshape(-1).tolist()
        flags = [1 if int(i) in self.sym_list else 0 for i in ids]
        sym = torch.tensor(flags, dtype=torch.uint8, device=pred_r.device)
        return _PoseLossFn.apply(pred_r.float(), pred_t.float(), targets.float(), model_points.float(), sym, flags)


class _VoteFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, p_emb, t_res):
        ctx.np_ = p_emb.size(1)
        return ops_metric.vote_translation(p_emb.contiguous(), t_res.contiguous())

    @staticmethod
    def backward(ctx, g):
        gn = g / ctx.np_
        return gn.unsqueeze(1).expand(-1, ctx.np_, -1), gn.unsqueeze(2).expand(-1, -1, ctx.np_)


def vote_translation(p_emb, t_res):
    """pred_t = (p_emb + t_res.permute(0, 2, 1)).mean(dim=1)  (lib/network/krrn.py:153)."""
    return _VoteFn.apply(p_emb.float(), t_res.float())


def choose_ransac_batch(coor_emb, uv_emb, K_matrix, n=125, cfg: RansacConfig | None = None):
    """lib/utils/utlis.py:55-74: numpy in, numpy int32 [bs,1,n] out (indices of RANSAC inliers,
    order-preserving random subset if more than n, wrap-padded if fewer)."""
    cfg = cfg or DEFAULT_RANSAC
    obj = torch.as_tensor(np.asarray(coor_emb, np.float32)).cuda()
    uv = torch.as_tensor(np.asarray(uv_emb, np.float32)).cuda()
    K = torch.as_tensor(np.asarray(K_matrix, np.float32)).cuda()
    bs, m, _ = obj.shape
    pts = torch.cat([obj.permute(0, 2, 1), uv.permute(0, 2, 1)], dim=1).contiguous()   # [bs,5,m]
    count = torch.full((bs,), m, dtype=torch.int32, device=pts.device)
    out = ops.pnp_ransac_pts(pts, count, batch_intrinsic_transform(K).contiguous(), cfg, want_bitmap=True)
    sel = ops.resample_inliers(out['inl_bitmap'], count, m, n, seed=cfg.seed)
    return sel.cpu().numpy().astype(np.int32).reshape(bs, 1, n)


# ---- lib/transform/umeyama.py ---------------------------------------------------------------------------
def _sim3_out(scale, R, t):
    """(Scale, Rotation, Translation, OutTransform) as the reference returns them (float64 numpy)."""
    S = float(scale)
    Rn = R.double().cpu().numpy()
    Tn = t.double().cpu().numpy()
    M = np.identity(4)
    M[:3, :3] = S * Rn
    M[:3, 3] = Tn
    return S, Rn, Tn, M


def estimateSimilarityTransform(source, target, verbose=False, cfg: ops_sim3.Sim3Config | None = None, device="cuda"):
    """lib/transform/umeyama.py:8-62: RANSAC similarity fit target ~ Scale * Rotation @ source + Translation.
    (cfg.rule = 'trans' runs the older rule of lib/transform/trans.py instead; see trans_estimateSimilarityTransform.)
    source / target: (N,3) arrays.  Returns (Scale, Rotation, Translation, OutTransform), or four ``None`` when the
    best inlier ratio is below 0.1 (the reference's failure return, :50-52).  The 5-point draws come from the
    counter RNG of ``cfg.seed`` instead of numpy's global state."""
    src = torch.as_tensor(np.asarray(source), dtype=torch.float32, device=device).reshape(1, -1, 3).contiguous()
    dst = torch.as_tensor(np.asarray(target), dtype=torch.float32, device=device).reshape(1, -1, 3).contiguous()
    if src.shape != dst.shape:
        raise AssertionError("Source and Target must have same number of points.")      # umeyama.py:13
    res = ops_sim3.sim3_ransac(src, dst, cfg=cfg)
    if int(res.status.item()) != 0:
        if verbose:
            print("[ WARN ] - Something is wrong. Small BestInlierRatio")
        return None, None, None, None
    return _sim3_out(res.scale[0], res.R[0], res.t[0])


def trans_estimateSimilarityTransform(source, target, verbose=False, cfg: ops_sim3.Sim3Config | None = None, device="cuda"):
    """lib/transform/trans.py:9-45 (the copy tools/script/eval.py imports): 100 iterations, the hypothesis with the
    smallest total residual wins, thresholds from the clouds' mean norms.  Returns what trans.py returns, quirks
    included: Scales as a length-3 array, Rotation = the TRANSPOSE of the rotation that maps source to target,
    OutTransform[:3,:3] = diag(Scales) @ Rotation, or four ``None`` below an inlier ratio of 0.1."""
    cfg = cfg or ops_sim3.Sim3Config(n_iter=100, rule="trans")
    if cfg.rule != "trans":
        raise ValueError("trans_estimateSimilarityTransform needs cfg.rule == 'trans'")
    src = torch.as_tensor(np.asarray(source), dtype=torch.float32, device=device).reshape(1, -1, 3).contiguous()
    dst = torch.as_tensor(np.asarray(target), dtype=torch.float32, device=device).reshape(1, -1, 3).contiguous()
    res = ops_sim3.sim3_ransac(src, dst, cfg=cfg)
    if int(res.status.item()) != 0:
        if verbose:
            print("[ WARN ] - Something is wrong. Small BestInlierRatio")
        return None, None, None, None
    S, Rn, Tn, M = _sim3_out(res.scale[0], res.R[0], res.t[0])
    return np.array([S, S, S]), Rn, Tn, M


def best_batch_fit_transform(A, B):
    """version/transparent/lib/declarative/leastsquares.py:10-48: least-squares rigid transform mapping A onto B,
    A, B [c,N,3] -> (R [c,3,3], t [c,3], centroid_A [c,3], centroid_B [c,3]) on the device.  (The reference subtracts
    its [c,3] centroids from [c,N,3] without a singleton axis, so it only runs for c == 1 or c == N; this is the
    per-problem fit it means.)"""
    R, t, _, cA, cB = ops_sim3.kabsch_fit(_dev(A).float(), _dev(B).float(), want_centroids=True)
    return R, t, cA, cB


def icp_refine(model_points, target, final_r, final_t, steps=10, compose="reference"):
    """The optimisation loop of version/transparent/scripts/train_declarative.py:80-100 for all hypotheses at once
    (transform -> knn -> gather -> best_batch_fit_transform -> compose); see ops_sim3.icp_refine."""
    R, t, _ = ops_sim3.icp_refine(_dev(model_points).float(), _dev(target).float(), _dev(final_r).float(),
                                  _dev(final_t).float(), n_steps=steps, compose=compose)
    return R, t


def uvd2cloud(uvd, K):
    """lib/transform/coordinate.py:91-101 -> xyz [bs,n,3] (the docstring's shape; the reference's torch.cat along the
    last axis returns [bs,3n])."""
    return ops_sim3.uvd2cloud(_dev(uvd).float(), _dev(K).float())


def estimateSimilarityUmeyama(SourceHom, TargetHom, device="cuda"):
    """lib/transform/umeyama.py:65-97: least-squares similarity over all columns of the (4,N) homogeneous arrays."""
    s = np.ascontiguousarray(np.asarray(SourceHom)[:3].T)
    t = np.ascontiguousarray(np.asarray(TargetHom)[:3].T)
    src = torch.as_tensor(s, dtype=torch.float32, device=device).reshape(1, -1, 3)
    dst = torch.as_tensor(t, dtype=torch.float32, device=device).reshape(1, -1, 3)
    scale, R, tt, status = ops_sim3.umeyama_fit(src, dst)
    if int(status.item()) != 0:
        raise RuntimeError("estimateSimilarityUmeyama: degenerate input")
    return _sim3_out(scale[0], R[0], tt[0])


# ---- lib/network/dnn/BPnP.py -----------------------------------------------------------------------------
class BPnP(torch.autograd.Function):
    """Back-propagatable PnP (BPnP.py:8-117).  pts2d [bs,n,2], pts3d [n,3], K [3,3] -> P_6d [bs,6] (angle-axis,
    translation), differentiable wrt pts2d, pts3d and K through the implicit function theorem.  Forward: RANSAC
    initial pose (3 px) unless ini_pose is given, then Levenberg-Marquardt on all points (pb_pnp_refine); backward:
    pb_bpnp_backward.  pts3d may also be [bs,n,3] (one 3-D set per sample; the docstring's BPnP_m3d case).
    Everything stays on the device: no per-sample loop, no host copy."""

    @staticmethod
    def forward(ctx, pts2d, pts3d, K, ini_pose=None):
        x = pts2d.detach().float().contiguous()
        z = pts3d.detach().float().contiguous()
        Kc = K.detach().float().contiguous()
        ini = None if ini_pose is None else ini_pose.detach().float().contiguous()
        P_6d, status = ops_bpnp.bpnp_forward(x, z, Kc, ini)
        ctx.save_for_backward(x, P_6d, z, Kc)
        ctx.status = status
        return P_6d

    @staticmethod
    def backward(ctx, grad_output):
        x, P_6d, z, Kc = ctx.saved_tensors
        gx, gz, gK, _ = ops_bpnp.bpnp_backward(x, z, Kc, P_6d, grad_output.float().contiguous())
        return gx, gz, gK, None


class BPnPModle(torch.nn.Module):
    """BPnP.py:120-124 (the reference's spelling)."""

    def forward(self, pts2d, pts3d, K, ini_pose=None):
        return BPnP.apply(pts2d, pts3d, K, ini_pose)


# ---- version/transparent/lib/networks/loss.py -------------------------------------------------------------
class _PossLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred_r, pred_t, pred_c, target, model_points, axis, target_r, nn_mode, w):
        args = [a.detach().float().contiguous() for a in (pred_r, pred_t, pred_c, target, model_points, axis, target_r)]
        vals, ws = ops_poss.poss_loss_fwd(args[0], args[1], args[2], args[3], args[4], nn_mode, w, args[5], args[6])
        ctx.save_for_backward(*args, ws)
        ctx.nn_mode, ctx.w = nn_mode, w
        return vals[0], vals[1], vals[2]

    @staticmethod
    def backward(ctx, g_loss, g_dis, g_rot):
        pr, pt, pc, tg, mp, ax, tr, ws = ctx.saved_tensors
        z = torch.zeros((), dtype=torch.float32, device=pr.device)
        g = torch.stack([z if x is None else x.float() for x in (g_loss, g_dis, g_rot)])
        gr, gt, gc = ops_poss.poss_loss_bwd(pr, pt, pc, tg, mp, ctx.nn_mode, ctx.w, ax, tr, g, ws)
        return gr, gt, gc, None, None, None, None, None, None


class PossLoss(torch.nn.Module):
    """version/transparent/lib/networks/loss.py:197-251.  Same constructor and forward arguments; returns
    (loss, dis[0][argmax pred_c[0]], loss_rotation) as 0-dim tensors with grad wrt pred_r, pred_t, pred_c.  ``knn`` is
    accepted for signature compatibility; the nearest-neighbour search is fused into the kernel.  ``idx`` should be a
    Python int (a tensor costs one host sync, as ``idx in self.sym_list`` does in the reference)."""

    def __init__(self, num_point_mesh, sym_list, knn=None):
        super().__init__()
        self.num_point_mesh = num_point_mesh
        self.sym_list = [int(s) for s in sym_list]

    def forward(self, pred_r, pred_t, pred_c, target, model_points, idx, refine, w, axis, target_r):
        nn_mode = (not refine) and (int(idx) in self.sym_list)
        return _PossLossFn.apply(pred_r, pred_t, pred_c, target, model_points, axis, target_r, bool(nn_mode), float(w))

    @staticmethod
    def predr2rotation(pred_r, bs, num_p):
        return _predr2rotation(pred_r, bs, num_p)


def _predr2rotation(pred_r, bs, num_p):
    """loss.py:253-267 (used on its own by scripts/train_declarative.py:74): [bs,num_p,4] -> [bs*num_p,3,3]."""
    r = pred_r / torch.norm(pred_r, dim=2).view(bs, num_p, 1)
    r0, r1, r2, r3 = r[..., 0], r[..., 1], r[..., 2], r[..., 3]
    rows = [1.0 - 2.0 * (r2 ** 2 + r3 ** 2), 2.0 * r1 * r2 - 2.0 * r0 * r3, 2.0 * r0 * r2 + 2.0 * r1 * r3,
            2.0 * r1 * r2 + 2.0 * r3 * r0, 1.0 - 2.0 * (r1 ** 2 + r3 ** 2), -2.0 * r0 * r1 + 2.0 * r2 * r3,
            -2.0 * r0 * r2 + 2.0 * r1 * r3, 2.0 * r0 * r1 + 2.0 * r2 * r3, 1.0 - 2.0 * (r1 ** 2 + r2 ** 2)]
    return torch.stack(rows, dim=2).contiguous().view(bs * num_p, 3, 3)
