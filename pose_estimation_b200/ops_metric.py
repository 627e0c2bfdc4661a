"""Tensor-level wrappers (continued): KNN / ADD(-S), per-frame metrics, accuracy table, pose loss,
T-vote.  Same rules as ops.py: CUDA tensors only, asynchronous on the current stream."""
from __future__ import annotations

import torch

from ._lib import call
from .ops import _chk, _ptr, _stream

TABLE_ROWS = 24
ROW_ALL_NUM, ROW_OBJ_NUM, ROW_VARIANT0 = 0, 1, 2
ROW_DIS_MASK, ROW_DIS_NORMAL, ROW_DIS_XYZ, ROW_FAILED = 20, 21, 22, 23
VARIANTS = ("base", "reg", "final")
VARIANT_COLS = ("succ_{}_rt", "dis_{}_rt", "succ_{}_r", "dis_{}_r", "succ_{}_t", "dis_{}_t")


TC_MIN_PAIRS = 1 << 18     # below this many pairs per problem the SIMT kernel wins (tile padding, prepare pass)
TC_MIN_DB = 768            # ... and below this many database points (a CTA sweeps too few tiles to pay for its set-up)


LAST_TC_STATS = None       # device int32 [4] of the last tensor-core call: timeout flag, fallback rows, exact evaluations


def set_knn_tc_impl(sweep="auto"):
    """Which sweep the tensor-core nearest-neighbour entry points use (process-wide): 'auto' (= 'tcgen05'), 'tcgen05',
    'mma' (warp-level mma.sync: a measured alternative, slower on B200).  Identical results."""
    call("pb_knn_tc_impl", {"auto": 0, "tcgen05": 1, "mma": 2}[sweep])


def knn1(x, y, want_d2=False, impl="auto"):
    """x [B,N,3] or [N,3]; y [B,M,3], [M,3] (shared) -> idx int64 [B,N] (and d2 f32).
    impl: "tc" = tcgen05 contraction + exact FP32 re-rank, "simt" = exact FP32 scan, "auto" picks by
    size; both return identical indices."""
    x = _chk(x, torch.float32, "x")
    y = _chk(y, torch.float32, "y")
    squeeze = x.dim() == 2
    if squeeze:
        x = x.unsqueeze(0)
    y_shared = 1 if y.dim() == 2 else 0
    if not y_shared and y.shape[0] != x.shape[0]:
        raise ValueError("batch mismatch between x and y")
    if x.shape[-1] != 3 or y.shape[-1] != 3:
        raise ValueError("points must be [...,3]")
    B, N, _ = x.shape
    M = y.shape[-2]
    idx = torch.empty((B, N), dtype=torch.int64, device=x.device)
    d2 = torch.empty((B, N), dtype=torch.float32, device=x.device) if want_d2 else None
    if impl not in ("auto", "tc", "simt"):
        raise ValueError("impl must be auto, tc or simt")
    use_tc = impl == "tc" or (impl == "auto" and N * M >= TC_MIN_PAIRS and M >= TC_MIN_DB)
    if use_tc:
        from . import _lib
        nb = _lib.load().pb_knn1_tc_workspace_bytes(B, N, M)
        ws = torch.empty(nb + 128, dtype=torch.uint8, device=x.device)
        off = (-ws.data_ptr()) % 128
        import ctypes as C
        call("pb_knn1_tc", _ptr(x), _ptr(y), B, N, M, y_shared, _ptr(idx), _ptr(d2),
             C.c_void_p(ws.data_ptr() + off), nb, _stream(idx))
        global LAST_TC_STATS
        LAST_TC_STATS = ws[off: off + 16].view(torch.int32)       # the workspace starts with int stats[4]
    else:
        call("pb_knn1", _ptr(x), _ptr(y), B, N, M, y_shared, _ptr(idx), _ptr(d2), _stream(idx))
    if squeeze:
        idx = idx[0]
        d2 = d2[0] if want_d2 else None
    return (idx, d2) if want_d2 else idx


def _aligned_ws(nbytes, device):
    """128-byte aligned device workspace -> (owner tensor, ctypes pointer)."""
    import ctypes as C
    ws = torch.empty(nbytes + 128, dtype=torch.uint8, device=device)
    off = (-ws.data_ptr()) % 128
    return ws, off, C.c_void_p(ws.data_ptr() + off)


def _sym_list(flags):
    """u8 [B] device flags -> (int32 device list of the set elements, count); one host sync."""
    lst = torch.nonzero(flags, as_tuple=False).flatten().to(torch.int32)
    return lst, int(lst.numel())


def add_adds(pred, target, want_adds=None, impl="auto"):
    """pred/target [B,N,3] -> add [B], adds [B]; want_adds u8 [B] or None (all).
    impl: "simt" exact FP32 scan, "tc" = N x N search on the tensor cores (same bits out), "auto" uses
    "tc" for large problems when every element wants ADD-S (no flag list to build)."""
    pred = _chk(pred, torch.float32, "pred")
    target = _chk(target, torch.float32, "target")
    want_adds = _chk(want_adds, torch.uint8, "want_adds", True)
    if pred.shape != target.shape:
        raise ValueError("pred and target must have the same shape")
    if impl not in ("auto", "tc", "simt"):
        raise ValueError("impl must be auto, tc or simt")
    B, N, _ = pred.shape
    add = torch.empty((B,), dtype=torch.float32, device=pred.device)
    adds = torch.empty((B,), dtype=torch.float32, device=pred.device)
    if impl == "tc" or (impl == "auto" and want_adds is None and N * N >= TC_MIN_PAIRS and N >= TC_MIN_DB):
        from . import _lib
        global LAST_TC_STATS
        lst, n = (None, 0) if want_adds is None else _sym_list(want_adds)
        nb = _lib.load().pb_add_adds_tc_workspace_bytes(B, N)
        ws, off, wptr = _aligned_ws(nb, pred.device)
        call("pb_add_adds_tc", _ptr(pred), _ptr(target), B, N, _ptr(want_adds), _ptr(lst), n, _ptr(add), _ptr(adds),
             wptr, nb, _stream(add))
        d2b = (B * N * 4 + 127) // 128 * 128
        LAST_TC_STATS = ws[off + d2b: off + d2b + 16].view(torch.int32)
    else:
        call("pb_add_adds", _ptr(pred), _ptr(target), B, N, _ptr(want_adds), _ptr(add), _ptr(adds), _stream(add))
    return add, adds


def transform_points(model, R, t):
    """model [B,N,3] or [N,3] (shared); R [B,3,3]; t [B,3] -> [B,N,3]."""
    model = _chk(model, torch.float32, "model")
    R = _chk(R, torch.float32, "R")
    t = _chk(t, torch.float32, "t")
    B = R.shape[0]
    shared = 1 if model.dim() == 2 else 0
    N = model.shape[-2]
    out = torch.empty((B, N, 3), dtype=torch.float32, device=R.device)
    call("pb_transform_points", _ptr(model), shared, _ptr(R), _ptr(t), B, N, _ptr(out), _stream(out))
    return out


def pose_metrics(R, t, R_gt, t_gt, model, cls, is_sym, target=None, out=None):
    """R,R_gt [F,3,3]; t,t_gt [F,3]; model [C,N,3]; cls i32 [F]; is_sym u8 [C]; target [F,N,3]|None
    -> add [F] (ADD-S for symmetric classes), r_deg [F], t_m [F] (written into ``out`` if given)."""
    R = _chk(R, torch.float32, "R"); t = _chk(t, torch.float32, "t")
    R_gt = _chk(R_gt, torch.float32, "R_gt"); t_gt = _chk(t_gt, torch.float32, "t_gt")
    model = _chk(model, torch.float32, "model")
    cls = _chk(cls, torch.int32, "cls")
    is_sym = _chk(is_sym, torch.uint8, "is_sym", True)
    target = _chk(target, torch.float32, "target", True)
    F = R.shape[0]
    Cn, N, _ = model.shape
    dev = R.device
    if out is not None:
        add, r_deg, t_m = (_chk(o, torch.float32, "out") for o in out)
    else:
        add = torch.empty((F,), dtype=torch.float32, device=dev)
        r_deg = torch.empty((F,), dtype=torch.float32, device=dev)
        t_m = torch.empty((F,), dtype=torch.float32, device=dev)
    call("pb_pose_metrics", _ptr(R), _ptr(t), _ptr(R_gt), _ptr(t_gt), _ptr(model), _ptr(target), _ptr(cls),
         _ptr(is_sym), F, N, Cn, _ptr(add), _ptr(r_deg), _ptr(t_m), _stream(add))
    return add, r_deg, t_m


def new_table(n_cls, device="cuda"):
    return torch.zeros((TABLE_ROWS, n_cls), dtype=torch.float64, device=device)


def accumulate_results(table, add, r_deg, t_m, cls, diameter, status=None, variant=0, add_frac=0.1,
                       r_thr_deg=5.0, t_thr_m=0.05, count_frames=True):
    """table [24,C] f64 += success counts / distance sums of this batch (tools/trainer.py:174-221)."""
    table = _chk(table, torch.float64, "table")
    diameter = _chk(diameter, torch.float32, "diameter")
    status = _chk(status, torch.int32, "status", True)
    F = add.shape[0]
    Cn = table.shape[1]
    call("pb_accumulate_results", _ptr(add), _ptr(r_deg), _ptr(t_m), _ptr(cls), _ptr(status), _ptr(diameter),
         F, Cn, variant, add_frac, r_thr_deg, t_thr_m, 1 if count_frames else 0, _ptr(table), _stream(table))
    return table


def table_to_result(table, objlist):
    """[24,C] table -> the reference's nested ``result`` dict (tools/trainer.py:479-509)."""
    tb = table.detach().cpu().numpy()
    res = {"all_num": {}, "obj_num": {}}
    for v in VARIANTS:
        for col in VARIANT_COLS:
            res[col.format(v)] = {}
    for k in ("dis_mask", "dis_normal", "dis_xyz"):
        res[k] = {}
    for c, obj in enumerate(objlist):
        res["all_num"][obj] = float(tb[ROW_ALL_NUM, c])
        res["obj_num"][obj] = float(tb[ROW_OBJ_NUM, c])
        for vi, v in enumerate(VARIANTS):
            for ci, col in enumerate(VARIANT_COLS):
                res[col.format(v)][obj] = float(tb[ROW_VARIANT0 + 6 * vi + ci, c])
        res["dis_mask"][obj] = float(tb[ROW_DIS_MASK, c])
        res["dis_normal"][obj] = float(tb[ROW_DIS_NORMAL, c])
        res["dis_xyz"][obj] = float(tb[ROW_DIS_XYZ, c])
    return res


def pose_loss_fwd(R, t, model, target, sym, impl="auto", sym_host=None):
    """-> loss_b [B], pred [B,N,3], nn_idx [B,N] (lib/network/loss.py:25-41).
    impl "tc": the symmetric samples' nearest-neighbour search runs on the tensor cores (same bits out);
    "auto" picks it for large N when the caller also passes the flags as a host list (sym_host), so the
    list of symmetric samples needs no device sync."""
    R = _chk(R, torch.float32, "R"); t = _chk(t, torch.float32, "t")
    model = _chk(model, torch.float32, "model"); target = _chk(target, torch.float32, "target")
    sym = _chk(sym, torch.uint8, "sym")
    if impl not in ("auto", "tc", "simt"):
        raise ValueError("impl must be auto, tc or simt")
    B, N, _ = model.shape
    dev = model.device
    pred = torch.empty((B, N, 3), dtype=torch.float32, device=dev)
    nn_idx = torch.empty((B, N), dtype=torch.int64, device=dev)
    loss_b = torch.empty((B,), dtype=torch.float32, device=dev)
    if impl == "tc" or (impl == "auto" and sym_host is not None and any(sym_host) and N * N >= TC_MIN_PAIRS
                       and N >= TC_MIN_DB):
        from . import _lib
        global LAST_TC_STATS
        sym_arg = sym
        if sym_host is None:
            lst, n = _sym_list(sym)
        elif all(sym_host):
            lst, n, sym_arg = None, 0, None
        else:
            ids = [i for i, f in enumerate(sym_host) if f]
            lst, n = torch.tensor(ids, dtype=torch.int32, device=dev), len(ids)
        nb = _lib.load().pb_knn1_tc_workspace_bytes(B, N, N)
        ws, off, wptr = _aligned_ws(nb, dev)
        call("pb_pose_loss_fwd_tc", _ptr(R), _ptr(t), _ptr(model), _ptr(target), _ptr(sym_arg), _ptr(lst), n, B, N,
             _ptr(pred), _ptr(nn_idx), _ptr(loss_b), wptr, nb, _stream(loss_b))
        LAST_TC_STATS = ws[off: off + 16].view(torch.int32)
    else:
        call("pb_pose_loss_fwd", _ptr(R), _ptr(t), _ptr(model), _ptr(target), _ptr(sym), B, N, _ptr(pred),
             _ptr(nn_idx), _ptr(loss_b), _stream(loss_b))
    return loss_b, pred, nn_idx


def pose_loss_bwd(pred, model, target, nn_idx, gscale):
    B, N, _ = pred.shape
    grad_R = torch.empty((B, 3, 3), dtype=torch.float32, device=pred.device)
    grad_t = torch.empty((B, 3), dtype=torch.float32, device=pred.device)
    call("pb_pose_loss_bwd", _ptr(pred), _ptr(model), _ptr(target), _ptr(nn_idx), B, N, float(gscale),
         _ptr(grad_R), _ptr(grad_t), _stream(grad_R))
    return grad_R, grad_t


def vote_translation(cloud, t_res):
    """cloud [B,NP,3], t_res [B,3,NP] -> [B,3] (lib/network/krrn.py:153)."""
    cloud = _chk(cloud, torch.float32, "cloud")
    t_res = _chk(t_res, torch.float32, "t_res")
    B, NP, _ = cloud.shape
    if t_res.shape != (B, 3, NP):
        raise ValueError("t_res must be [B,3,NP]")
    t = torch.empty((B, 3), dtype=torch.float32, device=cloud.device)
    call("pb_vote_translation", _ptr(cloud), _ptr(t_res), B, NP, _ptr(t), _stream(t))
    return t
