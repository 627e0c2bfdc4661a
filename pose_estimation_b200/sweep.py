"""Batched, sync-free evaluation sweep (the reference's test loop, tools/trainer.py:145-368, without
its per-frame D2H / H2D hops and .item() syncs) and its multi-GPU form.

Frames are independent, so ranks take contiguous frame shards with no data-path collective; the
only exchange is ONE all-reduce(SUM) of the [24, n_cls] float64 accuracy table (the reference's
``result`` dict, trainer.py:479-509) at the end of the sweep -- new functionality, the reference
never reduces evaluation metrics across ranks.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import ops, ops_metric
from .ops import RansacConfig


def shard_range(n_frames: int, rank: int, world: int):
    """Contiguous shard [lo, hi) of rank ``rank``; sizes differ by at most one frame."""
    base, rem = divmod(n_frames, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def reduce_table(table: torch.Tensor, group=None) -> torch.Tensor:
    """In-place all-reduce(SUM) of the accuracy table over the process group (no-op when
    torch.distributed is not initialised or the world has one rank)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(table, op=dist.ReduceOp.SUM, group=group)
    return table


def summarize(table: torch.Tensor, objlist):
    """Per-object and overall success rates / mean distances, the numbers the reference logs at
    tools/trainer.py:281-330."""
    tb = table.detach().cpu()
    n = tb[ops_metric.ROW_OBJ_NUM].clamp(min=1.0)
    out = {"objects": {}, "frames": float(tb[ops_metric.ROW_ALL_NUM].sum())}
    tot = max(out["frames"], 1.0)
    for vi, v in enumerate(ops_metric.VARIANTS):
        r0 = ops_metric.ROW_VARIANT0 + 6 * vi
        out[f"succ_{v}_rt"] = float(tb[r0].sum() / tot)
        out[f"dis_{v}_rt"] = float(tb[r0 + 1].sum() / tot)
        out[f"succ_{v}_r"] = float(tb[r0 + 2].sum() / tot)
        out[f"succ_{v}_t"] = float(tb[r0 + 4].sum() / tot)
    for c, obj in enumerate(objlist):
        r0 = ops_metric.ROW_VARIANT0
        out["objects"][obj] = {"num": float(tb[ops_metric.ROW_OBJ_NUM, c]),
                               "succ_base_rt": float(tb[r0, c] / n[c]), "dis_base_rt": float(tb[r0 + 1, c] / n[c]),
                               "succ_base_r": float(tb[r0 + 2, c] / n[c]), "succ_base_t": float(tb[r0 + 4, c] / n[c])}
    out["failed"] = float(tb[ops_metric.ROW_FAILED].sum())
    return out


class EvalSweep:
    """map -> pose -> ADD(-S)/r/t -> accuracy table, all on one stream, no host sync per batch."""

    def __init__(self, model_points, diameter, is_sym, F, H, W, cap, cfg: RansacConfig | None = None,
                 device="cuda", add_frac=0.1, r_thr_deg=5.0, t_thr_m=0.05):
        dev = torch.device(device)
        self.model = model_points.to(dev, torch.float32).contiguous()        # [C,N,3]
        self.diameter = diameter.to(dev, torch.float32).contiguous()          # [C]
        self.is_sym = is_sym.to(dev, torch.uint8).contiguous()                # [C]
        self.n_cls = self.model.shape[0]
        self.ransac = ops.MapRansac(F, H, W, cap, cfg, device=dev)
        self.table = ops_metric.new_table(self.n_cls, dev)
        self.add_frac, self.r_thr_deg, self.t_thr_m = add_frac, r_thr_deg, t_thr_m
        self.launches_per_step = None

    def step(self, xyz, mask, extent, lfborder, K4, cls, R_gt, t_gt, frame0=0, roi=None):
        """One batch of F frames (device tensors).  Returns (add, r_deg, t_m) [F] device tensors."""
        rs = self.ransac(xyz, mask, extent, lfborder, K4, roi=roi, frame0=frame0)
        add, r_deg, t_m = ops_metric.pose_metrics(rs.R, rs.t, R_gt, t_gt, self.model, cls, self.is_sym)
        ops_metric.accumulate_results(self.table, add, r_deg, t_m, cls, self.diameter, status=rs.status,
                                      variant=0, add_frac=self.add_frac, r_thr_deg=self.r_thr_deg,
                                      t_thr_m=self.t_thr_m, count_frames=True)
        return add, r_deg, t_m

    def finish(self, group=None):
        """All-reduce the table across ranks (on the compute stream) and return it."""
        return reduce_table(self.table, group)
