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


def format_report(table: torch.Tensor, objlist, opt_pose=False):
    """The per-object lines and the closing all-objects block that the reference's test loop logs
    (tools/trainer.py:281-330), character for character, from the [24,C] table.  Returns (list of per-object
    strings, all-objects string)."""
    res = ops_metric.table_to_result(table, objlist)
    n = lambda i: float(res["obj_num"][i])
    lines = []
    for i in objlist:
        q = lambda k: float(res[k][i]) / n(i)
        blk = lambda v: (f'ADD-S Dis:{q(f"dis_{v}_rt"):.04f} Succ:{q(f"succ_{v}_rt"):.03f}; '
                         f'5° Dis:{q(f"dis_{v}_r"):5.2f} Succ:{q(f"succ_{v}_r"):.03f}; ')
        if opt_pose:
            content = (f'Object {i} Fnl Dis:{q("dis_final_rt"):.04f} Succ:{q("succ_final_rt"):.03f}\n'
                       + 'Base ' + blk("base") + f'5CM Dis:{q("dis_base_t"):.04f} Succ:{q("succ_base_t"):.03f};' + '\n'
                       + 'Regs ' + blk("reg") + f'5CM Dis:{q("dis_reg_t"):.04f} Succ:{q("succ_reg_t"):.03f};' + '\n'
                       + 'Fnal ' + blk("final") + f'5CM Dis:{q("dis_final_t"):.04f} Succ:{q("succ_final_t"):.03f};' + '\n'
                       + f'XYZ Dis: {q("dis_xyz"):.04f}; NML Dis: {q("dis_normal"):.04f}; MASK Dis: {q("dis_mask"):.04f}; ')
        else:
            content = (f'Object {i} ' + 'Base ' + blk("base")
                       + f'5CM Dis:{q("dis_base_t"):.04g} Succ:{q("succ_base_t"):.03f}; '
                       + f'XYZ Dis: {q("dis_xyz"):.04f}; NML Dis: {q("dis_normal"):.04f}; MASK Dis: {q("dis_mask"):.04f}; ')
        lines.append(content)
    tot = float(sum(res["obj_num"].values()))
    g = lambda k: float(sum(res[k].values())) / tot
    ablk = lambda v: (f'ADD-S Dis:{g(f"dis_{v}_rt"):.04f} Succ:{g(f"succ_{v}_rt"):.03f}; '
                      f'5° Dis:{g(f"dis_{v}_r"):5.2f} Succ:{g(f"succ_{v}_r"):.03f}; '
                      f'5CM Dis:{g(f"dis_{v}_t"):.04f} Succ:{g(f"succ_{v}_t"):.03f};')
    all_content = ('\n' + 'Base ' + ablk("base") + '\n' + 'Regs ' + ablk("reg") + '\n' + 'Fnal ' + ablk("final") + '\n'
                   + f'XYZ Dis: {g("dis_xyz"):.04f}; NML Dis: {g("dis_normal"):.04f}; MASK Dis: {g("dis_mask"):.04f}; ')
    return lines, all_content


class EvalSweep:
    """map -> pose -> ADD(-S)/r/t -> accuracy table, all on one stream, no host sync per batch.

    Four ABI calls per batch of F frames: pb_compact_map (K0), pb_pnp_ransac_pts (K1-K4: score +
    finalize kernels), pb_pose_metrics (K6), pb_accumulate_results.  All buffers are persistent."""

    STAGES = ("compact", "ransac", "metrics", "accumulate")

    def __init__(self, model_points, diameter, is_sym, F, H, W, cap, cfg: RansacConfig | None = None,
                 device="cuda", add_frac=0.1, r_thr_deg=5.0, t_thr_m=0.05, lanes=1, free_running=False, deal="split"):
        """deal="alternate" (with lanes > 1; implies free_running): batches are not cut at all -- whole batches are dealt
        to the lanes in turn, so every kernel keeps its full-size, single-wave grid and batch i+1's K0 (HBM) runs under
        batch i's scoring / refit (FP32, latency) on another stream.  The fastest form for a sweep of many batches.
        free_running (with lanes > 1): ``step`` does not join the lanes back into the caller's stream, so lane A's
        kernels of batch i+1 follow its batch i at once while lane B is still on batch i; the join happens in
        ``finish`` / ``table`` / ``reset`` / ``join`` -- the per-frame tensors ``step`` returns are valid only after it.
        lanes > 1: the batch is cut into ``lanes`` contiguous parts that run on their own streams, so the
        HBM-bound K0 of one part overlaps the FP32-bound scoring and the latency-bound refit of another
        (results are identical: frames are independent and the per-lane tables are summed)."""
        dev = torch.device(device)
        self.dev = dev
        self.F, self.H, self.W, self.cap = F, H, W, cap
        self.cfg = cfg or RansacConfig()
        self.model = model_points.to(dev, torch.float32).contiguous()        # [C,N,3]
        self.diameter = diameter.to(dev, torch.float32).contiguous()          # [C]
        self.is_sym = is_sym.to(dev, torch.uint8).contiguous()                # [C]
        self.n_cls = self.model.shape[0]
        self.add_frac, self.r_thr_deg, self.t_thr_m = add_frac, r_thr_deg, t_thr_m
        self.lanes = []
        if deal not in ("split", "alternate"):
            raise ValueError("deal must be 'split' or 'alternate'")
        self.alternate = deal == "alternate" and lanes > 1
        self.free_running = (bool(free_running) or self.alternate) and lanes > 1
        self._unjoined = False
        self._turn = 0
        if lanes > 1:
            if F % lanes and not self.alternate:
                raise ValueError("F must be a multiple of lanes")
            Fl = F if self.alternate else F // lanes
            self.lanes = [EvalSweep(self.model, self.diameter, self.is_sym, Fl, H, W, cap, self.cfg, dev,
                                    add_frac, r_thr_deg, t_thr_m) for _ in range(lanes)]
            # the lanes' tables are rows of ONE buffer: the sweep's table is a single reduction over it (summed in
            # lane order, so the result does not depend on timing), ready for the one all-reduce of finish()
            self._tables = torch.zeros((lanes,) + tuple(self.lanes[0]._table.shape), dtype=torch.float64, device=dev)
            for k, l in enumerate(self.lanes):
                l._table = self._tables[k]
            self.lane_streams = [torch.cuda.Stream(device=dev) for _ in range(lanes)]
            self.ev_in = torch.cuda.Event()
            self.ev_out = [torch.cuda.Event() for _ in range(lanes)]
            self.out = torch.empty((lanes if self.alternate else 1, 3, F), dtype=torch.float32, device=dev)
            self.launches_per_step = self.lanes[0].launches_per_step if self.alternate else \
                sum(l.launches_per_step for l in self.lanes)
            return
        self.compact = ops.MapCompactor(F, H, W, cap, device=dev)
        self.gather = None                       # ops.ChooseGatherer, made on the first step(choose=...)
        self.ransac = ops.PtsRansac(F, cap, self.cfg, device=dev)
        self._table = ops_metric.new_table(self.n_cls, dev)
        self.launches_per_step = self.compact.launches(True) + self.ransac.launches + 2

    def reset(self):
        """Zero the accuracy table(s)."""
        self.join()
        if self.lanes:
            self._tables.zero_()
        else:
            self._table.zero_()

    def next_lane(self):
        """deal='alternate': index of the lane (``lane_streams[k]``) the next ``step`` will run on -- a caller that
        produces the batch on the device can do so on that very stream."""
        return self._turn % len(self.lanes) if self.lanes else 0

    def join(self):
        """Make the caller's stream wait for everything the lanes have been given (no-op without pending work)."""
        if self._unjoined:
            main = torch.cuda.current_stream(self.dev)
            for e in self.ev_out:
                main.wait_event(e)
            self._unjoined = False

    @property
    def table(self):
        if self.lanes:
            self.join()
            return self._tables.sum(dim=0)
        return self._table

    def _hold(self, stream, *tensors):
        """Tell the caching allocator that ``stream`` reads these tensors: a caller that drops its batch right after
        ``step`` returns must not have the memory handed out again while a lane is still reading it."""
        for t in tensors:
            if t is not None and t.is_cuda:
                t.record_stream(stream)

    def step(self, xyz, mask, extent, lfborder, K4, cls, R_gt, t_gt, frame0=0, roi=None, events=None, out=None,
             pred_t=None, cloud=None, t_res=None, choose=None, xmap=None, ymap=None, n_valid=None):
        """One batch of F frames (device tensors).  Returns (add, r_deg, t_m) [F] device tensors.
        choose [F,NP] int64 / int32 (+ xmap / ymap [F,NP] f32, n_valid [F] i32, all optional but ``choose``): the
        correspondences are gathered at the reference's ``choose`` lists (tools/trainer.py:406-421,
        pb_gather_choose) instead of being compacted from ``mask`` (then ignored).
        Lanes (lanes > 1): the batch is consumed on lane streams AFTER this call returns.  Contract: (1) the inputs may
        be dropped at once (they are recorded on the lane's stream); (2) they must not be OVERWRITTEN before
        ``self.last_event`` (recorded when the lane has finished this step) has been waited on, or before ``join`` /
        ``finish``; (3) the returned per-frame tensors alias a ring of ``lanes`` output buffers and are valid after
        ``join`` until ``lanes`` further steps have been issued.
        pred_t [F,3], or cloud [F,NP,3] + t_res [F,3,NP] (the T-offset vote, lib/network/krrn.py:153, is then taken
        here): the ``opt_pose`` branch of the test loop (tools/trainer.py:196-221) -- the 'reg' and 'final' variants
        pair the PnP rotation with the voted translation; their rows of the table are filled and their per-frame
        (add, r_deg, t_m) kept in ``self.last_reg``.
        ``events``: optional list that receives one (name, start, end) CUDA-event triple per stage
        (recorded on the current stream; used by bench.py for the per-kernel roofline)."""
        if self.lanes:
            main = torch.cuda.current_stream(self.dev)
            self.ev_in.record(main)
            if self.alternate:
                k = self._turn % len(self.lanes)
                self._turn += 1
                s = self.lane_streams[k]
                s.wait_event(self.ev_in)
                self._hold(s, xyz, mask, extent, lfborder, K4, cls, R_gt, t_gt, roi, pred_t, cloud, t_res)
                self._hold(s, choose, xmap, ymap, n_valid)
                with torch.cuda.stream(s):
                    self.lanes[k].step(xyz, mask, extent, lfborder, K4, cls, R_gt, t_gt, frame0=frame0, roi=roi,
                                       out=tuple(self.out[k, j] for j in range(3)), pred_t=pred_t, cloud=cloud, t_res=t_res,
                                       choose=choose, xmap=xmap, ymap=ymap, n_valid=n_valid)
                    self.ev_out[k].record(s)
                self.last_event = self.ev_out[k]
                self._unjoined = True
                return self.out[k, 0], self.out[k, 1], self.out[k, 2]
            Fl = self.F // len(self.lanes)
            for k, lane in enumerate(self.lanes):
                s = self.lane_streams[k]
                s.wait_event(self.ev_in)
                self._hold(s, xyz, mask, extent, lfborder, K4, cls, R_gt, t_gt, roi, pred_t, cloud, t_res)
                self._hold(s, choose, xmap, ymap, n_valid)
                lo, hi = k * Fl, (k + 1) * Fl
                cut = lambda t: None if t is None else t[lo:hi]
                with torch.cuda.stream(s):
                    lane.step(xyz[lo:hi], cut(mask), extent[lo:hi], lfborder[lo:hi],
                              K4[lo:hi], cls[lo:hi], R_gt[lo:hi], t_gt[lo:hi], frame0=frame0 + lo,
                              roi=cut(roi), out=tuple(self.out[0, j, lo:hi] for j in range(3)),
                              pred_t=cut(pred_t), cloud=cut(cloud), t_res=cut(t_res),
                              choose=cut(choose), xmap=cut(xmap), ymap=cut(ymap), n_valid=cut(n_valid))
                    self.ev_out[k].record(s)
            self.last_event = self.ev_out[-1]
            self._unjoined = True
            if not self.free_running:
                self.join()
            return self.out[0, 0], self.out[0, 1], self.out[0, 2]

        def mark():
            if events is None:
                return None
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            return e
        e0 = mark()
        if choose is not None:
            if self.gather is None:
                self.gather = ops.ChooseGatherer(self.F, self.H, self.W, self.cap, device=self.dev)
            pts, count = self.gather(xyz, choose, extent, lfborder, xmap=xmap, ymap=ymap, n_valid=n_valid, roi=roi)
        else:
            pts, count = self.compact(xyz, mask, extent, lfborder, roi=roi)
        e1 = mark()
        rs = self.ransac(pts, count, K4, frame0=frame0)
        e2 = mark()
        add, r_deg, t_m = ops_metric.pose_metrics(rs.R, rs.t, R_gt, t_gt, self.model, cls, self.is_sym, out=out)
        e3 = mark()
        ops_metric.accumulate_results(self._table, add, r_deg, t_m, cls, self.diameter, status=rs.status,
                                      variant=0, add_frac=self.add_frac, r_thr_deg=self.r_thr_deg,
                                      t_thr_m=self.t_thr_m, count_frames=True)
        e4 = mark()
        if events is not None:
            events.append((e0, e1, e2, e3, e4))
        if cloud is not None:
            pred_t = ops_metric.vote_translation(cloud, t_res)                       # krrn.py:153
        if pred_t is not None:                                                       # trainer.py:196-199: reg == final
            reg = ops_metric.pose_metrics(rs.R, pred_t.reshape(-1, 3), R_gt, t_gt, self.model, cls, self.is_sym)
            for variant in (1, 2):
                ops_metric.accumulate_results(self._table, *reg, cls, self.diameter, status=rs.status, variant=variant,
                                              add_frac=self.add_frac, r_thr_deg=self.r_thr_deg, t_thr_m=self.t_thr_m,
                                              count_frames=False)
            self.last_reg = reg
        return add, r_deg, t_m

    def finish(self, group=None):
        """All-reduce the table across ranks (on the compute stream) and return it."""
        return reduce_table(self.table, group)


SMALL_WIDTHS = (3, 3, 4, 9, 3)          # extent | lfborder | K4 | R_gt | t_gt
SMALL_COLS = sum(SMALL_WIDTHS)


def pack_small(extent, lfborder, K4, R_gt, t_gt, chunk):
    """Per-frame small inputs -> [F/chunk, 22*chunk] f32: one H2D copy per chunk, and every field is
    a contiguous view after it (extent [chunk,3] | lfborder [chunk,3] | K4 | R_gt | t_gt)."""
    F = extent.shape[0]
    parts = [torch.as_tensor(a, dtype=torch.float32).reshape(F // chunk, chunk * w)
             for a, w in zip((extent, lfborder, K4, R_gt, t_gt), SMALL_WIDTHS)]
    return torch.cat(parts, dim=1).contiguous()


def unpack_small(flat, chunk):
    out, off = [], 0
    for w in SMALL_WIDTHS:
        out.append(flat[off:off + chunk * w].view(chunk, w))
        off += chunk * w
    out[3] = out[3].view(chunk, 3, 3)
    return out


class HostEvalSweep:
    """The same pass for inputs that live in (pinned) HOST memory -- what a caller holding the
    network output on the host would run.  The batch is cut into chunks of ``chunk`` frames; the
    H2D copy of chunk i+1 (copy stream) overlaps K0..K6 of chunk i (compute stream) through two
    device staging buffers; the only D2H traffic is the [24,C] accuracy table at the end."""

    def __init__(self, model_points, diameter, is_sym, H, W, cap, chunk=128, cfg: RansacConfig | None = None,
                 device="cuda", zero_copy=False, **kw):
        """zero_copy=True (needs a mask): no staging copy of the maps at all -- the mask-first K0
        kernel reads the pinned host maps through their mapped (UVA) addresses, so only the mask
        bytes and the foreground pixels' 32-byte sectors cross PCIe (csrc/compact.cu).
        zero_copy="xyz": the mask chunk is staged by the copy engine (full-rate DMA) and only the
        coordinate planes are read in place."""
        dev = torch.device(device)
        self.dev, self.chunk, self.H, self.W = dev, chunk, H, W
        self.zero_copy = zero_copy
        # zero-copy chunks alternate between two compute lanes (own stream, buffers and table): K1..K6 of chunk i
        # overlap the PCIe-bound K0 of chunk i+1
        self.stage_mask = (not zero_copy) or zero_copy == "xyz"
        nl = 2 if zero_copy else 1
        self.sweeps = [EvalSweep(model_points, diameter, is_sym, chunk, H, W, cap, cfg, device=dev, **kw) for _ in range(nl)]
        self.sweep = self.sweeps[0]
        # the lanes' tables are rows of one buffer (one reduction gives the table, one memset resets it)
        self._tables = torch.zeros((nl,) + tuple(self.sweep._table.shape), dtype=torch.float64, device=dev)
        for k, sw_ in enumerate(self.sweeps):
            sw_._table = self._tables[k]
        self.lane_streams = [torch.cuda.Stream(device=dev) for _ in range(nl)] if nl > 1 else None
        self.lane_done = [torch.cuda.Event() for _ in range(nl)]
        self.start = torch.cuda.Event()
        self.copy_stream = torch.cuda.Stream(device=dev)
        zc = zero_copy
        self.stage = [dict(xyz=None if zc else torch.empty((chunk, 3, H, W), dtype=torch.float32, device=dev),
                           mask=torch.empty((chunk, H, W), dtype=torch.uint8, device=dev) if self.stage_mask else None,
                           small=torch.empty((SMALL_COLS * chunk,), dtype=torch.float32, device=dev),
                           cls=torch.empty((chunk,), dtype=torch.int32, device=dev),
                           roi=torch.empty((chunk, 2), dtype=torch.int32, device=dev),
                           ready=torch.cuda.Event(), free=torch.cuda.Event()) for _ in range(2)]
        self.table_host = torch.empty((ops_metric.TABLE_ROWS, self.sweep.n_cls), dtype=torch.float64).pin_memory()
        self.h2d_bytes = 0
        self.cap = cap
        self.list_stage = None                   # staging buffers of run_choose, made on first use

    @property
    def table(self):
        if len(self.sweeps) == 1:
            return self.sweep.table
        return self._tables.sum(dim=0)

    def reset(self):
        self._tables.zero_()

    def run(self, xyz_h, mask_h, small_h, cls_h, frame0=0, group=None, roi_h=None):
        """xyz_h [F,3,H,W] f32, mask_h [F,H,W] u8, small_h = pack_small(...) [F/chunk, 22*chunk] f32,
        cls_h [F] i32 -- pinned host tensors; F a multiple of ``chunk``.  roi_h [F,2] i32 (rmin, cmin), optional:
        the maps are CROPS of the image, as the reference's network emits them (``pred['xyz']`` is the
        ``get_square_bbox`` crop, dataset/linemod/batchdataset.py:890-961), and row / column + roi gives the image
        pixel.  Returns the table on the host (this call synchronises once, at the end)."""
        F = xyz_h.shape[0]
        if F % self.chunk:
            raise ValueError("F must be a multiple of the chunk size")
        comp = torch.cuda.current_stream(self.dev)
        n = F // self.chunk
        self.h2d_bytes = 0
        nl = len(self.sweeps)
        if nl > 1:
            self.start.record(comp)
            for s_ in self.lane_streams:
                s_.wait_event(self.start)               # table resets etc. queued on the caller's stream come first
        for i in range(n):
            st = self.stage[i & 1]
            sweep = self.sweeps[i % nl]
            lane = self.lane_streams[i % nl] if nl > 1 else comp
            lo, hi = i * self.chunk, (i + 1) * self.chunk
            with torch.cuda.stream(self.copy_stream):
                if i >= 2:
                    self.copy_stream.wait_event(st["free"])
                if not self.zero_copy:
                    st["xyz"].copy_(xyz_h[lo:hi], non_blocking=True)
                if self.stage_mask:
                    st["mask"].copy_(mask_h[lo:hi], non_blocking=True)
                st["small"].copy_(small_h[i], non_blocking=True)
                st["cls"].copy_(cls_h[lo:hi], non_blocking=True)
                if roi_h is not None:
                    st["roi"].copy_(roi_h[lo:hi], non_blocking=True)
                st["ready"].record(self.copy_stream)
            self.h2d_bytes += sum(st[k].numel() * st[k].element_size() for k in ("xyz", "mask", "small", "cls")
                                  if st[k] is not None) + (0 if roi_h is None else st["roi"].numel() * 4)
            roi = None if roi_h is None else st["roi"]
            lane.wait_event(st["ready"])
            ext, lfb, K4, Rg, tg = unpack_small(st["small"], self.chunk)
            with torch.cuda.stream(lane):
                if self.zero_copy:
                    if mask_h is None:
                        raise ValueError("zero_copy needs a mask (without one every pixel must be streamed)")
                    if not self.stage_mask:
                        self.h2d_bytes += mask_h[lo:hi].numel()  # + the gathered sectors, added by the caller
                    sweep.step(xyz_h[lo:hi], st["mask"] if self.stage_mask else mask_h[lo:hi], ext, lfb, K4, st["cls"],
                               Rg, tg, frame0=frame0 + lo, roi=roi)
                else:
                    sweep.step(st["xyz"], st["mask"], ext, lfb, K4, st["cls"], Rg, tg, frame0=frame0 + lo, roi=roi)
                st["free"].record(lane)
        if nl > 1:
            for k, s_ in enumerate(self.lane_streams):
                self.lane_done[k].record(s_)
                comp.wait_event(self.lane_done[k])
        table = self.table
        reduce_table(table, group)
        self.table_host.copy_(table, non_blocking=True)
        comp.synchronize()
        return self.table_host


    # ------------------------------------------------------------------------------------------------
    def _list_stage(self, NP, idx_dtype, with_xy):
        key = (NP, idx_dtype, with_xy)
        if self.list_stage is None or self.list_stage[0] != key:
            dev, c = self.dev, self.chunk
            mk = lambda: dict(choose=torch.empty((c, NP), dtype=idx_dtype, device=dev),
                              xmap=torch.empty((c, NP), dtype=torch.float32, device=dev) if with_xy else None,
                              ymap=torch.empty((c, NP), dtype=torch.float32, device=dev) if with_xy else None,
                              n_valid=torch.empty((c,), dtype=torch.int32, device=dev))
            self.list_stage = (key, [mk(), mk()])
        return self.list_stage[1]

    def run_choose(self, xyz, choose_h, small_h, cls_h, xmap_h=None, ymap_h=None, n_valid_h=None, frame0=0, group=None,
                   roi_h=None):
        """The pass for a caller that holds what the reference's test loop holds (tools/trainer.py:383-421): the
        coordinate maps and, per frame, the ``choose`` index list with its pixel-coordinate lists.
        xyz [F,3,H,W] f32: PINNED HOST memory (read in place by the gather kernel: only the chosen pixels' sectors
        cross PCIe) or a CUDA tensor (the network's output still on the device: only the lists cross);
        choose_h [F,NP] int64 (the reference's dtype) or int32, xmap_h / ymap_h [F,NP] f32 or None (coordinates
        from the index and roi_h), n_valid_h [F] i32 or None (NP entries each), small_h = pack_small(...),
        cls_h [F] i32 -- pinned host tensors.  Per chunk the lists go by the copy engine while the previous chunk is
        gathered and evaluated; chunks alternate between two compute lanes.  Returns the table on the host."""
        F = choose_h.shape[0]
        if F % self.chunk:
            raise ValueError("F must be a multiple of the chunk size")
        if len(self.sweeps) < 2:
            raise ValueError("run_choose needs the two-lane form: construct HostEvalSweep(..., zero_copy=True)")
        if not (xyz.is_cuda or xyz.is_pinned()):
            raise ValueError("xyz must be a CUDA tensor or a pinned host tensor")
        NP = choose_h.shape[1]
        with_xy = xmap_h is not None
        self._enqueue_choose(xyz, choose_h, small_h, cls_h, xmap_h, ymap_h, n_valid_h, frame0, roi_h)
        comp = torch.cuda.current_stream(self.dev)
        table = self.table
        reduce_table(table, group)
        self.table_host.copy_(table, non_blocking=True)
        comp.synchronize()
        return self.table_host

    def _enqueue_choose(self, xyz, choose_h, small_h, cls_h, xmap_h, ymap_h, n_valid_h, frame0, roi_h):
        """Everything run_choose puts on the streams, up to the lanes joining the caller's stream (no host sync)."""
        F = choose_h.shape[0]
        NP = choose_h.shape[1]
        with_xy = xmap_h is not None
        stg = self._list_stage(NP, choose_h.dtype, with_xy)
        comp = torch.cuda.current_stream(self.dev)
        n, nl = F // self.chunk, len(self.sweeps)
        self.h2d_bytes = 0
        self.start.record(comp)
        for s_ in self.lane_streams:
            s_.wait_event(self.start)
        self.copy_stream.wait_event(self.start)      # (under stream capture every stream must fork from the caller's)
        for i in range(n):
            st, ls = self.stage[i & 1], stg[i & 1]
            sweep, lane = self.sweeps[i % nl], self.lane_streams[i % nl]
            lo, hi = i * self.chunk, (i + 1) * self.chunk
            with torch.cuda.stream(self.copy_stream):
                if i >= 2:
                    self.copy_stream.wait_event(st["free"])
                ls["choose"].copy_(choose_h[lo:hi], non_blocking=True)
                moved = [ls["choose"], st["small"], st["cls"]]
                if with_xy:
                    ls["xmap"].copy_(xmap_h[lo:hi], non_blocking=True)
                    ls["ymap"].copy_(ymap_h[lo:hi], non_blocking=True)
                    moved += [ls["xmap"], ls["ymap"]]
                if n_valid_h is not None:
                    ls["n_valid"].copy_(n_valid_h[lo:hi], non_blocking=True)
                    moved.append(ls["n_valid"])
                st["small"].copy_(small_h[i], non_blocking=True)
                st["cls"].copy_(cls_h[lo:hi], non_blocking=True)
                if roi_h is not None:
                    st["roi"].copy_(roi_h[lo:hi], non_blocking=True)
                    moved.append(st["roi"])
                st["ready"].record(self.copy_stream)
            self.h2d_bytes += sum(t.numel() * t.element_size() for t in moved)
            lane.wait_event(st["ready"])
            ext, lfb, K4, Rg, tg = unpack_small(st["small"], self.chunk)
            with torch.cuda.stream(lane):
                sweep.step(xyz[lo:hi], None, ext, lfb, K4, st["cls"], Rg, tg, frame0=frame0 + lo,
                           roi=None if roi_h is None else st["roi"], choose=ls["choose"], xmap=ls["xmap"], ymap=ls["ymap"],
                           n_valid=None if n_valid_h is None else ls["n_valid"])
                st["free"].record(lane)
        for k, s_ in enumerate(self.lane_streams):
            self.lane_done[k].record(s_)
            comp.wait_event(self.lane_done[k])

    def capture_choose(self, xyz, choose_h, small_h, cls_h, xmap_h=None, ymap_h=None, n_valid_h=None, frame0=0, roi_h=None):
        """run_choose as ONE replayed CUDA graph, for a serving loop that refills the same pinned host buffers (and
        the same xyz tensor) every pass: the per-pass host work shrinks from ~100 launches / copies to one
        cudaGraphLaunch, the table all-reduce, one D2H copy and one sync.  Runs the pass once eagerly (warm-up), then
        captures it -- table reset, list copies on the copy stream, gather + K1..K6 on the two lanes, the table sum.
        Returns ``replay(group=None) -> table on the host``; the graph reads whatever the buffers hold at replay time
        (same shapes, same addresses)."""
        args = (xyz, choose_h, small_h, cls_h, xmap_h, ymap_h, n_valid_h, frame0, roi_h)
        for t in (choose_h, small_h, cls_h, xmap_h, ymap_h, n_valid_h, roi_h):
            if t is not None and not t.is_pinned():
                raise ValueError("capture_choose needs pinned host tensors (their addresses are baked into the graph)")
        self.reset()
        self.run_choose(xyz, choose_h, small_h, cls_h, xmap_h=xmap_h, ymap_h=ymap_h, n_valid_h=n_valid_h, frame0=frame0,
                        roi_h=roi_h)
        torch.cuda.synchronize(self.dev)
        tab_out = torch.zeros_like(self._tables[0])
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            self._tables.zero_()
            self._enqueue_choose(*args)
            tab_out.copy_(self._tables.sum(dim=0))
        self._graph_keep = (graph, tab_out, args)          # the graph reads these tensors: keep them alive

        done = torch.cuda.Event()

        def submit(group=None, stream=None):
            """Enqueue the pass (graph, table all-reduce, result copy) on ``stream`` (default: the current one); no sync."""
            s_ = stream if stream is not None else torch.cuda.current_stream(self.dev)
            with torch.cuda.stream(s_):
                graph.replay()
                reduce_table(tab_out, group)
                self.table_host.copy_(tab_out, non_blocking=True)
                done.record(s_)

        def wait():
            done.synchronize()
            return self.table_host

        def replay(group=None):
            submit(group)
            return wait()
        replay.submit, replay.wait = submit, wait
        return replay


class PipelinedPasses:
    """Captured passes in flight together -- the serving loop's double buffering.  ``replays`` are the callables
    HostEvalSweep.capture_choose returned, EACH FROM ITS OWN HostEvalSweep (own staging buffers, lanes and tables; the
    caller refills that instance's pinned host buffers before submitting it again).  They are replayed in turn on their own
    streams, so the list copies and the zero-copy gather of pass k+1 cross PCIe under the scoring tail, table reduction and
    result copy of pass k.  ``submit()`` enqueues the next pass and returns the table (a host copy) of the oldest pass
    still in flight once ``len(replays)`` are, else None; ``drain()`` returns the tables still outstanding, oldest first."""

    def __init__(self, replays, device="cuda"):
        if len(replays) < 2:
            raise ValueError("PipelinedPasses needs at least two captured passes")
        self.replays = list(replays)
        self.streams = [torch.cuda.Stream(device=torch.device(device)) for _ in self.replays]
        self.next = 0
        self.pending = []                       # indices of the passes in flight, oldest first

    def submit(self, group=None):
        k = self.next
        self.next = (k + 1) % len(self.replays)
        out = None
        if k in self.pending:                   # (only when the caller skipped the returned tables: depth exceeded)
            raise RuntimeError("pass still in flight")
        self.replays[k].submit(group, self.streams[k])
        self.pending.append(k)
        if len(self.pending) == len(self.replays):
            out = self.replays[self.pending.pop(0)].wait().clone()
        return out

    def drain(self):
        outs = [self.replays[k].wait().clone() for k in self.pending]
        self.pending = []
        return outs

