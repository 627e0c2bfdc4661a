#!/usr/bin/env python
"""bench_configs.py -- the other BASELINE.json configs (C1 ape batch, C3 KNN / ADD-S / symmetric-KNN loss, C4 dense
correspondences + T-vote, C5 65,536-frame sweep), measured with CUDA events on one B200.  bench.py stays the contract
benchmark (C2); this script only fills BASELINE.md's table.  One JSON line per config."""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from pose_estimation_b200 import ops, ops_metric, synth  # noqa: E402


def timed(fn, n, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def c3(args, dev):
    """eggbox / glue: B pose pairs, N = M = 2620 model points: KNN indices, ADD-S, PoseLoss value."""
    B, N = args.c3_batch, 2620
    rng = np.random.default_rng(3)
    models = synth.class_models(N)
    cls = np.where(np.arange(B) % 2 == 0, 7, 8)
    model = torch.from_numpy(models[cls]).to(dev)                        # [B,N,3]
    q = rng.normal(size=(B, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
    Rg = np.stack([synth.random_rotation(np.random.default_rng(i)) for i in range(B)]).astype(np.float32)
    tg = (rng.normal(size=(B, 3)) * 0.05 + [0, 0, 0.9]).astype(np.float32)
    ang = np.radians(rng.uniform(0, 10, B)); ax = rng.normal(size=(B, 3)); ax /= np.linalg.norm(ax, axis=1, keepdims=True)
    Kx = np.zeros((B, 3, 3)); Kx[:, 0, 1] = -ax[:, 2]; Kx[:, 0, 2] = ax[:, 1]; Kx[:, 1, 0] = ax[:, 2]
    Kx[:, 1, 2] = -ax[:, 0]; Kx[:, 2, 0] = -ax[:, 1]; Kx[:, 2, 1] = ax[:, 0]
    dR = np.eye(3)[None] + np.sin(ang)[:, None, None] * Kx + (1 - np.cos(ang))[:, None, None] * (Kx @ Kx)
    R = (dR @ Rg).astype(np.float32)
    t = (tg + rng.normal(size=(B, 3)) * 0.005).astype(np.float32)
    Rt, tt, Rgt, tgt = (torch.from_numpy(a).to(dev) for a in (R, t, Rg, tg))
    pred = ops_metric.transform_points(model, Rt, tt)
    target = ops_metric.transform_points(model, Rgt, tgt)
    sym = torch.ones(B, dtype=torch.uint8, device=dev)
    out = {}
    ms = timed(lambda: ops_metric.knn1(pred, target, impl="simt"), args.iters)
    out["knn1_simt"] = {"ms": ms, "pairs_per_s": B * N * N / (ms * 1e-3)}
    ms = timed(lambda: ops_metric.knn1(pred, target, impl="tc"), args.iters)
    st = ops_metric.LAST_TC_STATS.cpu().tolist()
    out["knn1_tc"] = {"ms": ms, "pairs_per_s": B * N * N / (ms * 1e-3), "fallback_rows": st[1],
                      "exact_evals_per_row": st[2] / (B * N), "timeout": st[0],
                      "chunks_looked_at_frac": st[3] / (B * ((N + 127) // 128) * 8 * ((N + 127) // 128) * 2)}
    same = bool(torch.equal(ops_metric.knn1(pred, target, impl="tc"), ops_metric.knn1(pred, target, impl="simt")))
    out["tc_equals_simt"] = same
    for impl in ("simt", "tc"):
        ms = timed(lambda: ops_metric.add_adds(pred, target, None, impl=impl), args.iters)
        out["add_adds_" + impl] = {"ms": ms, "pairs_per_s": B * N * N / (ms * 1e-3)}
        ms = timed(lambda: ops_metric.pose_loss_fwd(Rt, tt, model, target, sym, impl=impl, sym_host=[1] * B), args.iters)
        out["pose_loss_fwd_" + impl] = {"ms": ms, "pairs_per_s": B * N * N / (ms * 1e-3)}
    a_s = ops_metric.add_adds(pred, target, None, impl="simt"); a_t = ops_metric.add_adds(pred, target, None, impl="tc")
    l_s = ops_metric.pose_loss_fwd(Rt, tt, model, target, sym, impl="simt")
    l_t = ops_metric.pose_loss_fwd(Rt, tt, model, target, sym, impl="tc", sym_host=[1] * B)
    out["adds_tc_equals_simt"] = bool(torch.equal(a_s[1], a_t[1]) and torch.equal(a_s[0], a_t[0]))
    out["loss_tc_equals_simt"] = bool(torch.equal(l_s[0], l_t[0]) and torch.equal(l_s[2], l_t[2]))
    # roofline of the tensor-core search as DESIGN.md 4.1 states it: the minimum over the N*M scores on the ALU pipe,
    # 0.66 FMNMX per score at 16 lanes / clk / SMSP (the measured bound of the epilogue; the tensor pipe is at ~12 %)
    import ctypes as C
    from pose_estimation_b200 import _lib
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    clk = torch.cuda.get_device_properties(dev).clock_rate * 1e3 if hasattr(torch.cuda.get_device_properties(dev), "clock_rate") else 1.965e9
    alu_bound = sms * 4 * 16 * clk / 0.66
    out["roofline"] = {"kernel": "knn_tc_kernel", "bound": "alu (min epilogue)", "achieved": out["knn1_tc"]["pairs_per_s"],
                       "peak": alu_bound, "unit": "pairs/s", "frac": out["knn1_tc"]["pairs_per_s"] / alu_bound,
                       "note": "peak = SMs x 4 SMSP x 16 lanes x clock / 0.66 FMNMX per score"}
    if getattr(args, "cpu", False):
        # the reference's own N x N form (lib/utils/metric.py:24-32) on the host cores, a bounded sample of problems
        import time
        from oracle import ref_callsites as rc
        ph, th = pred[:4].cpu(), target[:4].cpu()
        rc.cal_adds_torch_nxn(ph[0], th[0])
        t0 = time.perf_counter()
        vals = [rc.cal_adds_torch_nxn(ph[k], th[k]) for k in range(4)]
        dt = (time.perf_counter() - t0) / 4
        got = ops_metric.add_adds(pred[:4], target[:4], None, impl="tc")[1].cpu().numpy()
        out["cpu_baseline"] = {"value": N * N / dt, "unit": "pairs/s", "s_per_problem": dt, "cores": torch.get_num_threads(),
                               "kind": "port", "sample": "4 problems of 2620 x 2620; oracle/ref_callsites.cal_adds_torch_nxn = the "
                               "[N,N,3] repeat / norm / min / mean of lib/utils/metric.py:24-32 in torch on the host",
                               "adds_equal_to_1e-5": bool(np.allclose(got, vals, rtol=1e-5))}
    res = {"config": "C3", "batch": B, "N": N, "pairs": B * N * N, **out}
    print(json.dumps(res), flush=True)
    return res


def c4(args, dev):
    """F frames, P = 10k correspondences, H = 1024 hypotheses, 5 re-score/refit rounds, T-vote on 10k points."""
    F, cap, H = args.c4_frames, 10240, 1024
    sps = synth.make_batch_sparse(0x5EED, 4, min(F, 16), p_range=(10000, 10000), outlier_frac=0.3,
                                  z_range=(0.30, 0.40), xy_frac=0.25)       # close-ups: silhouettes >= 10k px
    pts = np.zeros((F, 5, cap), np.float32); cnt = np.zeros(F, np.int32)
    for f in range(F):
        p = synth.compact_points(sps[f % len(sps)])
        n = min(p.shape[1], cap); pts[f, :, :n] = p[:, :n]; cnt[f] = n
    K4 = torch.from_numpy(np.tile(synth.LINEMOD_K4, (F, 1))).to(dev)
    pts_d, cnt_d = torch.from_numpy(pts).to(dev), torch.from_numpy(cnt).to(dev)
    rs = ops.PtsRansac(F, cap, ops.RansacConfig(n_hyp=H, lo_rounds=5), device=dev)
    ms = timed(lambda: rs(pts_d, cnt_d, K4), args.iters)
    hp = float(H) * float(cnt.sum())
    NP = 10000
    cloud = torch.randn(F, NP, 3, device=dev); tres = torch.randn(F, 3, NP, device=dev)
    ms_v = timed(lambda: ops_metric.vote_translation(cloud, tres), args.iters)
    res = {"config": "C4", "frames": F, "P_mean": float(cnt.mean()), "n_hyp": H, "lo_rounds": 5,
           "ransac_ms": ms, "frames_per_s": F / (ms * 1e-3), "hyp_pts_per_s": hp / (ms * 1e-3),
           "vote_ms": ms_v, "vote_GBps": F * NP * 24 / (ms_v * 1e-3) / 1e9}
    if getattr(args, "cpu", False):
        # the reference on the same kind of frame: get_pose still draws 256 of the 10 k pixels (tools/trainer.py:406-407)
        import bench
        spec = dict(config_id=4, kw=dict(p_range=(10000, 10000), outlier_frac=0.3, z_range=(0.30, 0.40), xy_frac=0.25))
        n_cpu = 4 * (os.cpu_count() or 1)
        times, workers, n_cpu, succ = bench.run_reference_cpu(n_cpu, 3, spec=spec)
        res["cpu_baseline"] = {"value": n_cpu / min(times[1:]), "unit": "frames/s", "cores": workers, "kind": "port",
                               "sample": f"{n_cpu} frames with 10 k masked px, best of 2 passes; oracle/ref_callsites.eval_frame "
                                         "(cv2.solvePnPRansac on 256 random px + ADD), one process per core",
                               "succ_add": succ[0] / n_cpu}
    print(json.dumps(res), flush=True)
    return res


def c1(args, dev):
    """LineMOD 'ape' batch: F = 1 and 64 frames of object 1, P ~ 1k masked px, 500-point model, H = 256."""
    import pose_estimation_b200 as pb
    out = {"config": "C1", "n_hyp": 256, "model_points": 500}
    H, W = synth.IMG_H, synth.IMG_W
    models = torch.from_numpy(synth.class_models(500))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    for F in (1, 64):
        sps = [synth.make_frame_sparse(0x5EED, 1, f, cls=0) for f in range(F)]
        xyz = np.zeros((F, 3, H, W), np.float32); mask = np.zeros((F, H, W), np.uint8)
        for i, sp in enumerate(sps):
            xyz[i], mask[i] = synth.dense_from_sparse(sp)
        cls = np.zeros(F, np.int32)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        inp = dict(xyz=t(xyz), mask=t(mask), cls=t(cls), extent=t(synth.EXTENT[cls].astype(np.float32)),
                   lfborder=t(synth.LFBORDER[cls].astype(np.float32)), K4=t(np.tile(synth.LINEMOD_K4, (F, 1))),
                   R_gt=t(np.stack([s["R_gt"] for s in sps])), t_gt=t(np.stack([s["t_gt"] for s in sps])))
        sw = pb.EvalSweep(models, diam, sym, F, H, W, 4096, ops.RansacConfig(n_hyp=256, seed=0x5EED), device=dev)
        def step():
            sw.reset()
            sw.step(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], inp["K4"], inp["cls"], inp["R_gt"], inp["t_gt"])
        ms = timed(step, max(args.iters, 20), warm=5)
        tb = sw.table.cpu().numpy()
        out[f"F{F}"] = {"ms_per_batch": ms, "frames_per_s": F / (ms * 1e-3), "P_mean": float(np.mean([len(s["rows"]) for s in sps])),
                        "add_success": float(tb[ops_metric.ROW_VARIANT0, 0] / max(tb[ops_metric.ROW_OBJ_NUM, 0], 1))}
    # the reference's own loop shape: one frame per call through the drop-in call sites (tools/trainer.py:166-172:
    # get_pose -> cal_dis, host tensors in, Python floats out, so every call synchronises) and the same step as a
    # replayed CUDA graph (one launch per frame instead of six)
    import time
    sp = synth.make_frame_sparse(0x5EED, 1, 0, cls=0)
    fr = synth.reference_frame_dict(sp, synth.class_models(500)[0])
    pred = {"xyz": torch.from_numpy(fr["xyz"])[None].to(dev)}
    data = {"extent": torch.from_numpy(fr["extent"])[None], "lfborder": torch.from_numpy(fr["lfborder"])[None],
            "choose": torch.from_numpy(fr["choose"])[None, None], "intrinsic": torch.from_numpy(fr["intrinsic"])[None],
            "x_map_choosed": torch.from_numpy(fr["x_map_choosed"])[None, :, None],
            "y_map_choosed": torch.from_numpy(fr["y_map_choosed"])[None, :, None],
            "target_r": torch.from_numpy(fr["target_r"])[None], "target_t": torch.from_numpy(fr["target_t"])[None],
            "target": torch.from_numpy(fr["target"])[None], "model_points": torch.from_numpy(fr["model_points"])[None],
            "cls_id": torch.LongTensor([fr["cls_id"]])}
    data = {k: v.to(dev) for k, v in data.items()}                 # dataloader tensors already on the device
    metric = pb.Metric(synth.SYM_CLS)
    for _ in range(20):
        R, t = pb.get_pose(pred, data); pb.cal_dis(metric, R, t, data)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 200
    for _ in range(n):
        R, t = pb.get_pose(pred, data)
        pb.cal_dis(metric, R, t, data)
    out["dropin_call_sites_ms_per_frame"] = (time.perf_counter() - t0) / n * 1e3
    sw = pb.EvalSweep(models, diam, sym, 1, H, W, 4096, ops.RansacConfig(n_hyp=256, seed=0x5EED), device=dev)
    one = {k: v[:1] for k, v in inp.items()}
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            sw.step(one["xyz"], one["mask"], one["extent"], one["lfborder"], one["K4"], one["cls"], one["R_gt"], one["t_gt"])
    torch.cuda.synchronize()
    gph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gph, stream=s):
        sw.step(one["xyz"], one["mask"], one["extent"], one["lfborder"], one["K4"], one["cls"], one["R_gt"], one["t_gt"])
    out["F1_graph_replay_ms"] = timed(gph.replay, 200, warm=10)
    if getattr(args, "cpu", False):
        import bench
        n_cpu = 4 * (os.cpu_count() or 1)
        times, workers, n_cpu, succ = bench.run_reference_cpu(n_cpu, 3, spec=dict(config_id=1, kw=dict(cls=0)))
        out["cpu_baseline"] = {"value": n_cpu / min(times[1:]), "unit": "frames/s", "cores": workers, "kind": "port",
                               "ms_per_frame_per_core": 1e3 * min(times[1:]) * workers / n_cpu,
                               "sample": f"{n_cpu} ape frames, best of 2 passes; oracle/ref_callsites.eval_frame (the exact "
                                         "reference call: cv2.solvePnPRansac(EPNP) on 256 random masked px + ADD), one process per core"}
    print(json.dumps(out), flush=True)
    return out


def c5(args, dev):
    """65,536 DISTINCT frames, sharded contiguously over the ranks (torchrun; one rank per GPU) and
    streamed in chunks of 2048 frames per GPU: whole chunks are dealt in turn to three streams, each of which erases
    its previous chunk from its own slab, has the device-side Philox generator (csrc/synthgen.cu) write the new one
    and evaluates it -- so one lane's generation runs under the other lanes' evaluation; the sweep ends with the ONE all-reduce of the accuracy table (NCCL when
    W > 1).  The timed region contains generation, evaluation and the all-reduce; CUDA events between barriers,
    maximum over ranks."""
    import os
    import torch.distributed as dist
    import bench
    import pose_estimation_b200 as pb
    from pose_estimation_b200 import sweep as sweep_mod
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    F, total = 2048, args.c5_frames
    lo, hi = sweep_mod.shard_range(total, rank, world)
    assert (hi - lo) % F == 0, "shard must be a whole number of chunks"
    nchunk = (hi - lo) // F
    H, W = synth.IMG_H, synth.IMG_W
    P = synth.c5_params(bench.SEED, lo, hi - lo)                  # host, untimed: 48 floats per frame
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    prm, cls, ext, lfb, K4, Rg, tg = (t(P[k]) for k in ("params", "cls", "extent", "lfborder", "K4", "R_gt", "t_gt"))
    bb = [int(((P["params"][c * F:(c + 1) * F, 33] - P["params"][c * F:(c + 1) * F, 32]) *
               (P["params"][c * F:(c + 1) * F, 35] - P["params"][c * F:(c + 1) * F, 34])).max()) for c in range(nchunk)]
    n_lanes = 3
    gens = [ops.FrameGenerator(F, H, W, bench.SEED, device=dev) for _ in range(n_lanes)]     # one slab per lane
    models = torch.from_numpy(synth.class_models(bench.N_MODEL))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    sw = pb.EvalSweep(models, diam, sym, F, H, W, bench.CAP, ops.RansacConfig(n_hyp=bench.N_HYP, seed=bench.SEED), device=dev,
                      lanes=n_lanes, deal="alternate")      # whole chunks dealt in turn to three streams, joined by finish()
    main = torch.cuda.current_stream(dev)

    def sweep():
        sw.reset()
        for c in range(nchunk):
            k = sw.next_lane()                                  # chunk c goes to lane k; lane k owns slab k, so erasing its
            sl = slice(c * F, (c + 1) * F)                      # previous chunk, writing this one and evaluating it are
            with torch.cuda.stream(sw.lane_streams[k]):         # one stream's work: no cross-stream wait in the sweep
                gens[k].erase()
                xyz, mask = gens[k].write(prm[sl], lo + c * F, bb[c])
            sw.step(xyz, mask, ext[sl], lfb[sl], K4[sl], cls[sl], Rg[sl], tg[sl], frame0=lo + c * F)
        tb = sw.finish()
        return tb

    def cleanup():
        torch.cuda.synchronize(dev)
        for g in gens:
            g.erase()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    sweep(); cleanup()
    best = None
    for _ in range(2):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        tb = sweep()
        e1.record()
        barrier()
        cleanup()
        tm = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        best = float(tm.item()) if best is None else min(best, float(tm.item()))
    summ = pb.summarize(tb, synth.OBJLIST)
    if rank == 0:
        print(json.dumps({"config": "C5", "frames": total, "chunk": F, "n_gpus": world, "sweep_ms": best,
                          "frames_per_s": total / (best * 1e-3),
                          "step_impl": f"EvalSweep(lanes={n_lanes}, deal='alternate'), each lane generating into its own slab (FrameGenerator)",
                          "mask_bytes_streamed": total * H * W, "dense_map_bytes_represented": total * (3 * H * W * 4 + H * W),
                          "frames_counted": summ["frames"], "succ_add": summ["succ_base_rt"], "failed": summ["failed"],
                          "data": "synthetic: every frame distinct, generated on the device (Philox4x32-10, csrc/synthgen.cu) "
                                  "inside the timed region"}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def dense(args, dev):
    """Dense, compact silhouettes (every pixel of the ellipsoid: P = 2 k .. 45 k, like a real LineMOD mask) through
    the mask-first K0 and the whole sweep -- the mask shape SURVEY.md 8(d) asked for; C2's mask is a subset of it."""
    import pose_estimation_b200 as pb
    F, cap = 256, 49152
    H, W = synth.IMG_H, synth.IMG_W
    sps = [synth.make_frame_sparse(0x5EED, 12, f % 64, dense=True, z_range=(0.45, 1.0)) for f in range(F)]
    xyz = torch.zeros((F, 3, H, W), dtype=torch.float32, device=dev); mask = torch.zeros((F, H, W), dtype=torch.uint8, device=dev)
    for i, sp in enumerate(sps):
        r, c = torch.from_numpy(sp["rows"].astype(np.int64)).to(dev), torch.from_numpy(sp["cols"].astype(np.int64)).to(dev)
        xyz[i, :, r, c] = torch.from_numpy(sp["xyzn"].T.copy()).to(dev)
        mask[i, r, c] = 1
    cls = np.array([s["cls"] for s in sps], np.int32)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    ext, lfb = t(synth.EXTENT[cls].astype(np.float32)), t(synth.LFBORDER[cls].astype(np.float32))
    K4, Rg, tg = t(np.tile(synth.LINEMOD_K4, (F, 1))), t(np.stack([s["R_gt"] for s in sps])), t(np.stack([s["t_gt"] for s in sps]))
    comp = ops.MapCompactor(F, H, W, cap, device=dev)
    ms_k0 = timed(lambda: comp(xyz, mask, ext, lfb), args.iters)
    n_px = int(sum(s["rows"].shape[0] for s in sps))
    strict = F * H * W + n_px * 32
    sw = pb.EvalSweep(torch.from_numpy(synth.class_models(500)), torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM),
                      F, H, W, cap, ops.RansacConfig(n_hyp=256, seed=0x5EED), device=dev)
    ms = timed(lambda: sw.step(xyz, mask, ext, lfb, K4, t(cls), Rg, tg), args.iters)
    sw.reset(); sw.step(xyz, mask, ext, lfb, K4, t(cls), Rg, tg)
    summ = pb.summarize(sw.finish(), synth.OBJLIST)
    res = {"config": "dense silhouettes", "frames": F, "P_mean": n_px / F, "P_max": max(s["rows"].shape[0] for s in sps),
           "k0_ms": ms_k0, "k0_GBps_strict_bytes": strict / (ms_k0 * 1e-3) / 1e9, "sweep_ms": ms, "frames_per_s": F / (ms * 1e-3),
           "hyp_pts_per_s": 256.0 * n_px / (ms * 1e-3), "succ_add": summ["succ_base_rt"], "failed": summ["failed"]}
    print(json.dumps(res), flush=True)
    return res


def knn_large(args, dev):
    """The large-shape nearest-neighbour call of the transparent-object loss (version/transparent/lib/networks/
    loss.py:228): 250,000 predicted points against a 500-point target, and a 1000 x (1000 x 1000) batch."""
    out = {"config": "KNN large shapes"}
    g = torch.Generator(device=dev).manual_seed(5)
    for name, B, N, M in (("250000x500", 1, 250000, 500), ("1000x1000x1000", 1000, 1000, 1000)):
        x = torch.rand(B, N, 3, device=dev, generator=g) * 0.2
        y = torch.rand(B, M, 3, device=dev, generator=g) * 0.2
        r = {}
        for impl in ("simt", "tc"):
            ms = timed(lambda: ops_metric.knn1(x, y, impl=impl), args.iters)
            r[impl + "_ms"] = ms
            r[impl + "_pairs_per_s"] = B * N * M / (ms * 1e-3)
        r["equal"] = bool(torch.equal(ops_metric.knn1(x, y, impl="tc"), ops_metric.knn1(x, y, impl="simt")))
        st = ops_metric.LAST_TC_STATS.cpu().tolist()
        r["fallback_rows"], r["exact_evals_per_row"] = st[1], st[2] / (B * N)
        out[name] = r
    print(json.dumps(out), flush=True)


def f3(args, dev):
    """B problems of n = 1000 correspondences, 128 iterations (lib/transform/umeyama.py defaults),
    40 % outliers; beside it the C port of the same algorithm on one host core (bounded sample)."""
    import time
    from pose_estimation_b200 import ops_sim3
    B, n = args.f3_problems, 1000
    rng = np.random.default_rng(33)
    src = ((rng.random((B, n, 3)) - 0.5) * 0.2).astype(np.float32)
    dst = np.empty_like(src)
    for b in range(B):
        R = synth.random_rotation(np.random.default_rng(b))
        d = src[b] @ R.T.astype(np.float32) + rng.normal(size=3).astype(np.float32) * 0.1
        out = rng.random(n) < 0.4
        d[out] = (rng.random((int(out.sum()), 3)) - 0.5).astype(np.float32)
        dst[b] = d
    sd, dd = torch.from_numpy(src).to(dev), torch.from_numpy(dst).to(dev)
    cnt = torch.full((B,), n, dtype=torch.int32, device=dev)
    ms = timed(lambda: ops_sim3.sim3_ransac(sd, dd, cnt), max(args.iters, 10))
    r = ops_sim3.sim3_ransac(sd, dd, cnt)
    ok = float((r.status == 0).float().mean())
    out = {"config": "f3 sim3", "problems": B, "pairs": n, "n_iter": 128, "ms": ms, "problems_per_s": B / (ms * 1e-3),
           "hyp_pairs_per_s": B * 128.0 * n / (ms * 1e-3), "status_ok_frac": ok}
    try:
        from oracle import cref
        m = min(B, 64)
        t0 = time.perf_counter()
        for b in range(m):
            cref.sim3_ransac(src[b], dst[b], prob=b)
        dt = time.perf_counter() - t0
        out["cpu_port"] = {"ms_per_problem": dt / m * 1e3, "problems_per_s_per_core": m / dt, "cores": 1,
                           "sample": f"{m} problems, C restatement of the same algorithm (all 128 iterations scored)"}
    except Exception as e:                                   # the oracle is optional here
        out["cpu_port"] = {"unavailable": repr(e)[:100]}
    print(json.dumps(out), flush=True)


def f4(args, dev):
    """back-propagatable PnP (lib/network/dnn/BPnP.py) on B keypoint sets of n points: forward
    (RANSAC + all-point LM), the LM step alone, backward; beside it the reference's algorithm (cv2 + the autograd
    double-backward loop, float32, as restated in oracle/bpnp_ref.py) on one host core over a bounded sample."""
    import time
    from pose_estimation_b200 import ops_bpnp
    K = np.array([[572.4114, 0.0, 325.2611], [0.0, 573.57043, 242.04899], [0.0, 0.0, 1.0]], np.float32)
    rows = []
    for B, n in ((args.f4_batch, 64), (args.f4_batch, 1000)):
        rng = np.random.default_rng(44)
        z = ((rng.random((n, 3)) - 0.5) * np.array([0.10, 0.08, 0.14])).astype(np.float32)
        x = np.empty((B, n, 2), np.float32)
        for b in range(B):
            R = synth.random_rotation(np.random.default_rng(b)).astype(np.float64)
            t = np.array([rng.uniform(-0.15, 0.15), rng.uniform(-0.1, 0.1), rng.uniform(0.6, 1.2)])
            p = (z.astype(np.float64) @ R.T + t) @ K.astype(np.float64).T
            x[b] = p[:, :2] / p[:, 2:3] + rng.normal(size=(n, 2)) * 0.5
        xd, zd, Kd = torch.from_numpy(x).to(dev), torch.from_numpy(z).to(dev), torch.from_numpy(K).to(dev)
        P, st = ops_bpnp.bpnp_forward(xd, zd, Kd)
        go = torch.randn((B, 6), device=dev)
        it = max(args.iters, 10)
        ms_f = timed(lambda: ops_bpnp.bpnp_forward(xd, zd, Kd), it)
        ms_r = timed(lambda: ops_bpnp.pnp_refine(xd, zd, Kd, pose6_init=P), it)
        ms_b = timed(lambda: ops_bpnp.bpnp_backward(xd, zd, Kd, P, go), it)
        row = {"batch": B, "points": n, "forward_ms": ms_f, "refine_only_ms": ms_r, "backward_ms": ms_b,
               "status_ok_frac": float((st == 0).float().mean()), "fwd_bwd_samples_per_s": B / ((ms_f + ms_b) * 1e-3)}
        try:
            from oracle import bpnp_ref
            m = 2 if n > 100 else 4
            t0 = time.perf_counter()
            Pc = bpnp_ref.bpnp_forward(x[:m], z, K)
            t1 = time.perf_counter()
            bpnp_ref.bpnp_backward(x[:m], z, K, Pc, go[:m].cpu().numpy(), dtype=torch.float32)
            t2 = time.perf_counter()
            row["cpu_port"] = {"forward_ms_per_sample": (t1 - t0) / m * 1e3, "backward_ms_per_sample": (t2 - t1) / m * 1e3,
                               "cores": torch.get_num_threads(), "sample": f"{m} samples, cv2 + torch autograd as the reference"}
        except Exception as e:
            row["cpu_port"] = {"unavailable": repr(e)[:100]}
        rows.append(row)
    print(json.dumps({"config": "f4 bpnp", "rows": rows}), flush=True)
    # PossLoss at the reference's own shape (loss.py:228: 250 000 predicted points against 500 target points)
    import pose_estimation_b200 as pb
    from pose_estimation_b200 import ops_metric
    num_p, N = 500, 500
    rng = np.random.default_rng(45)
    model = ((rng.random((1, N, 3)) - 0.5) * np.array([0.10, 0.08, 0.14])).astype(np.float32)
    Rg = synth.random_rotation(np.random.default_rng(1)).astype(np.float64)
    tg = np.array([0.05, -0.03, 0.8])
    target = (model.astype(np.float64) @ Rg.T + tg).astype(np.float32)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
    pr = T(rng.normal(size=(1, num_p, 4))).requires_grad_()
    pt = T(tg[None, None] + rng.normal(size=(1, num_p, 3)) * 0.01).requires_grad_()
    pc = T(rng.uniform(0.05, 0.95, size=(1, num_p, 1))).requires_grad_()
    md, td, ax, tr = T(model), T(target), T([[1.0, 0.0, 1.0]]), T(Rg[None])
    lossfn = pb.PossLoss(N, [7, 8])
    from pose_estimation_b200 import ops_poss
    gv = torch.tensor([1.0, 0.0, 1.0], device=dev)
    out = {"config": "f4 possloss", "hypotheses": num_p, "points": N, "pairs": num_p * N * N}
    for mode, name in ((1, "symmetric"), (0, "indexed")):
        vals, ws = ops_poss.poss_loss_fwd(pr.detach(), pt.detach(), pc.detach(), td, md, mode, 0.015, ax, tr)
        it = max(args.iters, 20)
        out[name + "_fwd_ms"] = timed(lambda: ops_poss.poss_loss_fwd(pr.detach(), pt.detach(), pc.detach(), td, md, mode, 0.015, ax, tr), it)
        out[name + "_bwd_ms"] = timed(lambda: ops_poss.poss_loss_bwd(pr.detach(), pt.detach(), pc.detach(), td, md, mode, 0.015, ax, tr, gv, ws), it)
    out["pairs_per_s"] = out["pairs"] / (out["symmetric_fwd_ms"] * 1e-3)

    def module_step():
        for v in (pr, pt, pc):
            v.grad = None
        loss, _, lrot = lossfn(pr, pt, pc, td, md, 8, False, 0.015, ax, tr)
        (loss + lrot).backward()
    out["module_fwd_bwd_ms"] = timed(module_step, it)

    def unfused():                                               # the same value from the separate kernels: R in torch, KNN op
        base = pb.PossLoss.predr2rotation(pr.detach(), 1, num_p)
        pred = torch.baddbmm(pt.detach().view(num_p, 1, 3), md.expand(num_p, N, 3), base.transpose(1, 2))
        idx = ops_metric.knn1(pred.reshape(1, -1, 3).contiguous(), td.reshape(1, N, 3))
        return (pred.reshape(-1, 3) - td[0][idx.view(-1)]).norm(dim=1).view(num_p, N).mean(1)
    out["unfused_fwd_ms_without_allo_to_ego"] = timed(unfused, it)
    print(json.dumps(out), flush=True)


def region(args, dev):
    """Region-coded maps (tools/script/eval.py:93-103, loss_utils.get_xyz): dense decode of F full-size frames
    (HBM-bound: (n + 3) * 4 B read + 12 B written per pixel) and the decode of the compacted foreground pixels only."""
    F, n, H, W = 16, 65, synth.IMG_H, synth.IMG_W
    g = torch.Generator(device=dev).manual_seed(1)
    off = torch.randn(F, 3, H, W, device=dev, generator=g) * 0.05
    lg = torch.randn(F, n, H, W, device=dev, generator=g)
    rp = torch.rand(F, n, 3, device=dev, generator=g)
    out = {"config": "region", "frames": F, "n_region": n, "map": [H, W]}
    byt = F * H * W * ((n + 3) * 4 + 12)
    for mode in ("hard", "soft"):
        ms = timed(lambda: ops.region_decode_map(off, lg, rp, mode=mode), args.iters)
        out["dense_" + mode] = {"ms": ms, "GBps": byt / (ms * 1e-3) / 1e9}
    sps = synth.make_batch_sparse(0x5EED, 2, F)
    mask = torch.zeros((F, H, W), dtype=torch.uint8, device=dev)
    for f, sp in enumerate(sps):
        mask[f, torch.from_numpy(sp["rows"]).long().to(dev), torch.from_numpy(sp["cols"]).long().to(dev)] = 1
    cls = np.array([s["cls"] for s in sps])
    ext = torch.from_numpy(synth.EXTENT[cls].astype(np.float32)).to(dev)
    lfb = torch.from_numpy(synth.LFBORDER[cls].astype(np.float32)).to(dev)
    pts, cnt = ops.compact_map(off, mask, ext, lfb, 1280)
    npx = int(cnt.sum())
    for mode in ("hard", "soft"):
        ms = timed(lambda: ops.region_decode_points(pts, cnt, off, lg, rp, ext, lfb, mode=mode), args.iters)
        out["points_" + mode] = {"ms": ms, "pixels": npx, "Mpx_per_s": npx / (ms * 1e-3) / 1e6}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--f4-batch", type=int, default=1024)
    ap.add_argument("--f3-problems", type=int, default=1024)
    ap.add_argument("--c5-frames", type=int, default=65536)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--c3-batch", type=int, default=4096)
    ap.add_argument("--c4-frames", type=int, default=256)
    ap.add_argument("--only", default="")
    ap.add_argument("--cpu", action="store_true", help="time the reference's CPU path beside C1 / C3 / C4 (bounded samples)")
    args = ap.parse_args()
    import os
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 and args.only != "c5":
        raise SystemExit("bench_configs.py: only --only c5 runs under torchrun")
    if args.only in ("", "c1"):
        c1(args, dev)
    if args.only in ("", "c3"):
        c3(args, dev)
    if args.only in ("", "c4"):
        c4(args, dev)
    if args.only in ("", "dense"):
        dense(args, dev)
    if args.only in ("", "c5"):
        c5(args, dev)
    if args.only in ("", "region"):
        region(args, dev)
    if args.only in ("", "knn"):
        knn_large(args, dev)
    if args.only in ("", "f3"):
        f3(args, dev)
    if args.only in ("", "f4"):
        f4(args, dev)


if __name__ == "__main__":
    main()
