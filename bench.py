#!/usr/bin/env python
"""bench.py -- frames/s of the PnP-RANSAC + ADD(-S) evaluation pass (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one evaluation pass over a batch of 1024 synthetic LineMOD-shaped frames per GPU
(BASELINE.json configs[1]: 13 objects, 480x640 coordinate map + mask, ~1k masked px, 256 RANSAC
hypotheses, ADD with 500-pt models, ADD-S for eggbox/glue).  The timed region is a sweep of K such
steps closed by the ONE all-reduce of the [24,13] accuracy table that the path has (NCCL when N > 1;
weak scaling: every rank evaluates its own 1024-frame shard per step).

  value     whole-job frames/s with the maps already resident in HBM (CUDA events, max over ranks): K steps of
            EvalSweep(lanes=3, deal="alternate") -- whole batches dealt in turn to three streams, joined once by
            finish() -- with no events inside the region; `instrumented` is the same K steps on one stream with
            events between the kernels, the pass the per-kernel times and the roofline come from
  e2e       the same pass through HostEvalSweep.run_choose with everything in PINNED HOST memory in the
            reference's own form (maps + int64 choose lists + f32 pixel-coordinate lists, tools/trainer.py:
            383-421), timed per step: the lists go by the copy engine, the gather kernel reads the maps in
            place (the bytes that cross PCIe are counted in h2d_bytes_per_step), the table is all-reduced and
            read back every step.  Variants beside it: maps already on the device (the network's output),
            maps + u8 masks on the host (round 1's headline), whole maps copied host->device
  roofline  the dominant kernel of the step (K3 hypothesis scoring, FP32-pipe bound) timed live with
            CUDA events against an FFMA probe run in the same process; roofline_all adds the HBM-bound
            K0 kernels (mask-first as used in the step -- against its strict algorithmic bytes: 1 B/px of
            mask + 12 B read + 20 B written per foreground pixel -- and the maskless full stream)
  cpu_baseline / --impl reference
            the reference's own per-frame CPU path (tools/trainer.py:166-191: 256-point
            cv2.solvePnPRansac + ADD(-S)) restated in oracle/ref_callsites.py around the real cv2,
            one worker process per host core, on the first >= 256 frames of the same batch (median of 5
            passes; the reference arm times K steps of one pass each); both arms print their success
            rates on that sample, and the GPU arm its verdict agreement with the reference's own run
            over all 1024 frames (tests/golden/c2_ref_verdicts.npz).

Nothing here reads /root/reference at run time.
"""
from __future__ import annotations

import argparse
import contextlib
import ctypes as C
import io
import json
import multiprocessing as mp
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SEED = 0x5EED
CONFIG_ID = 2                       # C2 of SURVEY.md section 8(d)
FRAMES_PER_GPU = 1024
N_HYP = 256
N_MODEL = 500
CAP = 1280                          # >= max masked px per frame (P ~ U{800..1200})
METRIC = "frames/sec PnP-RANSAC+ADD(-S) eval"
UNIT = "frames/s"
MAP_BYTES_PER_FRAME = 480 * 640 * (12 + 1)      # xyz 12 B + mask 1 B per pixel


# ------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's per-frame CPU path on the host cores
def _ref_worker(rank, frames, n_steps, barrier, out_q, cv2_threads=1, spec=None):
    """Worker process: builds its frames (untimed), then evaluates them once per step."""
    try:
        import numpy as np
        import cv2
        if cv2_threads is not None:
            cv2.setNumThreads(cv2_threads)          # one process per core; no nested OpenCV threads
        from oracle import ref_callsites as rc
        from pose_estimation_b200 import synth
        models = synth.class_models(N_MODEL)
        metric = rc.Metric(synth.SYM_CLS)
        fr = []
        spec = spec or {}
        for f in frames:
            sp = synth.make_frame_sparse(SEED, spec.get("config_id", CONFIG_ID), f, **spec.get("kw", {}))
            fr.append(synth.reference_frame_dict(sp, models[sp["cls"]]))
        rng = np.random.default_rng([SEED, rank])
        for _ in range(n_steps):
            barrier.wait()
            t0 = time.time()
            n_ok = [0, 0, 0]
            for d in fr:
                r = rc.eval_frame(metric, d, synth.DIAMETER, rng)
                n_ok[0] += int(r["succ_rt"]); n_ok[1] += int(r["succ_r"]); n_ok[2] += int(r["succ_t"])
            out_q.put((rank, t0, time.time(), n_ok))
    except Exception as e:                          # pragma: no cover
        out_q.put((rank, "error", repr(e), [0, 0, 0]))
        raise


def run_reference_cpu(n_frames, n_steps, n_workers=None, cv2_threads=1, spec=None):
    """Times the reference path on ``n_frames`` frames of the workload, ``n_steps`` times.
    cv2_threads=None leaves OpenCV's own threading at its default (the reference as it is launched: one process).
    Returns (list of seconds per step, workers, frames, [succ_add, succ_r, succ_t] counts of the last step)."""
    n_workers = n_workers or max(1, min(os.cpu_count() or 1, n_frames))
    ctx = mp.get_context("spawn")
    barrier = ctx.Barrier(n_workers + 1)
    q = ctx.Queue()
    shards = [list(range(w, n_frames, n_workers)) for w in range(n_workers)]
    procs = [ctx.Process(target=_ref_worker, args=(w, shards[w], n_steps, barrier, q, cv2_threads, spec), daemon=True)
             for w in range(n_workers)]
    for p in procs:
        p.start()
    times, succ = [], [0, 0, 0]
    for _ in range(n_steps):
        barrier.wait()
        res = [q.get() for _ in range(n_workers)]
        for r in res:
            if r[1] == "error":
                raise RuntimeError(f"reference worker failed: {r[2]}")
        times.append(max(r[2] for r in res) - min(r[1] for r in res))
        succ = [sum(r[3][k] for r in res) for k in range(3)]
    for p in procs:
        p.join(timeout=30)
    return times, n_workers, n_frames, succ


def cpu_sample_frames():
    """Bounded sample of the workload for the CPU legs: >= 256 frames and >= 16 per core (>= 200
    frames per measurement), the FIRST frames of the batch the GPU arm evaluates."""
    cores = os.cpu_count() or 1
    return min(FRAMES_PER_GPU, max(256, 16 * cores))


def accuracy_block(n_frames, succ):
    return {"frames": n_frames, "succ_add": succ[0] / n_frames, "succ_r5deg": succ[1] / n_frames,
            "succ_t5cm": succ[2] / n_frames}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_frames = cpu_sample_frames()
    times, workers, n_frames, succ = run_reference_cpu(n_frames, args.warmup + args.steps)
    timed = times[args.warmup:]
    total = sum(timed)
    fps = n_frames * len(timed) / total
    per_step = sorted(n_frames / t for t in timed)
    import cv2
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(timed),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "gpu_launches": 0,
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": workers, "kind": "port",
                         "median_step": statistics.median(per_step), "min_step": per_step[0], "max_step": per_step[-1],
                         "sample": f"the first {n_frames} frames of the workload per step, {len(timed)} timed steps; "
                                   f"oracle/ref_callsites.eval_frame = real cv2 {cv2.__version__} solvePnPRansac(EPNP) "
                                   "on 256 random masked px + numpy ADD/ADD-S; one process per host core"},
        "accuracy": accuracy_block(n_frames, succ),
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def workload_config(n_gpus):
    return {"workload": "C2: 13-object LineMOD-shaped eval pass, 1024 synthetic frames per GPU, 480x640 "
                        "coordinate map + mask, P~U{800..1200} masked px, 256 RANSAC hypotheses, 500-pt models "
                        "(ADD; ADD-S for eggbox/glue)",
            "frames_per_gpu": FRAMES_PER_GPU, "n_hyp": N_HYP, "model_points": N_MODEL,
            "map": [480, 640], "parallelism": f"frames sharded over {n_gpus} rank(s); one all-reduce of the [24,13] table",
            "l2_policy": "inputs (4.09 GB per step per GPU) exceed the 126 MB L2; no flush needed"}


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            p = [x.strip() for x in r.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1])); smax.append(float(p[2])); pw.append(float(p[3]))
            except ValueError:
                continue
            for nm, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def build_inputs(torch, dev, first_frame, n_frames):
    """Synthetic frames [first_frame, first_frame+n) -> device tensors (generation is untimed)."""
    import numpy as np
    from concurrent.futures import ThreadPoolExecutor
    from pose_estimation_b200 import synth
    H, W = synth.IMG_H, synth.IMG_W
    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as ex:
        sps = list(ex.map(lambda f: synth.make_frame_sparse(SEED, CONFIG_ID, f), range(first_frame, first_frame + n_frames)))
    xyz = torch.zeros((n_frames, 3, H, W), dtype=torch.float32, device=dev)
    mask = torch.zeros((n_frames, H, W), dtype=torch.uint8, device=dev)
    fidx = np.concatenate([np.full(s["rows"].shape[0], i, np.int64) for i, s in enumerate(sps)])
    rows = np.concatenate([s["rows"] for s in sps]).astype(np.int64)
    cols = np.concatenate([s["cols"] for s in sps]).astype(np.int64)
    vals = np.concatenate([s["xyzn"] for s in sps]).astype(np.float32)
    fi, ri, ci = (torch.from_numpy(a).to(dev) for a in (fidx, rows, cols))
    v = torch.from_numpy(vals).to(dev)
    for c in range(3):
        xyz[fi, c, ri, ci] = v[:, c]
    mask[fi, ri, ci] = 1
    cls = np.array([s["cls"] for s in sps], np.int32)
    out = dict(xyz=xyz, mask=mask, cls=torch.from_numpy(cls).to(dev),
               extent=torch.from_numpy(synth.EXTENT[cls].astype(np.float32)).to(dev),
               lfborder=torch.from_numpy(synth.LFBORDER[cls].astype(np.float32)).to(dev),
               K4=torch.from_numpy(np.tile(synth.LINEMOD_K4, (n_frames, 1))).to(dev),
               R_gt=torch.from_numpy(np.stack([s["R_gt"] for s in sps])).to(dev),
               t_gt=torch.from_numpy(np.stack([s["t_gt"] for s in sps])).to(dev),
               n_px=int(rows.shape[0]), counts=[int(s["rows"].shape[0]) for s in sps],
               rc=[(s["rows"], s["cols"]) for s in sps])
    return out


def k3_alternatives(torch, dev, inp, cfg, cap, steps):
    """The scoring stage (K3) of the same batch through the two tensor-core formulations kept in the library beside the
    FP32-pipe kernel the step uses (profiles/r02_k3_tc.md): stage time from the library's own CUDA events, and whether
    every integer and the pose equal the FP32-pipe kernel's."""
    import ctypes as C
    from pose_estimation_b200 import _lib, ops
    lib = _lib.load()
    pts, count = ops.compact_map(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], cap)
    F = pts.shape[0]
    out, ref = {}, None
    try:
        for impl in ("fp32", "tc", "mma"):
            ops.set_score_impl(impl)
            o = ops.pnp_ransac_pts(pts, count, inp["K4"], cfg, want_counts=True)
            pr = ops.PtsRansac(F, cap, cfg, device=dev)
            for _ in range(3):
                pr(pts, count, inp["K4"])
            torch.cuda.synchronize()
            lib.pb_stage_timing(1)
            for _ in range(steps):
                pr(pts, count, inp["K4"])
            ms3, ncalls = (C.c_float * 3)(), C.c_int(0)
            lib.pb_stage_timing_read(ms3, C.byref(ncalls))
            lib.pb_stage_timing(0)
            if ref is None:
                ref = o
            same = all(bool(torch.equal(o[k], ref[k])) for k in ("hyp_counts", "best_hyp", "n_inl", "status", "R", "t"))
            out[impl] = {"score_stage_ms": round(float(ms3[1]), 5), "identical_to_fp32": same}
            if impl != "fp32":
                st = ops.ransac_tc_stats(pr.ws, F, cap, cfg)
                out[impl].update(barrier_timeouts=st["timeout"], reevaluated_blocks=st["reevaluated_groups"])
    finally:
        ops.set_score_impl("auto")
    out["note"] = ("fp32 = ransac_score_bal_kernel (what the step runs); tc = tcgen05 + TMEM (csrc/ransac_score_tc.cu); "
                   "mma = warp-level mma.sync (csrc/ransac_score_mma.cu, its frame-constants kernel included); band x1")
    return out


def ours_arm(args):
    import torch
    import torch.distributed as dist
    import pose_estimation_b200 as pb
    from pose_estimation_b200 import _lib, ops, synth, sweep

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (pose_estimation_b200 has no CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    _lib.load()                                     # fail loudly if libposeb200.so is missing
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    F = FRAMES_PER_GPU
    H, W = synth.IMG_H, synth.IMG_W
    inp = build_inputs(torch, dev, rank * F, F)
    models = torch.from_numpy(synth.class_models(N_MODEL))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    cfg = ops.RansacConfig(n_hyp=N_HYP, seed=SEED)
    sw = pb.EvalSweep(models, diam, sym, F, H, W, CAP, cfg, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def one_step(events=None):
        sw.step(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], inp["K4"], inp["cls"], inp["R_gt"],
                inp["t_gt"], frame0=rank * F, events=events)

    # ---- device-resident throughput ------------------------------------------------------------
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()                             # samples warm-up + both timed regions
    for _ in range(max(args.warmup, 3)):
        one_step()
    sw.finish()
    barrier()
    evs = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    lib = _lib.load()
    lib.pb_stage_timing(1)                          # CUDA events around the three RANSAC launches
    barrier()
    sw.reset()
    e0.record()
    for _ in range(args.steps):                     # the timed region is one sweep of K batches ...
        one_step(evs)
    sw.finish()                                     # ... closed by the ONE all-reduce of the accuracy table
    e1.record()
    barrier()
    sw.reset()                                      # single-pass table (untimed): the e2e path must reproduce it
    per_frame = [x.clone() for x in sw.step(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], inp["K4"], inp["cls"],
                                            inp["R_gt"], inp["t_gt"], frame0=rank * F)]
    table = sw.finish().clone()
    ms3, ncalls = (C.c_float * 3)(), C.c_int(0)
    lib.pb_stage_timing_read(ms3, C.byref(ncalls))
    lib.pb_stage_timing(0)
    hyp_ms, score_ms, final_ms = float(ms3[0]), float(ms3[1]), float(ms3[2])
    # ---- the same sweep as the product runs it: whole batches dealt in turn to three streams that are not joined
    # between batches (EvalSweep(lanes=3, deal="alternate")), so every kernel keeps its full-size single-wave grid and
    # batch i+1's HBM-bound K0 runs under batch i's FP32-bound scoring / latency-bound refit; joined once by finish().
    # No events inside this region: `value` is this clean pass, the per-kernel times above come from the instrumented
    # single-stream pass (same K steps, same inputs, same process).
    n_lanes = int(os.environ.get("PB_LANES", "3"))
    sw2 = pb.EvalSweep(models, diam, sym, F, H, W, CAP, cfg, device=dev, lanes=n_lanes, deal="alternate")

    def lane_step():
        sw2.step(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], inp["K4"], inp["cls"], inp["R_gt"],
                 inp["t_gt"], frame0=rank * F)
    for _ in range(max(args.warmup, 3)):
        lane_step()
    sw2.reset()
    lane_step()
    same_lanes = bool(torch.equal(sw2.finish(), table))
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sw2.reset()
    c0.record()
    for _ in range(args.steps):
        lane_step()
    sw2.finish()
    c1.record()
    barrier()
    cmax = torch.tensor([c0.elapsed_time(c1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(cmax, op=dist.ReduceOp.MAX)
    ms_clean = float(cmax.item())
    ms_total = e0.elapsed_time(e1)
    tmax = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_total = float(tmax.item())
    value_instr = world * F * args.steps / (ms_total * 1e-3)
    value = world * F * args.steps / (ms_clean * 1e-3)
    stage_ms = [statistics.fmean(ev[i].elapsed_time(ev[i + 1]) for ev in evs) for i in range(4)]
    k0_ms = stage_ms[0]
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        hbm_peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        hbm_peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))

    def traffic_of(kernel):
        """DRAM bytes per launch from the committed ncu capture -- a MEASUREMENT OF ANOTHER RUN, so it is given with
        the commit the capture was taken at; null when the capture does not list this kernel."""
        v = traffic.get(kernel)
        if v is None:
            return None
        return {"dram_bytes_per_launch": v, "capture_commit": traffic.get("_commit"), "source": "profiles/ncu_traffic.json"}

    def timed(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(dev)
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(n):
            fn()
        a1.record()
        torch.cuda.synchronize(dev)
        return a0.elapsed_time(a1) / n

    # FP32 FMA peak of this GPU right now (scalar FFMA and packed FFMA2 probes; the larger one is the peak)
    probe_out = torch.zeros(4, dtype=torch.float32, device=dev)
    stream_ptr = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ctas, iters = 148 * 8, 4096
    fp32_peak = 0.0
    fp32_probe = {}
    for packed in (0, 1):
        ms = timed(lambda: lib.pb_probe_fma(ctas, iters, packed, C.c_void_p(probe_out.data_ptr()), stream_ptr), 5)
        tf = ctas * 256 * iters * 8 * 2 * (2 if packed else 1) / (ms * 1e-3) / 1e12
        fp32_probe["ffma2" if packed else "ffma"] = tf
        fp32_peak = max(fp32_peak, tf)
    # maskless K0 (every pixel streamed: the HBM-roofline kernel of SURVEY.md 8(d)), timed on the same maps
    dense = ops.MapCompactor(F, H, W, CAP, device=dev)
    dense_ms = timed(lambda: dense(inp["xyz"], None, inp["extent"], inp["lfborder"]), 10)
    dense_bytes = F * 480 * 640 * 12 + inp["n_px"] * 20
    sparse_bytes = F * 480 * 640 + inp["n_px"] * (12 + 20)       # 1 B/px mask + 12 B read + 20 B written per foreground px
    sparse_sector_bytes = F * 480 * 640 + inp["n_px"] * (3 * 32 + 20)   # what DRAM must move: one 32 B sector per plane
    score_flop = 27.0 * N_HYP * inp["n_px"]
    roof_score = {"kernel": "ransac_score_bal_kernel (K3: reprojection scoring of every masked pixel x 256 hypotheses; CTAs take equal ranges of the batch's point list)",
                  "bound": "fp32", "achieved": score_flop / (score_ms * 1e-3) / 1e12, "peak": fp32_peak,
                  "unit": "TFLOP/s", "frac": score_flop / (score_ms * 1e-3) / 1e12 / fp32_peak,
                  "traffic": traffic_of("ransac_score_bal_kernel"),
                  "peak_source": "FFMA/FFMA2 probe kernel timed in this run (MEASURED_PEAKS.json has no FP32 entry); "
                                 "nominal 148 SM x 128 lanes x 2 x 1.965 GHz = 74.4",
                  "algorithmic_flop_per_launch": score_flop, "flop_per_hyp_pt": 27, "launch_ms": score_ms,
                  "note": "FP32-pipe bound (SURVEY.md 8(d)); the kernel executes 14 FMA-pipe instructions per "
                          "hypothesis*point = 28 flop-slots, so its pipe occupancy is frac*28/27"}
    roof_k0 = {"kernel": "compact_stream_kernel (K0 mask-first: mask stream by ticketed tiles + foreground gather)",
               "bound": "hbm", "achieved": sparse_bytes / (k0_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
               "frac": sparse_bytes / (k0_ms * 1e-3) / 1e9 / hbm_peak, "traffic": traffic_of("compact_stream_kernel"),
               "peak_source": peak_src, "algorithmic_bytes_per_launch": sparse_bytes, "launch_ms": k0_ms,
               "sector_bytes_per_launch": sparse_sector_bytes,
               "frac_of_sector_bytes": sparse_sector_bytes / (k0_ms * 1e-3) / 1e9 / hbm_peak,
               "note": "algorithmic bytes = 1 B/px mask + per foreground px 12 B read + 20 B written; the scattered "
                       "foreground pixels of this workload cost one 32 B sector per plane (sector_bytes)"}
    roof_dense = {"kernel": "compact_kernel (K0 maskless: all three planes streamed; not part of the timed step)",
                  "bound": "hbm", "achieved": dense_bytes / (dense_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                  "frac": dense_bytes / (dense_ms * 1e-3) / 1e9 / hbm_peak, "traffic": traffic_of("compact_kernel"),
                  "peak_source": peak_src, "algorithmic_bytes_per_launch": dense_bytes, "launch_ms": dense_ms}

    # ---- end to end from pinned host memory --------------------------------------------------------
    import numpy as np
    xyz_h = torch.empty((F, 3, H, W), dtype=torch.float32).pin_memory()
    mask_h = torch.empty((F, H, W), dtype=torch.uint8).pin_memory()
    xyz_h.copy_(inp["xyz"]); mask_h.copy_(inp["mask"])
    cls_h = inp["cls"].cpu().pin_memory()
    small_src = [inp[k].cpu() for k in ("extent", "lfborder", "K4", "R_gt", "t_gt")]
    # the reference's host-side lists (dataset/linemod/batchdataset.py:673-687, 712-716): choose = flat indices of
    # the masked pixels (row-major), x / y_map_choosed = their pixel coordinates; ragged -> n_valid
    NPL = (max(inp["counts"]) + 15) // 16 * 16
    choose_np = np.zeros((F, NPL), np.int64)
    for f, (r_, c_) in enumerate(inp["rc"]):
        choose_np[f, :r_.shape[0]] = r_.astype(np.int64) * W + c_
    xm_np = (choose_np % W).astype(np.float32)
    ym_np = (choose_np // W).astype(np.float32)
    nv_np = np.asarray(inp["counts"], np.int32)
    sectors = sum(np.unique(choose_np[f, :nv_np[f]] >> 3).shape[0] for f in range(F))   # distinct 32 B sectors per plane
    choose_h, xm_h, ym_h, nv_h = (torch.from_numpy(a_).pin_memory() for a_ in (choose_np, xm_np, ym_np, nv_np))
    choose32_h = torch.from_numpy(choose_np.astype(np.int32)).pin_memory()
    torch.cuda.synchronize(dev)

    def run_e2e(kind, chunk, steps):
        zc = {"staged": False, "mask": "xyz"}.get(kind, True)
        hs = sweep.HostEvalSweep(models, diam, sym, H, W, CAP, chunk=chunk, cfg=cfg, device=dev, zero_copy=zc)
        small_h = sweep.pack_small(*small_src, chunk).pin_memory()

        replay = None
        if kind == "choose_graph":     # the same pass as "choose", captured once and replayed as a CUDA graph
            replay = hs.capture_choose(xyz_h, choose_h, small_h, cls_h, xmap_h=xm_h, ymap_h=ym_h, n_valid_h=nv_h, frame0=rank * F)
        elif kind == "choose_dev_graph":
            replay = hs.capture_choose(inp["xyz"], choose_h, small_h, cls_h, xmap_h=xm_h, ymap_h=ym_h, n_valid_h=nv_h, frame0=rank * F)
        elif kind == "choose_dev32_graph":
            replay = hs.capture_choose(inp["xyz"], choose32_h, small_h, cls_h, n_valid_h=nv_h, frame0=rank * F)
        pipe = None
        if kind == "choose_graph_pipe":  # two captured passes in flight (each with its own sweep object): a serving loop's double buffering
            hs2 = sweep.HostEvalSweep(models, diam, sym, H, W, CAP, chunk=chunk, cfg=cfg, device=dev, zero_copy=zc)
            pipe = sweep.PipelinedPasses([h_.capture_choose(xyz_h, choose_h, small_h, cls_h, xmap_h=xm_h, ymap_h=ym_h, n_valid_h=nv_h,
                                                            frame0=rank * F) for h_ in (hs, hs2)], device=dev)
        last = [None]

        def once():
            if pipe is not None:
                t_ = pipe.submit()
                if t_ is not None:
                    last[0] = t_
                return last[0]
            if replay is not None:
                return replay()
            hs.reset()
            if kind == "choose":           # the reference's interface: maps + int64 choose + f32 x / y lists, all on the host
                return hs.run_choose(xyz_h, choose_h, small_h, cls_h, xmap_h=xm_h, ymap_h=ym_h, n_valid_h=nv_h, frame0=rank * F)
            if kind == "choose_dev":       # the network's output still on the device; the lists from the host
                return hs.run_choose(inp["xyz"], choose_h, small_h, cls_h, xmap_h=xm_h, ymap_h=ym_h, n_valid_h=nv_h, frame0=rank * F)
            if kind == "choose_dev32":     # ... with int32 indices and the coordinates taken from the index
                return hs.run_choose(inp["xyz"], choose32_h, small_h, cls_h, n_valid_h=nv_h, frame0=rank * F)
            return hs.run(xyz_h, mask_h, small_h, cls_h, frame0=rank * F)
        tb = None
        for _ in range(max(args.warmup, 3)):
            once()
        if pipe is not None:
            pipe.drain()
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(steps):
            tb = once()
        if pipe is not None:               # every pass's table is read inside the timed region
            tb = pipe.drain()[-1]
        g1.record()
        barrier()
        ms = torch.tensor([g0.elapsed_time(g1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return world * F * steps / (float(ms.item()) * 1e-3), hs.h2d_bytes, tb.clone()

    nchunk = int(os.environ.get("PB_ZC_SPLIT", "4"))
    eager_value, list_bytes, tb_eager = run_e2e("choose", F // nchunk, args.steps)
    e2e_value, _, tb_host = run_e2e("choose_graph", F // nchunk, args.steps)
    e2e_bytes = list_bytes + sectors * 3 * 32
    pipe_value, _, tb_pipe = run_e2e("choose_graph_pipe", F // nchunk, args.steps)
    ndev = int(os.environ.get("PB_DEV_SPLIT", "2"))      # maps on the device: two chunks (one per lane) measure best
    dev_eager, dev_bytes, tb_dev = run_e2e("choose_dev", F // ndev, args.steps)
    dev32_eager, dev32_bytes, tb_dev32 = run_e2e("choose_dev32", F // ndev, args.steps)
    dev_value, _, tb_devg = run_e2e("choose_dev_graph", F // ndev, args.steps)
    dev32_value, _, tb_dev32g = run_e2e("choose_dev32_graph", F // ndev, args.steps)
    zc_value, zc_bytes, tb_zc = run_e2e("mask", F // nchunk, min(args.steps, 10))
    zc_bytes += sectors * 3 * 32
    staged_value, staged_bytes, tb_staged = run_e2e("staged", 128, min(args.steps, 5))
    tcpu = table.cpu()
    same = all(bool(torch.equal(t_, tcpu)) for t_ in (tb_host, tb_pipe, tb_eager, tb_dev, tb_dev32, tb_devg, tb_dev32g, tb_zc, tb_staged))

    clocks = sampler.stop() if sampler else None
    if rank == 0:
        summ = pb.summarize(table, synth.OBJLIST)
        cpu_base = None
        n_cpu = cpu_sample_frames()
        if world == 1 and not args.no_cpu_baseline:
            times, workers, n_frames, succ = run_reference_cpu(n_cpu, 6)
            fps = sorted(n_frames / t_ for t_ in times[1:])            # 5 timed passes after 1 warm-up
            import cv2
            cpu_base = {"value": statistics.median(fps), "unit": UNIT, "cores": workers, "kind": "port",
                        "min": fps[0], "max": fps[-1],
                        "sample": f"the first {n_frames} frames of the workload, median of 5 timed passes after 1 warm-up "
                                  f"(min / max beside it); oracle/ref_callsites.eval_frame around real cv2 {cv2.__version__} "
                                  "solvePnPRansac(EPNP, 256 px) + numpy ADD/ADD-S; one process per host core",
                        "accuracy": accuracy_block(n_frames, succ)}
            # the reference as it is launched: ONE process, OpenCV's default threading (SURVEY.md 8(d) variant i)
            t1, _, n1, _ = run_reference_cpu(32, 2, n_workers=1, cv2_threads=None)
            cpu_base["single_process"] = {"value": n1 / t1[1], "unit": UNIT, "cv2_threads": cv2.getNumThreads(),
                                          "sample": f"{n1} frames, second of 2 passes"}
        # verdicts of the GPU pass: on the CPU legs' sample, and against the reference's own run over all 1024 frames
        pf = torch.stack(per_frame, dim=1).double().cpu().numpy()
        thr = 0.1 * synth.DIAMETER[inp["cls"].cpu().numpy()]
        verd = np.stack([pf[:, 0] < thr, pf[:, 1] < 5.0, pf[:, 2] < 0.05], axis=1)
        parity = None
        gpath = os.path.join(ROOT, "tests", "golden", "c2_ref_verdicts.npz")
        if rank == 0 and os.path.exists(gpath):
            g = np.load(gpath)
            ra, rb = g["verdict"].astype(bool), g["verdict_b"].astype(bool)
            parity = {"frames": int(ra.shape[0]),
                      "verdict_disagreements_vs_reference_run_a": int((verd != ra).any(axis=1).sum()),
                      "verdict_disagreements_vs_reference_run_b": int((verd != rb).any(axis=1).sum()),
                      "reference_run_a_vs_run_b": int((ra != rb).any(axis=1).sum()),
                      "reference_succ_add_r_t": ra.mean(axis=0).round(4).tolist(),
                      "source": "tests/golden/c2_ref_verdicts.npz: the reference's own Trainer.get_pose + cal_dis run twice "
                                "(two randperm seeds) over these 1024 frames by tests/golden/gen_golden_c2_verdicts.py"}
        # the other BASELINE.json configs that fit one GPU, each beside a bounded run of the reference's CPU path
        # (bench_configs.py; C5 needs 25 GB of slabs and runs on its own: profiles/r02_c5_n*.json)
        other = None
        if world == 1 and not args.no_configs:
            import types
            import bench_configs
            cargs = types.SimpleNamespace(iters=5, c3_batch=4096, c4_frames=256, cpu=not args.no_cpu_baseline)
            with contextlib.redirect_stdout(io.StringIO()):
                other = {"C1": bench_configs.c1(cargs, dev), "C3": bench_configs.c3(cargs, dev), "C4": bench_configs.c4(cargs, dev)}
            other["K3_tensor_core_alternatives"] = k3_alternatives(torch, dev, inp, cfg, CAP, args.steps)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_clean / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(world),
            "hyp_pts_per_s": world * N_HYP * inp["n_px"] * args.steps / (ms_clean * 1e-3),
            "step_impl": f"EvalSweep(lanes={n_lanes}, deal='alternate'): whole batches dealt in turn to {n_lanes} streams, "
                         "joined once by finish() before the one table all-reduce; table identical to the single-stream "
                         "step: " + str(same_lanes),
            "instrumented": {"ms_per_step": ms_total / args.steps, "value": value_instr,
                             "note": "the same K steps on one stream with CUDA events between the kernels (11 per step) -- "
                                     "the pass kernel_ms / stage_ms / roofline.launch_ms are measured in"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": e2e_bytes,
                    "d2h_bytes_per_step": tb_host.numel() * tb_host.element_size(),
                    "api": "HostEvalSweep.capture_choose(...) -> replay(): run_choose captured once as a CUDA graph and replayed per "
                           "step (one cudaGraphLaunch + the table all-reduce + D2H + sync per step; the graph's memcpy nodes "
                           "and the gather kernel read whatever the pinned buffers hold).  run_choose: everything in PINNED HOST memory in the reference's own form "
                           "(tools/trainer.py:383-421) -- coordinate maps [F,3,480,640] f32, choose [F,NP] int64, "
                           "x / y_map_choosed [F,NP] f32, per-frame extent / lfborder / K / GT pose / class.  Per chunk the "
                           "lists go by the copy engine, pb_gather_choose reads the maps in place (only the chosen "
                           "pixels' 32 B sectors cross PCIe: counted exactly, distinct sectors x 3 planes), K1..K6 on two "
                           "alternating lanes, table D2H + sync every step",
                    "table_equals_device_path": same,
                    "pipelined_variant": {"value": pipe_value, "unit": UNIT, "h2d_bytes_per_step": e2e_bytes,
                                          "api": "sweep.PipelinedPasses over two captured passes (two HostEvalSweep objects): pass k+1 is "
                                                 "submitted before pass k's table is waited for, so its list copies and gather cross PCIe "
                                                 "under pass k's scoring tail; every pass's inputs are copied and every pass's table is "
                                                 "read on the host inside the timed region (the last ones by drain())"},
                    "eager_variant": {"value": eager_value, "unit": UNIT, "h2d_bytes_per_step": e2e_bytes,
                                      "api": "HostEvalSweep.run_choose called every step (no graph): ~100 launches / copies from Python"},
                    "device_map_variant": {"value": dev_value, "eager_value": dev_eager, "unit": UNIT, "h2d_bytes_per_step": dev_bytes,
                                           "api": "capture_choose / replay with pred['xyz'] still on the device (it is the "
                                                  "network's output); the int64 / f32 lists and per-frame data from the host; "
                                                  "two chunks; eager_value = run_choose called every step"},
                    "device_map_int32_variant": {"value": dev32_value, "eager_value": dev32_eager, "unit": UNIT,
                                                 "h2d_bytes_per_step": dev32_bytes,
                                                 "api": "the same with int32 choose and the pixel coordinates taken from "
                                                        "the index (no x / y lists)"},
                    "mask_zero_copy_variant": {"value": zc_value, "unit": UNIT, "h2d_bytes_per_step": zc_bytes,
                                               "api": "HostEvalSweep.run(zero_copy='xyz'): maps + u8 masks on the host (round 1's "
                                                      "headline): the 307 kB mask per frame is staged by the copy engine, the "
                                                      "mask-first K0 reads the planes in place"},
                    "staged_copy_variant": {"value": staged_value, "unit": UNIT, "h2d_bytes_per_step": staged_bytes,
                                            "api": "HostEvalSweep.run(zero_copy=False): whole maps copied H2D in "
                                                   "128-frame chunks overlapped with compute"}},
            "gpu_launches": args.steps * sw2.launches_per_step,
            "kernels_per_step": ["compact_stream_kernel", "ransac_hypothesis_kernel", "ransac_score_bal_kernel",
                                 "ransac_finalize_kernel", "pose_metrics_kernel", "accumulate_kernel"],
            "stage_ms": dict(zip(sw.STAGES, stage_ms)),
            "kernel_ms": {"compact_stream_kernel": k0_ms, "ransac_hypothesis_kernel": hyp_ms,
                          "ransac_score_bal_kernel": score_ms, "ransac_finalize_kernel": final_ms,
                          "pose_metrics_kernel": stage_ms[2], "accumulate_kernel": stage_ms[3]},
            "roofline": roof_score,                 # the dominant kernel of the step
            "roofline_all": [roof_score, roof_k0, roof_dense],
            "fp32_probe_tflops": fp32_probe,
            "cpu_baseline": cpu_base,
            "other_configs": other,
            "clocks": clocks,
            "accuracy": {"succ_add": summ["succ_base_rt"], "succ_r5deg": summ["succ_base_r"],
                         "succ_t5cm": summ["succ_base_t"], "failed": summ["failed"], "frames": summ["frames"],
                         "on_cpu_sample": accuracy_block(n_cpu, verd[:n_cpu].sum(axis=0).tolist()),
                         "parity": parity},
        }
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_JSON_FD = None


def claim_stdout():
    """stdout carries exactly ONE line, the JSON result: anything a library prints on fd 1 (NCCL's version
    banner, for one) is sent to stderr instead."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1 / C3 / C4 block (other_configs)")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours_arm(args)


if __name__ == "__main__":
    main()
