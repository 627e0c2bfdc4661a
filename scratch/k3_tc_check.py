"""Tensor-core scoring kernel (csrc/ransac_score_tc.cu) against the FP32-pipe kernels on a C2-shaped batch:
bit-equality of the inlier counts / chosen hypothesis / pose, the re-evaluation statistics, the calibration pass
and the time of the scoring stage.  Usage: python scratch/k3_tc_check.py [F] [out.json]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
from pose_estimation_b200 import _lib, ops


def stage_ms(fn, reps=20):
    lib = _lib.load()
    fn(); torch.cuda.synchronize()
    lib.pb_stage_timing(1)
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    import ctypes as C
    ms = (C.c_float * 3)(); calls = C.c_int()
    lib.pb_stage_timing_read(ms, C.byref(calls))
    lib.pb_stage_timing(0)
    return [round(float(x), 5) for x in ms]


def main():
    F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    out_path = sys.argv[2] if len(sys.argv) > 2 else None
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    inp = bench.build_inputs(torch, dev, 0, F)
    cap = 1280
    pts, count = ops.compact_map(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], cap)
    cfg = ops.RansacConfig()
    res = {"F": F, "cap": cap, "n_hyp": cfg.n_hyp, "thresh_px": cfg.thresh_px, "points": int(count.sum())}
    outs = {}
    TC = os.environ.get("K3_IMPL", "tc")                      # "tc" (tcgen05) or "mma" (mma.sync)
    for impl in ("fp32", TC):
        ops.set_score_impl(impl)
        o = ops.pnp_ransac_pts(pts, count, inp["K4"], cfg, want_counts=True)
        torch.cuda.synchronize()
        outs[impl] = {k: v.clone() for k, v in o.items()}
    a, b = outs["fp32"], outs[TC]
    res["impl"] = TC
    res["counts_equal"] = bool(torch.equal(a["hyp_counts"], b["hyp_counts"]))
    res["counts_diff"] = int((a["hyp_counts"] != b["hyp_counts"]).sum())
    if res["counts_diff"]:
        d = (a["hyp_counts"] != b["hyp_counts"]).nonzero()[:8].cpu().numpy().tolist()
        res["first_diffs"] = [(f, h, int(a["hyp_counts"][f, h]), int(b["hyp_counts"][f, h])) for f, h in d]
    res["best_equal"] = bool(torch.equal(a["best_hyp"], b["best_hyp"]) and torch.equal(a["n_inl"], b["n_inl"]))
    res["pose_equal"] = bool(torch.equal(a["R"], b["R"]) and torch.equal(a["t"], b["t"]))
    # persistent launcher: stats and timing
    pr = ops.PtsRansac(F, cap, cfg, device=dev)
    for impl in ("fp32", TC):
        ops.set_score_impl(impl)
        res[f"stage_ms_{impl}"] = stage_ms(lambda: pr(pts, count, inp["K4"]))
        if impl == TC:
            res["tc_stats"] = ops.ransac_tc_stats(pr.ws, F, cap, cfg)
    for kappa in (0.5, 0.25, 0.0625):
        ops.set_score_impl(TC, kappa)
        o = ops.pnp_ransac_pts(pts, count, inp["K4"], cfg, want_counts=True)
        torch.cuda.synchronize()
        res[f"kappa_{kappa}"] = dict(counts_diff=int((a["hyp_counts"] != o["hyp_counts"]).sum()),
                                     stage_ms=stage_ms(lambda: pr(pts, count, inp["K4"])),
                                     stats=ops.ransac_tc_stats(pr.ws, F, cap, cfg))
    ops.set_score_impl(TC + "_calib", 1.0)
    pr(pts, count, inp["K4"])
    res["calib"] = ops.ransac_tc_stats(pr.ws, F, cap, cfg)
    if TC == "tc":
        ops.set_score_impl("tc_prof", 1.0)
        pr(pts, count, inp["K4"])
        res["prof"] = ops.ransac_tc_stats(pr.ws, F, cap, cfg)["clocks"]
    ops.set_score_impl("auto", 1.0)
    print(json.dumps(res))
    if out_path:
        with open(out_path, "w") as fh:
            json.dump(res, fh, indent=1)


if __name__ == "__main__":
    main()
