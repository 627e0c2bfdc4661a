"""Stage times of the RANSAC call with a library variant (scratch/_so/*.so built with other PB_BAL_WAVES / experiment
macros).  Usage: python scratch/k3_waves.py scratch/_so/w2.so"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from pose_estimation_b200 import _lib
_lib.SO_PATH = os.path.abspath(sys.argv[1])
import bench
from pose_estimation_b200 import ops


def stage_ms(fn, reps=30):
    lib = _lib.load()
    fn(); torch.cuda.synchronize()
    lib.pb_stage_timing(1)
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    ms = (C.c_float * 3)(); calls = C.c_int()
    lib.pb_stage_timing_read(ms, C.byref(calls))
    lib.pb_stage_timing(0)
    return [round(float(x), 5) for x in ms]


def main():
    F = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    inp = bench.build_inputs(torch, dev, 0, F)
    cap = 1280
    pts, count = ops.compact_map(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], cap)
    cfg = ops.RansacConfig()
    pr = ops.PtsRansac(F, cap, cfg, device=dev)
    ops.set_score_impl("fp32")
    ref = ops.pnp_ransac_pts(pts, count, inp["K4"], cfg, want_counts=True)
    res = {"so": sys.argv[1], "F": F, "counts_sum": int(ref["hyp_counts"].clamp(min=0).sum()),
           "stage_ms": [stage_ms(lambda: pr(pts, count, inp["K4"])) for _ in range(3)]}
    lib = _lib.load()
    if hasattr(lib, "pb_debug_k3_timeline"):
        n = 4096
        buf = (C.c_ulonglong * (4 * n))()
        lib.pb_debug_k3_timeline(buf, n)
        a = np.frombuffer(buf, dtype=np.uint64).reshape(n, 4).astype(np.int64)
        a = a[(a[:, 0] > 0) & (a[:, 1] > 0)]
        t0 = a[:, 0].min()
        en = a[:, 1] - t0
        res["ctas"] = int(len(a))
        res["end_ns"] = [int(np.percentile(en, q)) for q in (0, 10, 25, 50, 75, 90, 100)]
    print(json.dumps(res))


if __name__ == "__main__":
    main()
