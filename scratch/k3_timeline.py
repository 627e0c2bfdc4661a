"""Per-CTA timeline of the balanced scoring kernel (library built with PB_EXTRA_NVCC=-DPB_K3_TIMELINE=1):
start / end (globaltimer, ns), SM id, cycles of every CTA of the last launch.  Usage: python scratch/k3_timeline.py [out.json]"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
from pose_estimation_b200 import _lib, ops


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else None
    F = 1024
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    inp = bench.build_inputs(torch, dev, 0, F)
    cap = 1280
    pts, count = ops.compact_map(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], cap)
    cfg = ops.RansacConfig()
    pr = ops.PtsRansac(F, cap, cfg, device=dev)
    ops.set_score_impl("fp32")
    for _ in range(5):
        pr(pts, count, inp["K4"])
    torch.cuda.synchronize()
    lib = _lib.load()
    n = 888
    buf = (C.c_ulonglong * (4 * n))()
    rc = lib.pb_debug_k3_timeline(buf, n)
    assert rc == 0
    a = np.frombuffer(buf, dtype=np.uint64).reshape(n, 4).astype(np.int64)
    t0 = a[:, 0].min()
    st, en, sm, cyc = a[:, 0] - t0, a[:, 1] - t0, a[:, 2], a[:, 3]
    live = en > 0
    res = {"ctas": int(n), "live": int(live.sum()),
           "start_ns": {"min": int(st.min()), "p50": int(np.median(st)), "max": int(st.max())},
           "end_ns": {"min": int(en[live].min()), "p10": int(np.percentile(en[live], 10)), "p50": int(np.median(en[live])),
                      "p90": int(np.percentile(en[live], 90)), "max": int(en[live].max())},
           "dur_ns": {"min": int((en - st)[live].min()), "p50": int(np.median((en - st)[live])), "max": int((en - st)[live].max())},
           "cycles": {"min": int(cyc[live].min()), "p50": int(np.median(cyc[live])), "max": int(cyc[live].max())},
           "ctas_per_sm": np.bincount(sm, minlength=148).tolist()}
    # per SM: when its last CTA ends
    last = np.zeros(148, np.int64); first_end = np.full(148, 1 << 60, np.int64)
    for i in range(n):
        if live[i]:
            last[sm[i]] = max(last[sm[i]], en[i]); first_end[sm[i]] = min(first_end[sm[i]], en[i])
    res["sm_last_end_ns"] = {"min": int(last.min()), "p50": int(np.median(last)), "max": int(last.max())}
    res["sm_first_end_ns"] = {"min": int(first_end.min()), "p50": int(np.median(first_end)), "max": int(first_end.max())}
    res["points"] = int(count.sum())
    res["hist_end_ns"] = np.histogram(en[live], bins=12)[0].tolist()
    res["hist_end_edges"] = [int(x) for x in np.histogram(en[live], bins=12)[1]]
    print(json.dumps(res))
    if out_path:
        with open(out_path, "w") as fh:
            json.dump(res, fh, indent=1)


if __name__ == "__main__":
    main()
