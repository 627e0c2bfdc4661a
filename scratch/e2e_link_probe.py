"""What the PCIe link gives the end-to-end pass (maps + lists in pinned host memory): the list copies alone, the zero-copy
gather alone (one launch / four chunks / two streams), both together, against the whole graph pass.
Usage: python scratch/e2e_link_probe.py [out.json]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
from pose_estimation_b200 import ops, synth


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else None
    F, cap = 1024, 1280
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    H, W = synth.IMG_H, synth.IMG_W
    inp = bench.build_inputs(torch, dev, 0, F)
    xyz_h = torch.empty((F, 3, H, W), dtype=torch.float32).pin_memory()
    xyz_h.copy_(inp["xyz"])
    NPL = (max(inp["counts"]) + 15) // 16 * 16
    choose_np = np.zeros((F, NPL), np.int64)
    for f, (r_, c_) in enumerate(inp["rc"]):
        choose_np[f, :r_.shape[0]] = r_.astype(np.int64) * W + c_
    xm_np = (choose_np % W).astype(np.float32)
    ym_np = (choose_np // W).astype(np.float32)
    nv_np = np.asarray(inp["counts"], np.int32)
    sectors = sum(np.unique(choose_np[f, :nv_np[f]] >> 3).shape[0] for f in range(F))
    choose_h, xm_h, ym_h, nv_h = (torch.from_numpy(a_).pin_memory() for a_ in (choose_np, xm_np, ym_np, nv_np))
    choose_d, xm_d, ym_d, nv_d = (torch.empty_like(t, device=dev) for t in (choose_h, xm_h, ym_h, nv_h))
    list_mb = sum(t.numel() * t.element_size() for t in (choose_h, xm_h, ym_h, nv_h)) / 1e6
    gat_mb = sectors * 3 * 32 / 1e6
    res = {"F": F, "NPL": int(NPL), "list_MB": list_mb, "gather_MB": gat_mb}

    def copy_lists(sl=slice(None)):
        choose_d[sl].copy_(choose_h[sl], non_blocking=True)
        xm_d[sl].copy_(xm_h[sl], non_blocking=True)
        ym_d[sl].copy_(ym_h[sl], non_blocking=True)
        nv_d[sl].copy_(nv_h[sl], non_blocking=True)

    ms = timed(copy_lists)
    res["lists_alone"] = {"ms": ms, "GBps": list_mb / ms}
    copy_lists(); torch.cuda.synchronize()

    g_all = ops.ChooseGatherer(F, H, W, cap, device=dev)
    ms = timed(lambda: g_all(xyz_h, choose_d, inp["extent"], inp["lfborder"], xm_d, ym_d, nv_d))
    res["gather_1x1024"] = {"ms": ms, "GBps": gat_mb / ms}
    # device-resident maps for comparison
    ms = timed(lambda: g_all(inp["xyz"], choose_d, inp["extent"], inp["lfborder"], xm_d, ym_d, nv_d))
    res["gather_1x1024_device_maps"] = {"ms": ms}

    for nchunk in (2, 4, 8):
        c = F // nchunk
        gs = [ops.ChooseGatherer(c, H, W, cap, device=dev) for _ in range(2)]
        def seq():
            for i in range(nchunk):
                sl = slice(i * c, (i + 1) * c)
                gs[0](xyz_h[sl], choose_d[sl], inp["extent"][sl], inp["lfborder"][sl], xm_d[sl], ym_d[sl], nv_d[sl])
        ms = timed(seq)
        res[f"gather_{nchunk}_chunks_one_stream"] = {"ms": ms, "GBps": gat_mb / ms}
        s = [torch.cuda.Stream(dev), torch.cuda.Stream(dev)]
        ev = torch.cuda.Event()
        def two():
            ev.record()
            for k in range(2):
                s[k].wait_event(ev)
            for i in range(nchunk):
                sl = slice(i * c, (i + 1) * c)
                with torch.cuda.stream(s[i & 1]):
                    gs[i & 1](xyz_h[sl], choose_d[sl], inp["extent"][sl], inp["lfborder"][sl], xm_d[sl], ym_d[sl], nv_d[sl])
            for k in range(2):
                torch.cuda.current_stream().wait_stream(s[k])
        ms = timed(two)
        res[f"gather_{nchunk}_chunks_two_streams"] = {"ms": ms, "GBps": gat_mb / ms}

    # lists and gather together (gather of the whole batch on one stream, copies on another)
    cs = torch.cuda.Stream(dev)
    ev2 = torch.cuda.Event()
    def both():
        ev2.record()
        cs.wait_event(ev2)
        with torch.cuda.stream(cs):
            copy_lists()
        g_all(xyz_h, choose_d, inp["extent"], inp["lfborder"], xm_d, ym_d, nv_d)
        torch.cuda.current_stream().wait_stream(cs)
    ms = timed(both)
    res["lists_and_gather_concurrently"] = {"ms": ms, "GBps": (gat_mb + list_mb) / ms}
    print(json.dumps(res))
    if out_path:
        with open(out_path, "w") as fh:
            json.dump(res, fh, indent=1)


if __name__ == "__main__":
    main()
