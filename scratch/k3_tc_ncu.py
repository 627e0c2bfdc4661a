"""One launch of the tensor-core scoring kernel on a C2-shaped batch (for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from pose_estimation_b200 import ops
F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dev = torch.device("cuda", 0)
inp = bench.build_inputs(torch, dev, 0, F)
pts, count = ops.compact_map(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], 1280)
pr = ops.PtsRansac(F, 1280, ops.RansacConfig(), device=dev)
ops.set_score_impl(sys.argv[2] if len(sys.argv) > 2 else "tc", float(sys.argv[3]) if len(sys.argv) > 3 else 1.0)
for _ in range(3):
    pr(pts, count, inp["K4"])
torch.cuda.synchronize()
