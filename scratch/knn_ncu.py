"""A few launches of the tensor-core nearest-neighbour kernel on a C3-shaped batch (for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pose_estimation_b200 import ops_metric
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
g = torch.Generator(device="cuda").manual_seed(3)
x = torch.rand(B, 2620, 3, device="cuda", generator=g) * 0.2 - 0.1
y = x + 0.002 * torch.randn(B, 2620, 3, device="cuda", generator=g)
for _ in range(3):
    idx = ops_metric.knn1(x, y, impl="tc")
torch.cuda.synchronize()
