"""Developer probe: EvalSweep step time vs number of stream lanes (C2, 1024 frames)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import pose_estimation_b200 as pb
from pose_estimation_b200 import ops, synth
dev = torch.device("cuda:0")
F = bench.FRAMES_PER_GPU
inp = bench.build_inputs(torch, dev, 0, F)
models = torch.from_numpy(synth.class_models(bench.N_MODEL))
diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
cfg = ops.RansacConfig(n_hyp=bench.N_HYP, seed=bench.SEED)
ref = None
for lanes in (1, 2, 4, 8):
    sw = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, bench.CAP, cfg, device=dev, lanes=lanes)
    def step():
        sw.reset()
        sw.step(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], inp["K4"], inp["cls"], inp["R_gt"], inp["t_gt"])
    for _ in range(5): step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50): step()
    b.record(); torch.cuda.synchronize()
    t = sw.table.clone()
    if ref is None: ref = t
    print(f"lanes {lanes}: {a.elapsed_time(b) / 50:.4f} ms/step  table equal to 1-lane: {bool(torch.equal(t, ref))}")

print("--- CUDA graph replay")
for lanes in (1, 2, 4):
    sw = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, bench.CAP, cfg, device=dev, lanes=lanes)
    def step():
        sw.reset()
        sw.step(inp["xyz"], inp["mask"], inp["extent"], inp["lfborder"], inp["K4"], inp["cls"], inp["R_gt"], inp["t_gt"])
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3): step()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g, stream=s):
            step()
    except Exception as e:
        print("capture failed:", repr(e)[:300]); continue
    for _ in range(5): g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50): g.replay()
    b.record(); torch.cuda.synchronize()
    t = sw.table.clone()
    print(f"graph lanes {lanes}: {a.elapsed_time(b) / 50:.4f} ms/step  table equal: {bool(torch.equal(t, ref))}")
