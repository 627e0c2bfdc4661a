"""Developer probe: free-running lanes -- L independent EvalSweeps of F/L frames, each on its own stream, stepping K
times without a join between steps (joined once at the end), plain launches and as one captured graph."""
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
KEYS = ("xyz", "mask", "extent", "lfborder", "K4", "cls", "R_gt", "t_gt")
K = 20
for L in (2, 3, 4):
    if F % L and L != 3:
        continue
    Fl = F // L if L != 3 else 341
    sws = [pb.EvalSweep(models, diam, sym, Fl, synth.IMG_H, synth.IMG_W, bench.CAP, cfg, device=dev) for _ in range(L)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(L)]
    start, done = torch.cuda.Event(), [torch.cuda.Event() for _ in range(L)]

    def sweep(k_steps):
        main = torch.cuda.current_stream(dev)
        start.record(main)
        for l in range(L):
            streams[l].wait_event(start)
        for _ in range(k_steps):
            for l in range(L):
                lo = l * Fl
                with torch.cuda.stream(streams[l]):
                    sws[l].step(*(inp[k][lo:lo + Fl] for k in KEYS), frame0=lo)
        for l in range(L):
            done[l].record(streams[l])
            main.wait_event(done[l])
    sweep(3); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); sweep(K); b.record(); torch.cuda.synchronize()
    print(f"free lanes {L} (x{Fl} frames): {a.elapsed_time(b) / K * F / (L * Fl):.4f} ms per 1024 frames (plain launches)")
    side = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(side):
        sweep(2)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        sweep(K // 2)
    g.replay(); torch.cuda.synchronize()
    a.record(); g.replay(); g.replay(); b.record(); torch.cuda.synchronize()
    print(f"free lanes {L}: {a.elapsed_time(b) / K * F / (L * Fl):.4f} ms per 1024 frames (graph of {K // 2} steps, replayed twice)")
