"""Developer probe: two EvalSweeps of the FULL batch on two streams, batches dealt alternately (no split)."""
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
for L in (2, 3):
    sws = [pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, bench.CAP, cfg, device=dev) for _ in range(L)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(L)]
    start, done = torch.cuda.Event(), [torch.cuda.Event() for _ in range(L)]

    def sweep(k_steps):
        main = torch.cuda.current_stream(dev)
        start.record(main)
        for l in range(L):
            streams[l].wait_event(start)
        for k in range(k_steps):
            with torch.cuda.stream(streams[k % L]):
                sws[k % L].step(*(inp[key] for key in KEYS), frame0=0)
        for l in range(L):
            done[l].record(streams[l])
            main.wait_event(done[l])
    sweep(4); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); sweep(K); b.record(); torch.cuda.synchronize()
    print(f"alternating full batches over {L} streams: {a.elapsed_time(b) / K:.4f} ms per 1024 frames")
