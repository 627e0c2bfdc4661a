"""Developer probe: where the C5 sweep's time goes (generator write / erase alone, evaluation alone, pipeline)."""
import os, sys, torch, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import pose_estimation_b200 as pb
from pose_estimation_b200 import ops, synth
dev = torch.device("cuda:0")
F, H, W = 2048, synth.IMG_H, synth.IMG_W
P = synth.c5_params(bench.SEED, 0, 4 * F)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
prm, cls, ext, lfb, K4, Rg, tg = (t(P[k]) for k in ("params", "cls", "extent", "lfborder", "K4", "R_gt", "t_gt"))
bb = P["bbox_max"]
print("bbox_max", bb, "mean bbox", float(((P["params"][:, 33] - P["params"][:, 32]) * (P["params"][:, 35] - P["params"][:, 34])).mean()))
gen = ops.FrameGenerator(F, H, W, bench.SEED, device=dev)

def timed(fn, n=10):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n

def wr():
    gen.erase(); gen.write(prm[:F], 0, bb)
print("erase + write per chunk: %.3f ms" % timed(wr))
gen.erase(); gen.write(prm[:F], 0, bb)
def er():
    gen._launch(True)
print("erase alone: %.3f ms" % timed(er))
gen.params = prm[:F]; gen._launch(False)
models = torch.from_numpy(synth.class_models(bench.N_MODEL)); diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
cfg = ops.RansacConfig(n_hyp=bench.N_HYP, seed=bench.SEED)
for lanes in (1, 3):
    sw = pb.EvalSweep(models, diam, sym, F, H, W, bench.CAP, cfg, device=dev, **({} if lanes == 1 else dict(lanes=lanes, deal="alternate")))
    def ev():
        sw.step(gen.xyz, gen.mask, ext[:F], lfb[:F], K4[:F], cls[:F], Rg[:F], tg[:F])
    ms = timed(ev, 12)
    sw.finish(); torch.cuda.synchronize()
    print("eval of a generated chunk, lanes=%d: %.3f ms per chunk" % (lanes, ms))
print("foreground px per frame: %.1f" % float(gen.mask.sum().item() / F))
