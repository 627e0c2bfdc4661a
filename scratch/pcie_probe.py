"""Developer probe: where the end-to-end time of the zero-copy path goes (C2, 1024 frames in pinned host memory).
(a) DMA of the mask alone, (b) K0 reading everything in place, (c) K0 with the mask on the device and the planes in
place, (d) (a) and (c) of different chunks concurrently."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from pose_estimation_b200 import ops, synth
dev = torch.device("cuda:0")
F = bench.FRAMES_PER_GPU
inp = bench.build_inputs(torch, dev, 0, F)
H, W = synth.IMG_H, synth.IMG_W
xyz_h = torch.empty((F, 3, H, W), dtype=torch.float32).pin_memory(); xyz_h.copy_(inp["xyz"])
mask_h = torch.empty((F, H, W), dtype=torch.uint8).pin_memory(); mask_h.copy_(inp["mask"])
mask_d = torch.empty((F, H, W), dtype=torch.uint8, device=dev)
torch.cuda.synchronize()

def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n

ms = timed(lambda: mask_d.copy_(mask_h, non_blocking=True))
print(f"(a) mask DMA {mask_h.numel()/1e6:.1f} MB: {ms:.3f} ms = {mask_h.numel()/ms/1e6:.1f} GB/s")
comp = ops.MapCompactor(F, H, W, bench.CAP, device=dev)
ms = timed(lambda: comp(xyz_h, mask_h, inp["extent"], inp["lfborder"]))
print(f"(b) K0 all in place: {ms:.3f} ms")
ms = timed(lambda: comp(xyz_h, mask_d, inp["extent"], inp["lfborder"]))
print(f"(c) K0 mask on device, planes in place: {ms:.3f} ms   (n_px {inp['n_px']}, 96 B/px = {inp['n_px']*96/1e6:.1f} MB)")
ms = timed(lambda: comp(inp["xyz"], mask_d, inp["extent"], inp["lfborder"]))
print(f"    K0 all on device: {ms:.3f} ms")
for nchunk in (2, 4, 8, 16):
    c = F // nchunk
    comps = [ops.MapCompactor(c, H, W, bench.CAP, device=dev) for _ in range(2)]
    cs = torch.cuda.Stream()
    evs = [torch.cuda.Event() for _ in range(nchunk)]
    def piped():
        main = torch.cuda.current_stream()
        cs.wait_stream(main)
        with torch.cuda.stream(cs):
            for i in range(nchunk):
                mask_d[i*c:(i+1)*c].copy_(mask_h[i*c:(i+1)*c], non_blocking=True)
                evs[i].record(cs)
        for i in range(nchunk):
            main.wait_event(evs[i])
            comps[i & 1](xyz_h[i*c:(i+1)*c], mask_d[i*c:(i+1)*c], inp["extent"][i*c:(i+1)*c], inp["lfborder"][i*c:(i+1)*c])
    ms = timed(piped)
    print(f"(d) {nchunk} chunks, mask DMA || K0 gather: {ms:.3f} ms")
