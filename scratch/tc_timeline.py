"""Developer probe (needs the debug build of scratch/tc_debug.cu): per-CTA timeline of knn_tc_kernel."""
import ctypes as C, numpy as np, torch
from pose_estimation_b200 import ops_metric, _lib
B, N = 256, 2620
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.rand(B, N, 3, device="cuda", generator=g); y = torch.rand(B, N, 3, device="cuda", generator=g)
for _ in range(2):
    ops_metric.knn1(x, y, impl="tc")
torch.cuda.synchronize()
lib = _lib.load() if hasattr(_lib, "load") else _lib.LIB
buf = (C.c_ulonglong * (256 * 8))()
lib.pb_tc_debug_read.argtypes = [C.c_void_p]; lib.pb_tc_debug_read.restype = C.c_int
print("rc", lib.pb_tc_debug_read(buf))
t = np.frombuffer(buf, dtype=np.uint64).reshape(256, 8).astype(np.int64)
t0 = t[:, 0].min()
names = ["entry", "alloc+sync", "A loaded", "mma0 issued", "last mma issued", "epi tile1", "epi loop end", "exit"]
d = t - t[:, :1]
print("median ns since CTA entry:")
for i, n in enumerate(names): print(f"  {n:16s} {np.median(d[:, i]):9.0f}   p10 {np.percentile(d[:, i],10):9.0f} p90 {np.percentile(d[:, i],90):9.0f}")
print("kernel span of sampled CTAs (us)", (t[:, 7].max() - t0) / 1e3)
o = np.argsort(t[:, 0])
for k in o[:6].tolist() + o[-3:].tolist(): print(k, (t[k] - t0).tolist())

if hasattr(lib, "pb_tc_debug_read_cycles"):
    lib.pb_tc_debug_read_cycles.argtypes = [C.c_void_p]; lib.pb_tc_debug_read_cycles.restype = C.c_int
    lib.pb_tc_debug_read_cycles(buf)
    c = np.frombuffer(buf, dtype=np.uint64).reshape(256, 8).astype(np.int64)
    print("warp 0 cycles over the tile loop [mbar_wait, fence_after, tmem ld+wait+min/push, fence+arrive]:", np.median(c[:, :4], axis=0).tolist())
    print("issuer cycles [wait load, wait free, mma+commit issue, first commit->barrier latency]:", np.median(c[:, 4:8], axis=0).tolist())
