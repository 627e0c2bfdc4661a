"""Developer probe (needs the debug build of scratch/tc_debug.cu): tcgen05.mma issue/execution rate per operand layout."""
import ctypes as C, torch
from pose_estimation_b200 import _lib
torch.zeros(1, device="cuda")
lib = _lib.load()
f = lib.pb_tc_probe_mma
f.argtypes = [C.c_int, C.c_int, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_void_p]; f.restype = C.c_int
out = (C.c_longlong * 2)()
reps = 512
cases = [("none  lbo2048 sbo128 (current)", 0, 2048, 128, 4096),
         ("none  lbo128 sbo256 (k chunks adjacent)", 0, 128, 256, 256),
         ("sw32  sbo256", 6, 16, 256, 32),
         ("sw64  sbo512", 4, 16, 512, 32),
         ("sw128 sbo1024", 2, 16, 1024, 32)]
for ctas in (1, 148):
    for n in (64, 128, 256):
        for nacc in (1, 2):
            if n * nacc > 512: continue
            for name, layout, lbo, sbo, kstep in cases:
                rc = f(ctas, n, layout, lbo, sbo, kstep, reps, nacc, out)
                print(f"ctas {ctas:3d} N {n:3d} nacc {nacc} {name:42s} rc {rc} issue {out[0] / (2 * reps):7.1f} cyc/mma, complete {out[1] / (2 * reps):7.1f} cyc/mma")
