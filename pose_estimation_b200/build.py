"""Build libposeb200.so (sm_100a only) in-tree with nvcc.

``-fmad=false``: every fused multiply-add in the bit-defined kernels is an explicit ``fmaf``; the
compiler must not contract the remaining ``a*b+c`` (the CPU oracle is built -ffp-contract=off).
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libposeb200.so")
SOURCES = ["abi.cu", "cloud.cu", "compact.cu", "ransac.cu", "ransac_score.cu", "ransac_score_tc.cu", "ransac_score_mma.cu", "knn.cu", "knn_tc.cu", "metrics.cu", "sim3.cu", "fps.cu", "bpnp.cu", "possloss.cu", "region.cu", "synthgen.cu", "probe.cu"]
HEADERS = ["common.cuh", "p3p.cuh", "nn.cuh", "ransac_plan.h", "tc.cuh", "score_tc_common.cuh", os.path.join("..", "..", "include", "poseb200.h")]
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build() -> bool:
    if not os.path.exists(SO):
        return True
    so_t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    deps += [os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.exists(d) and os.path.getmtime(d) > so_t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return SO
    nvcc = _nvcc()
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        if not os.path.exists(src):
            continue
        obj = os.path.join(HERE, "build", s.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, *os.environ.get("PB_EXTRA_NVCC", "").split(), "-c", src, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write(f"--- nvcc {s} failed ---\n{out}\n")
        elif verbose or out.strip():
            sys.stderr.write(f"--- nvcc {s} ---\n{out}\n")
    if failed:
        raise RuntimeError("nvcc failed")
    tmp = SO + ".tmp"
    subprocess.check_call([nvcc, "-shared", "-o", tmp, *objs, "-lcudart", "-lcuda"])
    os.replace(tmp, SO)
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
