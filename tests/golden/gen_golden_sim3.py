"""Generate tests/golden/sim3_ref.npz by RUNNING THE REFERENCE: imports /root/reference/lib/transform/umeyama.py
(pure numpy) in this container, feeds it seeded synthetic correspondences, records the index sets its
np.random.randint calls produced, the per-iteration inlier counts (recomputed with the reference's own
estimateSimilarityUmeyama) and the returned transform.  Run here only; the GPU box has no /root/reference.

    python tests/golden/gen_golden_sim3.py
"""
import importlib.util
import os
import sys

import numpy as np

REF = "/root/reference/lib/transform/umeyama.py"
spec = importlib.util.spec_from_file_location("ref_umeyama", REF)
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)


def rot(rng):
    q = rng.normal(size=4); q /= np.linalg.norm(q)
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def make_case(rng, n, outlier_frac, noise, scale):
    src = (rng.random((n, 3)) - 0.5) * np.array([0.10, 0.08, 0.14])
    R = rot(rng)
    t = rng.normal(size=3) * 0.1 + np.array([0.0, 0.0, 0.8])
    dst = scale * src @ R.T + t + rng.normal(size=(n, 3)) * noise
    out = rng.random(n) < outlier_frac
    dst[out] = t + (rng.random((int(out.sum()), 3)) - 0.5) * 0.3
    return src.astype(np.float32), dst.astype(np.float32), R, t


def run_reference(src, dst, seed):
    """reference estimateSimilarityTransform with np.random seeded; logs every randint draw."""
    log = []
    orig = np.random.randint
    def spy(*a, **k):
        r = orig(*a, **k)
        log.append(np.array(r, dtype=np.int32))
        return r
    np.random.seed(seed)
    np.random.randint = spy
    try:
        res = ref.estimateSimilarityTransform(src.astype(np.float64), dst.astype(np.float64))
    finally:
        np.random.randint = orig
    idx = np.stack(log)                                     # [iterations actually run, 5]
    n = src.shape[0]
    SH = np.vstack([src.T.astype(np.float64), np.ones(n)]); TH = np.vstack([dst.T.astype(np.float64), np.ones(n)])
    cen = SH[:3].mean(axis=1)
    inlier_t = 2 * np.amax(np.linalg.norm(SH[:3] - cen[:, None], axis=0)) / 10.0
    counts = []
    for r in idx:                                           # per-iteration inlier counts, the reference's own math
        sc, _, _, M = ref.estimateSimilarityUmeyama(SH[:, r], TH[:, r])
        res_v = np.linalg.norm((TH - M @ SH)[:3], axis=0)
        counts.append(int((res_v < sc * inlier_t).sum()))
    return idx, np.array(counts, np.int32), res


def main():
    rng = np.random.default_rng(20240507)
    out = {}
    cases = [(600, 0.30, 5e-4, 1.0), (1000, 0.50, 1e-3, 1.0), (257, 0.10, 2e-4, 0.5), (2000, 0.65, 1e-3, 2.0), (64, 0.0, 0.0, 1.0)]
    for k, (n, of, noise, sc) in enumerate(cases):
        src, dst, R, t = make_case(rng, n, of, noise, sc)
        idx, counts, res = run_reference(src, dst, 1000 + k)
        S, Rr, Tr, M = res
        out[f"src{k}"], out[f"dst{k}"] = src, dst
        out[f"idx{k}"], out[f"counts{k}"] = idx, counts
        out[f"scale{k}"] = np.float64(S); out[f"R{k}"] = Rr; out[f"t{k}"] = Tr
        print(f"case {k}: n={n} iterations run={idx.shape[0]} best count={counts.max()} scale={S:.6f}")
    # direct Umeyama fits (no RANSAC): reference estimateSimilarityUmeyama on clean / noisy / reflected inputs
    for k in range(4):
        n = 50 + 37 * k
        src, dst, R, t = make_case(rng, n, 0.0, 1e-3 * k, 0.7 + 0.3 * k)
        if k == 3:
            dst = dst * np.array([1, 1, -1], np.float32)           # a reflection: the det correction must fire
        SH = np.vstack([src.T.astype(np.float64), np.ones(n)]); TH = np.vstack([dst.T.astype(np.float64), np.ones(n)])
        S, Rr, Tr, M = ref.estimateSimilarityUmeyama(SH, TH)
        out[f"u_src{k}"], out[f"u_dst{k}"] = src, dst
        out[f"u_scale{k}"] = np.float64(S); out[f"u_R{k}"] = Rr; out[f"u_t{k}"] = Tr
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "sim3_ref.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
