"""Generate tests/golden/fps_ref.npz by RUNNING THE REFERENCE: imports farthest_point_sampling from
/root/reference/tools/script/sample_model.py (numpy) on seeded clouds.  Run here only.

    python tests/golden/gen_golden_fps.py
"""
import importlib.util
import os

import numpy as np

spec = importlib.util.spec_from_file_location("ref_sample_model", "/root/reference/tools/script/sample_model.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

rng = np.random.default_rng(608)
out = {}
for k, (n, m) in enumerate(((1500, 300), (777, 777), (64, 100), (2620, 64), (1, 4))):
    pts = (rng.random((n, 3)) * np.array([0.10, 0.08, 0.12])).astype(np.float32)
    if k == 1:
        pts[100:110] = pts[50:60]                       # exact duplicates: np.argmax's first-maximum rule decides
    out[f"pts{k}"] = pts
    out[f"idx{k}"] = ref.farthest_point_sampling(pts, m).astype(np.int32)
    print(k, n, m, out[f"idx{k}"][:6])
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fps_ref.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path))
