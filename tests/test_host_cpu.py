"""CPU tests: the C-ABI library loads and exports every declared symbol (no compute calls), host-side
logic (sharding, table handling, AUC, intrinsics), and the world_size-2 gloo path."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

from pose_estimation_b200 import _lib, api, ops_metric, sweep, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "poseb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(pb_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    from pose_estimation_b200 import build
    build.build()
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name)
    assert lib.pb_abi_version() == 2            # 2: pb_sim3_params gained `rule`
    # size queries are host-only and safe without a GPU
    import ctypes as C
    prm = _lib.RansacParams(256, 1.0, 6, 0, 5, 1, 0)
    assert lib.pb_ransac_workspace_bytes(1024, 1280, C.byref(prm)) >= 1024 * 256 * 48
    assert lib.pb_compact_workspace_bytes(1024, 480, 640, 1280) == 256 + 1024 * 5 * 4   # ticket + one state word per 64 KB tile
    assert lib.pb_compact_workspace_bytes(4, 480, 640, 1280) > 0
    assert lib.pb_pnp_ransac_map_workspace_bytes(8, 480, 640, 1280, C.byref(prm)) > 8 * 5 * 1280 * 4
    # argument errors come back as codes + message, never as exceptions/crashes
    assert lib.pb_knn1(None, None, 1, 4, 0, 0, None, None, None) == -1
    assert b"pb_knn1" in lib.pb_last_error()


def test_kernel_switches_validate_their_arguments():
    """The opt-in kernel switches (pb_ransac_score_impl, pb_knn_tc_impl) touch no device: out-of-range selectors are
    refused with a message, valid ones are accepted and the default is restored."""
    from pose_estimation_b200 import ops
    lib = _lib.load()
    assert lib.pb_ransac_score_impl(7, 1.0) == -1 and b"pb_ransac_score_impl" in lib.pb_last_error()
    assert lib.pb_ransac_score_impl(-1, 1.0) == -1
    assert lib.pb_ransac_score_impl(2, 0.0) == -1 and b"band_scale" in lib.pb_last_error()
    assert lib.pb_knn_tc_impl(3) == -1 and b"pb_knn_tc_impl" in lib.pb_last_error()
    with pytest.raises(KeyError):
        ops.set_score_impl("no-such-kernel")
    for name in ("fp32", "tc", "tc_calib", "mma", "mma_calib", "auto"):
        ops.set_score_impl(name, 0.5)
    ops.set_score_impl("auto", 1.0)
    for name in ("mma", "tcgen05", "auto"):
        ops_metric.set_knn_tc_impl(name)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "SO_PATH", "/nonexistent/libposeb200.so")
    with pytest.raises(_lib.PoseB200Error, match="no CPU"):
        _lib.load()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "pose_estimation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("CPU oracle", "").replace("the oracle", ""), f


def test_shard_range_partitions():
    for n in (0, 1, 7, 1024, 65536, 65537):
        for w in (1, 2, 3, 4, 8):
            spans = [sweep.shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_intrinsics_auc_and_table_helpers():
    K4 = torch.tensor([[572.4114, 573.57043, 325.2611, 242.04899], [700., 700., 320., 240.]])
    K33 = api.batch_intrinsic_transform(K4)
    assert K33.shape == (2, 3, 3) and K33[0, 0, 2] == K4[0, 2] and K33[1, 2, 2] == 1
    assert torch.equal(api.batch_intrinsic_transform(K33), K4)
    with pytest.raises(Exception, match="Wrong Shape"):
        api.batch_intrinsic_transform(torch.zeros(3))
    from oracle import ref_callsites as rc
    rng = np.random.default_rng(1)
    for n in (1, 2, 50):
        d = list(rng.random(n) * 0.2)
        assert abs(api.Metric([]).cal_auc(d) - rc.Metric([]).cal_auc(d)) < 1e-9
    tb = torch.zeros((ops_metric.TABLE_ROWS, synth.N_CLS), dtype=torch.float64)
    tb[0] = 2; tb[1] = 2; tb[2] = 1; tb[3] = 0.01
    res = ops_metric.table_to_result(tb, synth.OBJLIST)
    assert set(res) >= {"all_num", "obj_num", "succ_base_rt", "dis_final_t", "dis_xyz"} and len(res) == 23
    s = sweep.summarize(tb, synth.OBJLIST)
    assert s["frames"] == 26 and abs(s["succ_base_rt"] - 0.5) < 1e-12


_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from pose_estimation_b200 import sweep, ops_metric, synth
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank = dist.get_rank()
lo, hi = sweep.shard_range(101, rank, 2)
tb = torch.zeros((ops_metric.TABLE_ROWS, synth.N_CLS), dtype=torch.float64)
for f in range(lo, hi):                       # stand-in for per-rank accumulation
    tb[0, f % synth.N_CLS] += 1; tb[2, f % synth.N_CLS] += f % 2; tb[3, f % synth.N_CLS] += 0.001 * f
sweep.reduce_table(tb)
exp = torch.zeros_like(tb)
for f in range(101):
    exp[0, f % synth.N_CLS] += 1; exp[2, f % synth.N_CLS] += f % 2; exp[3, f % synth.N_CLS] += 0.001 * f
assert torch.allclose(tb, exp, atol=1e-12), rank
assert tb[0].sum() == 101
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_table_allreduce_gloo_world2(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    for p in procs:
        out, _ = p.communicate(timeout=120)
        assert p.returncode == 0 and "ok" in out, out


# ---- model / fixture ingestion (pose_estimation_b200/models.py): host-side parsing
def _write_ply(path, pts, binary):
    with open(path, "wb") as f:
        fmt = "binary_little_endian" if binary else "ascii"
        f.write((f"ply\nformat {fmt} 1.0\ncomment made by a test\nelement vertex {len(pts)}\n"
                 "property float x\nproperty float y\nproperty float z\nproperty uchar red\n"
                 "element face 1\nproperty list uchar int vertex_indices\nend_header\n").encode())
        if binary:
            for p in pts:
                f.write(np.asarray(p, "<f4").tobytes() + bytes([7]))
            f.write(bytes([3]) + np.asarray([0, 1, 2], "<i4").tobytes())
        else:
            for p in pts:
                f.write((" ".join(repr(float(v)) for v in p) + " 7\n").encode())
            f.write(b"3 0 1 2\n")


def test_ply_reader_ascii_and_binary(tmp_path):
    from pose_estimation_b200 import models
    pts = np.random.default_rng(0).normal(size=(50, 3)).astype(np.float32) * 100
    for binary in (False, True):
        p = tmp_path / f"m{int(binary)}.ply"
        _write_ply(p, pts, binary)
        got = models.ply_vtx(p)
        assert got.dtype == np.float32 and np.array_equal(got, pts)
    bad = tmp_path / "bad.ply"
    bad.write_bytes(b"plx\n")
    with pytest.raises(ValueError):
        models.ply_vtx(bad)


def test_models_info_and_model_bank(tmp_path):
    """models_info.yml in the reference's format (millimetres) -> the metre arrays the kernels take; the values
    are those of dataset/linemod/dataset_config/models_info.yml, held in synth.py."""
    from pose_estimation_b200 import models, synth
    yml = tmp_path / "models_info.yml"
    with open(yml, "w") as f:
        for o, (d, mn, sz) in synth._MODELS_INFO_MM.items():
            f.write(f"{o}: {{diameter: {d:.8f}, min_x: {mn[0]:.8f}, min_y: {mn[1]:.8f}, min_z: {mn[2]:.8f}, "
                    f"size_x: {sz[0]:.8f}, size_y: {sz[1]:.8f}, size_z: {sz[2]:.8f}}}\n")
    info = models.load_models_info(yml)
    assert sorted(info) == sorted(synth.OBJLIST)
    rng = np.random.default_rng(1)
    clouds = {o: (info[o]["lfborder"] + rng.random((3000 if o != 9 else 100, 3)) * info[o]["extent"]) for o in synth.OBJLIST}
    bank = models.ModelBank.from_arrays(clouds, info, num_points=2600, seed=3)
    assert bank.model.shape == (13, 2600, 3) and bank.model.dtype == np.float32
    assert np.allclose(bank.extent, synth.EXTENT) and np.allclose(bank.lfborder, synth.LFBORDER)
    assert np.allclose(bank.diameter, synth.DIAMETER) and np.array_equal(bank.is_sym, synth.IS_SYM)
    k = synth.OBJLIST.index(9)                                   # fewer points than asked for: wrap-padded
    assert np.array_equal(bank.model[k][:100], clouds[9].astype(np.float32))
    assert np.array_equal(bank.model[k][100:200], clouds[9].astype(np.float32))
    j = synth.OBJLIST.index(1)                                   # a subset of the class's own points, no repeats
    assert len({tuple(r) for r in bank.model[j]}) == 2600
    again = models.ModelBank.from_arrays(clouds, info, num_points=2600, seed=3)
    assert np.array_equal(again.model, bank.model)
