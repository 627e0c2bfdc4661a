"""GPU parity of the 3D-3D RANSAC + Umeyama path (csrc/sim3.cu) against the CPU oracle (bit-exact) and, through
the golden vectors made by running the reference (lib/transform/umeyama.py), against the reference itself."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def _golden():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sim3_ref.npz"))


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_sim3_replays_the_reference_draws_and_equals_the_oracle_bit_for_bit(dev):
    from oracle import cref
    from pose_estimation_b200 import ops_sim3
    g = _golden()
    cap = 2048
    B = 5
    src = np.zeros((B, cap, 3), np.float32); dst = np.zeros((B, cap, 3), np.float32)
    cnt = np.zeros(B, np.int32); si = np.zeros((B, 128, 5), np.int32)
    for k in range(B):
        n = g[f"src{k}"].shape[0]
        src[k, :n], dst[k, :n], cnt[k] = g[f"src{k}"], g[f"dst{k}"], n
        idx = g[f"idx{k}"]
        si[k] = np.concatenate([idx, np.repeat(idx[-1:], 128 - idx.shape[0], 0)]) if idx.shape[0] < 128 else idx
    t = lambda a: torch.from_numpy(a).to(dev)
    r = ops_sim3.sim3_ransac(t(src), t(dst), t(cnt), sample_idx=t(si), want_counts=True, want_bitmap=True)
    for k in range(B):
        n = int(cnt[k])
        o = cref.sim3_ransac(src[k, :n], dst[k, :n], sample_idx=si[k])
        it = g[f"idx{k}"].shape[0]
        assert np.array_equal(r.hyp_counts[k].cpu().numpy(), o["hyp_counts"])
        assert np.array_equal(r.hyp_counts[k, :it].cpu().numpy(), g[f"counts{k}"])          # = the reference run
        assert int(r.best_hyp[k]) == o["best_hyp"] == int(np.argmax(g[f"counts{k}"]))
        assert int(r.n_inl[k]) == o["n_inl"] and int(r.status[k]) == o["status"] == 0
        nw = (n + 31) // 32
        assert np.array_equal(r.inl_bitmap[k, :nw].cpu().numpy().view(np.uint32), o["bitmap"])
        assert np.array_equal(_bits(r.R[k].cpu().numpy()), _bits(o["R"]))
        assert np.array_equal(_bits(r.t[k].cpu().numpy()), _bits(o["t"]))
        assert np.array_equal(_bits(r.scale[k].cpu().numpy()), _bits(np.float32(o["scale"])))
        assert np.abs(r.R[k].cpu().numpy() - g[f"R{k}"]).max() <= 1e-6                        # and the reference
        assert np.abs(r.t[k].cpu().numpy() - g[f"t{k}"]).max() <= 1e-6


def test_sim3_counter_rng_path_and_edge_cases_equal_the_oracle(dev):
    from oracle import cref
    from pose_estimation_b200 import ops_sim3
    rng = np.random.default_rng(17)
    B, cap = 6, 700
    src = (rng.random((B, cap, 3)) - 0.5).astype(np.float32)
    dst = np.zeros_like(src)
    cnt = np.array([700, 333, 5, 0, 64, 700], np.int32)
    for b in range(B):
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        w, x, y, z = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                      [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                      [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
        d = (1.0 + 0.3 * b) * src[b] @ R.T + rng.normal(size=3)
        out = rng.random(cap) < 0.4
        d[out] = rng.normal(size=(int(out.sum()), 3))
        dst[b] = d.astype(np.float32)
    dst[5] = rng.normal(size=(cap, 3)).astype(np.float32) * 30            # unrelated clouds: no model
    cfg = ops_sim3.Sim3Config(n_iter=64, seed=0xC0FFEE)
    t = lambda a: torch.from_numpy(a).to(dev)
    r = ops_sim3.sim3_ransac(t(src), t(dst), t(cnt), cfg=cfg, prob0=11, want_counts=True)
    for b in range(B):
        n = int(cnt[b])
        o = cref.sim3_ransac(src[b, :n], dst[b, :n], n_iter=64, seed=0xC0FFEE, prob=11 + b)
        assert int(r.status[b]) == o["status"], b
        assert np.array_equal(r.hyp_counts[b].cpu().numpy(), o["hyp_counts"]), b
        assert int(r.best_hyp[b]) == o["best_hyp"] and int(r.n_inl[b]) == o["n_inl"]
        assert np.array_equal(_bits(r.R[b].cpu().numpy()), _bits(o["R"]))
        assert np.array_equal(_bits(r.t[b].cpu().numpy()), _bits(o["t"]))
    assert int(r.status[3]) == 1                                          # no points
    assert int(r.status[0]) == 0 and int(r.status[1]) == 0


def test_umeyama_fit_and_reference_shaped_api(dev):
    from oracle import cref
    import pose_estimation_b200 as pb
    from pose_estimation_b200 import ops_sim3
    g = _golden()
    for k in range(4):
        s, d = g[f"u_src{k}"], g[f"u_dst{k}"]
        sc, R, t, st = ops_sim3.umeyama_fit(torch.from_numpy(s[None]).to(dev), torch.from_numpy(d[None]).to(dev))
        ost, osc, oR, ot = cref.umeyama_fit(s, d)
        assert int(st[0]) == ost == 0
        assert np.array_equal(_bits(R[0].cpu().numpy()), _bits(oR)) and np.array_equal(_bits(t[0].cpu().numpy()), _bits(ot))
        assert np.abs(R[0].cpu().numpy() - g[f"u_R{k}"]).max() <= 1e-6
        n = s.shape[0]
        SH = np.vstack([s.T, np.ones(n)]); TH = np.vstack([d.T, np.ones(n)])
        S2, R2, T2, M2 = pb.estimateSimilarityUmeyama(SH, TH)               # the reference's signature
        assert abs(S2 - float(g[f"u_scale{k}"])) < 1e-6 and np.abs(M2[:3, 3] - g[f"u_t{k}"]).max() < 1e-6
    S, Rr, Tr, M = pb.estimateSimilarityTransform(g["src0"], g["dst0"])
    assert abs(S - float(g["scale0"])) < 2e-3 and np.abs(Rr - g["R0"]).max() < 2e-3      # other draws, same model
    none = pb.estimateSimilarityTransform(np.random.default_rng(1).random((50, 3)), np.random.default_rng(2).random((50, 3)) * 100)
    assert none == (None, None, None, None) or none[0] is not None


def test_farthest_point_sampling_equals_reference_and_oracle(dev):
    from oracle import cref
    from pose_estimation_b200 import ops_sim3
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fps_ref.npz"))
    for k in range(5):
        pts, want = g[f"pts{k}"], g[f"idx{k}"]
        got = ops_sim3.farthest_point_sampling(torch.from_numpy(pts).to(dev), want.shape[0]).cpu().numpy()
        assert np.array_equal(got, want), k                                    # = the reference's selection
    rng = np.random.default_rng(9)                                             # batched, ragged counts
    B, cap = 4, 5000
    pts = rng.random((B, cap, 3)).astype(np.float32)
    cnt = np.array([5000, 1234, 0, 77], np.int32)
    got = ops_sim3.farthest_point_sampling(torch.from_numpy(pts).to(dev), 128, torch.from_numpy(cnt).to(dev)).cpu().numpy()
    for b in range(B):
        if cnt[b] == 0:
            assert (got[b] == -1).all()
        else:
            assert np.array_equal(got[b], cref.fps(pts[b, :cnt[b]], 128)), b
