"""GPU tests of the tcgen05 (tensor-core) nearest-neighbour path."""
import ctypes as C

import numpy as np
import pytest
import torch

from pose_estimation_b200 import _lib

pytestmark = pytest.mark.gpu


def _canon(mat_bf16):
    """[128, 32] bf16 (as uint16) -> canonical tile layout [kc 0..3][row][8]."""
    return np.ascontiguousarray(mat_bf16.reshape(128, 4, 8).transpose(1, 0, 2))


def test_tcgen05_selftest_matches_numpy(dev):
    rng = np.random.default_rng(0)
    A = torch.from_numpy(rng.normal(size=(128, 32)).astype(np.float32)).bfloat16()
    B = torch.from_numpy(rng.normal(size=(128, 32)).astype(np.float32)).bfloat16()
    want = A.double() @ B.double().T
    a = torch.from_numpy(_canon(A.view(torch.int16).numpy())).to(dev)
    b = torch.from_numpy(_canon(B.view(torch.int16).numpy())).to(dev)
    D = torch.zeros((128, 128), dtype=torch.float32, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    _lib.call("pb_tc_selftest", C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), C.c_void_p(D.data_ptr()),
              C.c_void_p(err.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert int(err.item()) == 0
    assert torch.allclose(D.cpu().double(), want, atol=1e-4, rtol=1e-5)


def _brute(x, y):
    """FP32 SqDist argmin with the oracle's operation order (dx*dx, fma dy, fma dz), lowest j on ties."""
    from oracle import cref
    return cref.knn1(x, y)


@pytest.fixture(params=["mma", "tcgen05"])
def sweep(request):
    """Both tensor-core sweeps (tcgen05 + TMEM = the default, warp-level mma.sync) must pass every test below."""
    from pose_estimation_b200 import ops_metric
    ops_metric.set_knn_tc_impl(request.param)
    yield request.param
    ops_metric.set_knn_tc_impl("auto")


@pytest.mark.parametrize("B,N,M,scale", [(3, 500, 500, 0.1), (2, 2620, 2620, 0.1), (1, 1000, 130, 1.0),
                                         (4, 37, 1, 0.1), (1, 129, 257, 5.0)])
def test_knn_tc_indices_equal_exact_fp32(dev, sweep, B, N, M, scale):
    from pose_estimation_b200 import ops_metric
    rng = np.random.default_rng(N * 7 + M)
    x = (rng.random((B, N, 3)) * scale + 0.7).astype(np.float32)          # offset: the centring must cope
    y = (rng.random((B, M, 3)) * scale + 0.7).astype(np.float32)
    if N > 10 and M > 10:
        y[0, 5] = y[0, 3]                       # exact duplicate database points: lowest index wins
        x[0, 1] = y[0, 3]                       # a query sitting exactly on them
        y[0, 9] = y[0, 3] + np.float32(1e-7)    # and a near tie
    xt, yt = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
    idx_tc, d2_tc = ops_metric.knn1(xt, yt, want_d2=True, impl="tc")
    idx_si, d2_si = ops_metric.knn1(xt, yt, want_d2=True, impl="simt")
    stats = ops_metric.LAST_TC_STATS.cpu().numpy()
    assert stats[0] == 0                                                 # no barrier timeout
    if M >= 100 and scale <= 1.0:                                        # the tensor path did the work:
        assert stats[1] <= 0.01 * B * N                                  #   hardly any exact-scan fallback rows
        assert stats[2] <= (160 if sweep == "mma" else 32) * B * N       #   and a few groups (tcgen05: 8 columns, mma: a 128-column tile) of exact evaluations per row
    assert torch.equal(idx_tc, idx_si) and torch.equal(d2_tc, d2_si)     # the two device paths: bit-identical
    for b in range(B):
        ref_idx, ref_d2 = _brute(x[b], y[b])
        assert np.array_equal(idx_tc[b].cpu().numpy(), ref_idx)
        assert np.array_equal(d2_tc[b].cpu().numpy(), ref_d2)


def test_knn_tc_shared_database_and_adversarial_order(dev, sweep):
    """y shared by all batch elements; database sorted by decreasing distance from the queries (every
    column a new running minimum: candidate lists overflow and rows fall back to the exact scan)."""
    from pose_estimation_b200 import ops_metric
    rng = np.random.default_rng(5)
    N, M = 300, 1500
    y = np.zeros((M, 3), np.float32)
    y[:, 0] = np.linspace(2.0, 0.01, M)
    x = (rng.normal(size=(2, N, 3)) * 1e-3).astype(np.float32)
    xt, yt = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
    idx_tc = ops_metric.knn1(xt, yt, impl="tc")
    idx_si = ops_metric.knn1(xt, yt, impl="simt")
    assert torch.equal(idx_tc, idx_si)
    assert (idx_tc == M - 1).float().mean() > 0.9


def test_add_adds_and_pose_loss_on_tensor_cores_equal_simt_bits(dev, sweep):
    """pb_add_adds_tc / pb_pose_loss_fwd_tc: mixed symmetric flags, results bit-identical to the SIMT entries
    and to the CPU oracle's ADD-S."""
    from pose_estimation_b200 import ops_metric
    from oracle import cref
    rng = np.random.default_rng(11)
    B, N = 5, 700
    model = (rng.normal(size=(B, N, 3)) * 0.05).astype(np.float32)
    R = np.stack([np.linalg.qr(rng.normal(size=(3, 3)))[0] for _ in range(B)]).astype(np.float32)
    t = (rng.normal(size=(B, 3)) * 0.02 + [0, 0, 0.8]).astype(np.float32)
    Rg = np.stack([np.linalg.qr(rng.normal(size=(3, 3)))[0] for _ in range(B)]).astype(np.float32)
    mt, Rt, tt, Rgt = (torch.from_numpy(a).to(dev) for a in (model, R, t, Rg))
    pred = ops_metric.transform_points(mt, Rt, tt)
    target = ops_metric.transform_points(mt, Rgt, tt)
    for flags in ([1, 0, 1, 1, 0], [1, 1, 1, 1, 1], [0, 0, 0, 0, 0]):
        want = torch.tensor(flags, dtype=torch.uint8, device=dev)
        a_s, s_s = ops_metric.add_adds(pred, target, want, impl="simt")
        a_t, s_t = ops_metric.add_adds(pred, target, want, impl="tc")
        assert torch.equal(a_s, a_t) and torch.equal(s_s, s_t)
        l_s, p_s, i_s = ops_metric.pose_loss_fwd(Rt, tt, mt, target, want, impl="simt")
        l_t, p_t, i_t = ops_metric.pose_loss_fwd(Rt, tt, mt, target, want, impl="tc")
        assert torch.equal(i_s, i_t) and torch.equal(l_s, l_t) and torch.equal(p_s, p_t)
        l_h, _, i_h = ops_metric.pose_loss_fwd(Rt, tt, mt, target, want, impl="tc", sym_host=flags)
        assert torch.equal(i_s, i_h) and torch.equal(l_s, l_h)
    a_all, s_all = ops_metric.add_adds(pred, target, None, impl="tc")
    a_ref, s_ref = ops_metric.add_adds(pred, target, None, impl="simt")
    assert torch.equal(a_all, a_ref) and torch.equal(s_all, s_ref)
    ra, rs = cref.add_adds(pred[0].cpu().numpy(), target[0].cpu().numpy())
    assert np.isclose(float(a_all[0]), ra, rtol=1e-5) and np.isclose(float(s_all[0]), rs, rtol=1e-5)


@pytest.mark.parametrize("gap_log2", [-11, -13, -14, -15, -17])
def test_knn_tc_adversarial_corners_need_no_tie(dev, sweep, gap_log2):
    """ADVICE (round 1): exactness of the tensor-core argmin rests on  T = 2^-12 R^2 > 2E  with
    E <= 1.06 * 2^-14 R^2 for |p||q| <= R^2 (header of csrc/knn_tc.cu).  The worst case for the bf16 h + m split is
    queries and database points at OPPOSITE corners of the bounding box (|p| ~ |q| ~ R, p anti-parallel to q) whose
    two nearest candidates differ by about 2^gap R^2 in squared distance -- below, at and above the error bound.
    The indices must equal the exact FP32 scan in every case, without relying on an exact tie."""
    from pose_estimation_b200 import ops_metric
    rng = np.random.default_rng(100 - gap_log2)
    N, M = 384, 512
    lo, hi = np.array([0.31, -0.52, 0.83]), np.array([1.37, 0.61, 1.94])      # awkward mantissas on purpose
    R2 = float(((hi - lo) ** 2).sum()) / 4.0
    gap = 2.0 ** gap_log2 * R2
    x = np.empty((1, N, 3), np.float64)
    y = np.empty((1, M, 3), np.float64)
    x[0] = lo + rng.random((N, 3)) * 1e-3 * (hi - lo)                         # queries hug one corner
    y[0, :2] = lo                                                             # keeps the bounding box (and the centre)
    y[0, 2:] = hi - rng.random((M - 2, 3)) * 1e-3 * (hi - lo)                 # database hugs the opposite corner
    y[0, 0] = lo + 0.9 * (hi - lo); y[0, 1] = lo + 0.9 * (hi - lo)            # two far-corner candidates ...
    d = np.linalg.norm(y[0, 0] - x[0].mean(0))
    y[0, 1] += (hi - lo) / np.linalg.norm(hi - lo) * (gap / (2.0 * d))        # ... 2^gap R^2 apart in squared distance
    xt, yt = torch.from_numpy(x.astype(np.float32)).to(dev), torch.from_numpy(y.astype(np.float32)).to(dev)
    idx_tc, d2_tc = ops_metric.knn1(xt, yt, want_d2=True, impl="tc")
    idx_si, d2_si = ops_metric.knn1(xt, yt, want_d2=True, impl="simt")
    assert int(ops_metric.LAST_TC_STATS[0]) == 0
    assert torch.equal(idx_tc, idx_si) and torch.equal(d2_tc, d2_si)
    ref_i, _ = _brute(xt[0].cpu().numpy(), yt[0].cpu().numpy())
    assert np.array_equal(idx_tc[0].cpu().numpy(), ref_i)
