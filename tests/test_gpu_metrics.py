"""GPU parity: KNN / ADD(-S) / metrics / accuracy table / pose loss / T-vote / inlier resampling."""
import numpy as np
import pytest
import torch

from pose_estimation_b200 import ops, ops_metric, synth

pytestmark = pytest.mark.gpu


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _clouds(rng, B, N, cls=7):
    model = synth.LFBORDER[cls] + rng.random((B, N, 3)) * synth.EXTENT[cls]
    return model.astype(np.float32)


@pytest.mark.parametrize("B,N,M", [(1, 1, 1), (3, 500, 500), (2, 2620, 2620), (1, 5000, 333), (2, 37, 1500)])
def test_knn1_indices_bit_exact(cref, dev, B, N, M):
    rng = np.random.default_rng(N + M)
    x, y = _clouds(rng, B, N), _clouds(rng, B, M)
    if N > 10 and M > 10:
        y[:, 5] = y[:, 3]            # exact ties: the lowest index must win
        x[:, 0] = y[:, 3]
    idx, d2 = ops_metric.knn1(_t(x, dev), _t(y, dev), want_d2=True)
    idx, d2 = idx.cpu().numpy(), d2.cpu().numpy()
    for b in range(B):
        ri, rd = cref.knn1(x[b], y[b])
        assert np.array_equal(idx[b], ri)
        assert np.array_equal(d2[b], rd)
    # shared database
    idx_s = ops_metric.knn1(_t(x, dev), _t(y[0], dev)).cpu().numpy()
    assert np.array_equal(idx_s[0], idx[0])


def test_add_adds_and_transform(cref, dev):
    rng = np.random.default_rng(1)
    B, N = 4, 2620
    model = _clouds(rng, B, N)
    R = np.stack([synth.random_rotation(rng) for _ in range(B)]).astype(np.float32)
    t = rng.normal(size=(B, 3)).astype(np.float32) * 0.1
    Rg = np.stack([synth.random_rotation(rng) for _ in range(B)]).astype(np.float32)
    pred = ops_metric.transform_points(_t(model, dev), _t(R, dev), _t(t, dev))
    tgt = ops_metric.transform_points(_t(model, dev), _t(Rg, dev), _t(t, dev))
    for b in range(B):
        assert np.array_equal(pred[b].cpu().numpy(), cref.transform(model[b], R[b], t[b]))
    want = np.array([1, 0, 1, 1], np.uint8)
    add, adds = ops_metric.add_adds(pred, tgt, _t(want, dev))
    add, adds = add.cpu().numpy(), adds.cpu().numpy()
    p, g = pred.cpu().numpy(), tgt.cpu().numpy()
    for b in range(B):
        ra, rs = cref.add_adds(p[b], g[b], want_adds=bool(want[b]))
        assert abs(add[b] - ra) <= 1e-5 * ra
        assert abs(adds[b] - rs) <= 1e-5 * rs
        assert adds[b] <= add[b] * (1 + 1e-6)          # ADD-S <= ADD


def test_pose_metrics_and_table(cref, dev):
    rng = np.random.default_rng(2)
    F, N = 64, 500
    models = synth.class_models(N)
    cls = (np.arange(F) % synth.N_CLS).astype(np.int32)
    Rg = np.stack([synth.random_rotation(rng) for _ in range(F)]).astype(np.float32)
    tg = (rng.normal(size=(F, 3)) * 0.1 + [0, 0, 0.9]).astype(np.float32)
    # perturbed predictions: small for most, large for some
    R = Rg.copy(); t = tg.copy()
    for f in range(F):
        ang = np.radians(rng.uniform(0, 8)); ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
        Kx = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
        dR = np.eye(3) + np.sin(ang) * Kx + (1 - np.cos(ang)) * Kx @ Kx
        R[f] = (dR @ Rg[f]).astype(np.float32)
        t[f] += (rng.normal(size=3) * 0.02).astype(np.float32)
    add, r_deg, t_m = ops_metric.pose_metrics(_t(R, dev), _t(t, dev), _t(Rg, dev), _t(tg, dev), _t(models, dev),
                                              _t(cls, dev), _t(synth.IS_SYM, dev))
    add_h, r_h, t_h = add.cpu().numpy(), r_deg.cpu().numpy(), t_m.cpu().numpy()
    for f in range(F):
        p = cref.transform(models[cls[f]], R[f], t[f]); g = cref.transform(models[cls[f]], Rg[f], tg[f])
        ra, rs = cref.add_adds(p, g, want_adds=True)
        ref = rs if synth.IS_SYM[cls[f]] else ra
        assert abs(add_h[f] - ref) <= 1e-5 * ref
        assert abs(r_h[f] - cref.angular_deg(R[f], Rg[f])) <= 2e-3       # degrees (acosf implementations)
        assert abs(t_h[f] - np.linalg.norm(t[f] - tg[f])) <= 1e-6
    # explicit target gives the same numbers
    tgt = np.stack([cref.transform(models[cls[f]], Rg[f], tg[f]) for f in range(F)])
    add2, _, _ = ops_metric.pose_metrics(_t(R, dev), _t(t, dev), _t(Rg, dev), _t(tg, dev), _t(models, dev),
                                         _t(cls, dev), _t(synth.IS_SYM, dev), target=_t(tgt, dev))
    assert torch.equal(add, add2)
    # accuracy table (tools/trainer.py:174-191) against numpy, two batches accumulate
    status = np.zeros(F, np.int32); status[5] = 2
    table = ops_metric.new_table(synth.N_CLS, dev)
    for _ in range(2):
        ops_metric.accumulate_results(table, add, r_deg, t_m, _t(cls, dev), _t(synth.DIAMETER, dev), _t(status, dev))
    tb = table.cpu().numpy()
    for c in range(synth.N_CLS):
        m = cls == c
        assert tb[0, c] == 2 * m.sum() and tb[1, c] == 2 * m.sum()
        assert tb[2, c] == 2 * (add_h[m] < np.float32(0.1) * synth.DIAMETER[c]).sum()
        assert abs(tb[3, c] - 2 * add_h[m].astype(np.float64).sum()) < 1e-9
        assert tb[4, c] == 2 * (r_h[m] < 5.0).sum() and tb[6, c] == 2 * (t_h[m] < 0.05).sum()
        assert tb[23, c] == 2 * (status[m] != 0).sum()
    assert tb[8:20].sum() == 0
    res = ops_metric.table_to_result(table, synth.OBJLIST)
    assert res["all_num"][1] == tb[0, 0] and res["succ_base_rt"][15] == tb[2, 12]


def test_vote_translation_and_resample(cref, dev):
    rng = np.random.default_rng(4)
    B, NP = 3, 10000
    cloud = rng.normal(size=(B, NP, 3)).astype(np.float32)
    tres = rng.normal(size=(B, 3, NP)).astype(np.float32)
    t = ops_metric.vote_translation(_t(cloud, dev), _t(tres, dev)).cpu().numpy()
    for b in range(B):
        assert np.abs(t[b] - cref.vote_translation(cloud[b], tres[b])).max() < 2e-6
    # inlier resampling: exact against the oracle (integer work)
    F, cap = 4, 300
    bm = rng.integers(0, 2**32, (F, (cap + 31) // 32), dtype=np.uint64).astype(np.uint32)
    bm[1] = 0; bm[2, 1:] = 0; bm[2, 0] = 0b1011
    count = np.array([300, 300, 300, 77], np.int32)
    for f in range(F):           # bits beyond count must be ignored
        for i in range(count[f], cap):
            bm[f, i >> 5] &= ~np.uint32(1 << (i & 31))
    for n in (125, 10):
        out = ops.resample_inliers(_t(bm.view(np.int32), dev), _t(count, dev), cap, n, seed=99, frame0=3).cpu().numpy()
        for f in range(F):
            assert np.array_equal(out[f], cref.resample_inliers(bm[f], int(count[f]), n, seed=99, frame=3 + f))


def test_pose_loss_forward_backward(dev):
    from pose_estimation_b200 import PoseLoss
    rng = np.random.default_rng(6)
    B, N = 4, 700
    model = torch.from_numpy(_clouds(rng, B, N)).to(dev)
    Rg = torch.from_numpy(np.stack([synth.random_rotation(rng) for _ in range(B)]).astype(np.float32)).to(dev)
    tg = torch.randn(B, 1, 3, device=dev) * 0.1
    targets = model @ Rg.permute(0, 2, 1) + tg
    idxs = torch.tensor([[7], [0], [8], [3]])
    R = Rg.clone().requires_grad_(True)
    t = (tg + 0.01 * torch.randn_like(tg)).requires_grad_(True)
    loss = PoseLoss([7, 8])(R, t, targets, model, idxs)
    loss.backward()
    # plain torch fp32 restatement of lib/network/loss.py:25-41
    R2 = Rg.clone().requires_grad_(True); t2 = t.detach().clone().requires_grad_(True)
    pp = model @ R2.permute(0, 2, 1) + t2
    tg_list = []
    for b in range(B):
        g = targets[b]
        if int(idxs[b]) in (7, 8):
            d = ((pp[b].detach()[:, None, :] - g[None, :, :]) ** 2).sum(-1)
            g = g[d.argmin(1)]
        tg_list.append(g)
    ref = torch.norm(pp - torch.stack(tg_list), dim=2).mean(1).mean()
    ref.backward()
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    assert torch.allclose(R.grad, R2.grad, rtol=1e-3, atol=1e-5)
    assert torch.allclose(t.grad, t2.grad, rtol=1e-3, atol=1e-6)
