"""GPU parity through the reference-shaped Python API against the cv2-generated golden vectors
(tests/golden/pnp_cv2.npz, metric.npz).  Reads like the reference's call sites:
get_pose(pred, data) -> cal_dis(...) -> success tests (tools/trainer.py:166-191)."""
import os

import numpy as np
import pytest
import torch

import pose_estimation_b200 as pb
from pose_estimation_b200 import synth
from tests.test_oracle_golden import GOLD, load_frames
from tests.util import batch_arrays, rot_angle

pytestmark = pytest.mark.gpu


def _datas(sp, models):
    fr = synth.reference_frame_dict(sp, models[sp["cls"]])
    pred = {"xyz": torch.from_numpy(fr["xyz"])[None]}
    data = {"extent": torch.from_numpy(fr["extent"])[None], "lfborder": torch.from_numpy(fr["lfborder"])[None],
            "choose": torch.from_numpy(fr["choose"])[None, None], "intrinsic": torch.from_numpy(fr["intrinsic"])[None],
            "x_map_choosed": torch.from_numpy(fr["x_map_choosed"])[None, :, None],
            "y_map_choosed": torch.from_numpy(fr["y_map_choosed"])[None, :, None],
            "target_r": torch.from_numpy(fr["target_r"])[None], "target_t": torch.from_numpy(fr["target_t"])[None],
            "target": torch.from_numpy(fr["target"])[None], "model_points": torch.from_numpy(fr["model_points"])[None],
            "cls_id": torch.LongTensor([fr["cls_id"]])}
    return pred, data


def test_get_pose_cal_dis_vs_cv2_golden():
    g = np.load(os.path.join(GOLD, "pnp_cv2.npz"))
    models = synth.class_models(500)
    metric = pb.Metric(synth.SYM_CLS)
    for tag in ("clean", "noisy"):
        for f, sp in enumerate(load_frames(g, tag)):
            pred, data = _datas(sp, models)
            R, t = pb.get_pose(pred, data)
            assert R.shape == (1, 3, 3) and t.shape == (1, 3) and not R.is_cuda and R.dtype == torch.float32
            if tag == "clean":      # clear consensus: same pose as cv2 to 1e-4 rad / 1e-4 m
                assert rot_angle(R[0].numpy(), g["clean_R"][f]) <= 1e-4
                assert np.abs(t[0].numpy() - g["clean_t"][f]).max() <= 1e-4
            add, r_dis, t_dis = pb.cal_dis(metric, R, t, data)
            assert isinstance(add, float)
            succ = [add < 0.1 * synth.DIAMETER[sp["cls"]], r_dis < 5.0, t_dis < 0.05]
            assert succ == g[f"{tag}_succ"][f].tolist()          # identical pass/fail
            if tag == "clean":
                assert abs(add - g["clean_add"][f]) < 2e-5


def test_metric_knn_loss_vs_golden(dev):
    g = np.load(os.path.join(GOLD, "metric.npz"))
    for cls in (7, 8, 0):
        model, R, t, Rg, tg = (torch.from_numpy(g[f"c{cls}_{k}"]).to(dev) for k in ("model", "R", "t", "Rg", "tg"))
        pred = model @ R.T + t
        tgt = model @ Rg.T + tg
        vals = g[f"c{cls}_vals"]
        add, adds = pb.Metric([cls]).cal_adds_cuda(pred, tgt, cls)
        assert add == adds and abs(adds - vals[1]) <= 1e-5 * vals[1]
        add_ns, adds_ns = pb.Metric([]).cal_adds_cuda(pred.cpu(), tgt.cpu(), cls)     # CPU tensors accepted like the ref
        assert add_ns == adds_ns and abs(add_ns - vals[0]) <= 1e-5 * vals[0]
        r = pb.Metric.angular_distance(R[None], Rg[None])
        assert r.shape == (1, 1) and abs(r.item() - vals[2]) < 2e-3
        assert abs(pb.Metric.translation_distance(t[None], tg[None]).item() - vals[3]) < 1e-6
        idx = pb.knn(pred.contiguous(), tgt.contiguous())
        assert idx.shape == (pred.shape[0], 1) and idx.dtype == torch.int64
        ref_idx = g[f"c{cls}_knn"]
        diff = np.nonzero(idx.cpu().numpy().reshape(-1) != ref_idx)[0]
        assert len(diff) <= 2                                    # near-ties of differently-rounded FP32 formulas only
        loss = pb.PoseLoss([7, 8])(R[None], t[None, None], tgt[None], model[None], torch.tensor([[cls]]))
        assert abs(loss.item() - float(g[f"c{cls}_loss"])) <= 1e-5 * float(g[f"c{cls}_loss"])


def test_auc_matches_reference_restatement():
    from oracle import ref_callsites as rc
    rng = np.random.default_rng(0)
    d = list(rng.random(200) * 0.15)
    assert abs(pb.Metric([]).cal_auc(d) - rc.Metric([]).cal_auc(d)) < 1e-9
    assert pb.Metric([]).cal_auc([0.5, 0.7]) == 0


def test_choose_ransac_batch_and_vote(dev):
    sps = synth.make_batch_sparse(0x5EED, 2, 3, p_range=(400, 500))
    m = min(s["rows"].shape[0] for s in sps)
    coor = np.stack([synth.compact_points(s)[:3, :m].T for s in sps])
    uv = np.stack([synth.compact_points(s)[3:, :m].T for s in sps])
    K = np.tile(np.array([[572.4114, 0, 325.2611], [0, 573.57043, 242.04899], [0, 0, 1]], np.float32), (3, 1, 1))
    out = pb.choose_ransac_batch(coor, uv, K, n=125)
    assert out.shape == (3, 1, 125) and out.dtype == np.int32
    for b in range(3):
        sel = out[b, 0]
        u = np.unique(sel)
        assert sel.min() >= 0 and sel.max() < m
        if len(u) == 125:
            assert np.all(np.diff(sel) > 0)                  # order-preserving random subset of the inliers
        else:
            assert np.array_equal(sel, np.resize(u, 125))    # fewer inliers than n: np.pad(..., 'wrap')
        assert sps[b]["inlier"][sel].mean() > 0.95
    out_pad = pb.choose_ransac_batch(coor, uv, K, n=m)        # fewer inliers than n -> wrap padding
    assert out_pad.shape == (3, 1, m)
    p = torch.randn(2, 500, 3, device=dev, requires_grad=True)
    r = torch.randn(2, 3, 500, device=dev, requires_grad=True)
    v = pb.vote_translation(p, r)
    ref = (p + r.permute(0, 2, 1)).mean(dim=1)
    assert torch.allclose(v, ref, atol=1e-6)
    v.sum().backward()
    assert torch.allclose(p.grad, torch.full_like(p, 1 / 500)) and torch.allclose(r.grad, torch.full_like(r, 1 / 500))


def test_eval_sweep_table(dev):
    F = 26
    sps = synth.make_batch_sparse(0x5EED, 2, F)
    b = batch_arrays(sps)
    models = synth.class_models(500)
    sw = pb.EvalSweep(torch.from_numpy(models), torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM),
                      F, synth.IMG_H, synth.IMG_W, 1280, device=dev)
    t = lambda a: torch.from_numpy(a).to(dev)
    add, r_deg, t_m = sw.step(t(b["xyz"]), t(b["mask"]), t(b["extent"]), t(b["lfborder"]), t(b["K4"]), t(b["cls"]),
                              t(b["R_gt"]), t(b["t_gt"]))
    table = sw.finish()
    s = pb.summarize(table, synth.OBJLIST)
    assert s["frames"] == F and s["failed"] == 0
    assert s["succ_base_rt"] >= 0.9 and s["objects"][1]["num"] == 2
    assert (pb.batch_intrinsic_transform(pb.batch_intrinsic_transform(t(b["K4"]))) == t(b["K4"])).all()


def test_host_eval_sweep_staged_and_zero_copy_equal_device(dev):
    """HostEvalSweep (pinned host maps; staged H2D or zero-copy) gives the device path's table bit for bit."""
    from pose_estimation_b200 import sweep
    F, chunk = 16, 8
    sps = synth.make_batch_sparse(0x5EED, 2, F)
    b = batch_arrays(sps)
    models = torch.from_numpy(synth.class_models(500))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    t = lambda a: torch.from_numpy(a).to(dev)
    sw = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev)
    sw.step(t(b["xyz"]), t(b["mask"]), t(b["extent"]), t(b["lfborder"]), t(b["K4"]), t(b["cls"]), t(b["R_gt"]), t(b["t_gt"]))
    want = sw.finish().cpu()
    sw2 = pb.EvalSweep(models, diam, sym, 2 * F, synth.IMG_H, synth.IMG_W, 1280, device=dev)
    d = lambda k: torch.cat([t(b[k]), t(b[k])])
    sw2.step(d("xyz"), d("mask"), d("extent"), d("lfborder"), d("K4"), d("cls"), d("R_gt"), d("t_gt"))
    want2 = sw2.finish().cpu()
    xyz_h, mask_h = torch.from_numpy(b["xyz"]).pin_memory(), torch.from_numpy(b["mask"]).pin_memory()
    cls_h = torch.from_numpy(b["cls"]).pin_memory()
    for zc in (False, True, "xyz"):
        hs = sweep.HostEvalSweep(models, diam, sym, synth.IMG_H, synth.IMG_W, 1280, chunk=chunk, device=dev, zero_copy=zc)
        small_h = sweep.pack_small(b["extent"], b["lfborder"], b["K4"], b["R_gt"], b["t_gt"], chunk).pin_memory()
        got = hs.run(xyz_h, mask_h, small_h, cls_h)
        assert torch.equal(got, want), zc
        assert hs.h2d_bytes > 0
        hs.reset()                                   # a second pass over 4 chunks exercises both lanes / staging buffers
        got = hs.run(torch.cat([xyz_h, xyz_h]).pin_memory(), torch.cat([mask_h, mask_h]).pin_memory(),
                     torch.cat([small_h, small_h]).pin_memory(), torch.cat([cls_h, cls_h]).pin_memory())
        assert torch.equal(got, want2), zc


def test_host_eval_sweep_on_object_crops_equals_full_frames(dev):
    """The reference's network emits an object crop of the map (get_square_bbox, batchdataset.py:890-961) and the
    crop's offset; HostEvalSweep.run(roi_h=...) on 240x240 crops gives the full-frame table bit for bit."""
    from pose_estimation_b200 import sweep
    F, chunk, S = 16, 8, 240
    sps = synth.make_batch_sparse(0x5EED, 2, F)
    b = batch_arrays(sps)
    models = torch.from_numpy(synth.class_models(500))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    t = lambda a: torch.from_numpy(a).to(dev)
    sw = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev)
    sw.step(t(b["xyz"]), t(b["mask"]), t(b["extent"]), t(b["lfborder"]), t(b["K4"]), t(b["cls"]), t(b["R_gt"]), t(b["t_gt"]))
    want = sw.finish().cpu()
    xc = np.zeros((F, 3, S, S), np.float32); mc = np.zeros((F, S, S), np.uint8); roi = np.zeros((F, 2), np.int32)
    for f, sp in enumerate(sps):
        r0 = int(np.clip((sp["rows"].min() + sp["rows"].max()) // 2 - S // 2, 0, synth.IMG_H - S))
        c0 = int(np.clip((sp["cols"].min() + sp["cols"].max()) // 2 - S // 2, 0, synth.IMG_W - S))
        assert sp["rows"].max() - r0 < S and sp["cols"].max() - c0 < S
        xc[f] = b["xyz"][f][:, r0:r0 + S, c0:c0 + S]; mc[f] = b["mask"][f][r0:r0 + S, c0:c0 + S]; roi[f] = (r0, c0)
    small_h = sweep.pack_small(b["extent"], b["lfborder"], b["K4"], b["R_gt"], b["t_gt"], chunk).pin_memory()
    pin = lambda a: torch.from_numpy(a).pin_memory()
    for zc in (False, "xyz"):
        hs = sweep.HostEvalSweep(models, diam, sym, S, S, 1280, chunk=chunk, device=dev, zero_copy=zc)
        got = hs.run(pin(xc), pin(mc), small_h, pin(b["cls"]), roi_h=pin(roi))
        assert torch.equal(got, want), zc


def test_eval_sweep_stream_lanes_and_graph_replay_equal_single_stream(dev):
    """EvalSweep(lanes=2) (two streams) and a CUDA-graph replay of the step give the single-stream table and
    per-frame outputs bit for bit."""
    F = 16
    sps = synth.make_batch_sparse(0x5EED, 2, F)
    b = batch_arrays(sps)
    models = torch.from_numpy(synth.class_models(500))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    t = lambda a: torch.from_numpy(a).to(dev)
    args = [t(b[k]) for k in ("xyz", "mask", "extent", "lfborder", "K4", "cls", "R_gt", "t_gt")]
    sw1 = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev)
    add1, r1, t1 = (x.clone() for x in sw1.step(*args))
    want = sw1.finish().clone()
    sw2 = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev, lanes=2)
    add2, r2, t2 = sw2.step(*args)
    torch.cuda.synchronize()
    assert torch.equal(sw2.finish(), want)
    assert torch.equal(add2, add1) and torch.equal(r2, r1) and torch.equal(t2, t1)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        sw2.reset(); sw2.step(*args)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        sw2.reset()
        sw2.step(*args)
    g.replay(); g.replay()
    torch.cuda.synchronize()
    assert torch.equal(sw2.table, want)


def test_cpu_tensors_have_no_fallback():
    from pose_estimation_b200 import ops_metric
    from pose_estimation_b200._lib import PoseB200Error
    with pytest.raises(PoseB200Error):
        ops_metric.knn1(torch.zeros(4, 3), torch.zeros(4, 3))


def test_eval_sweep_opt_pose_variants(dev):
    """tools/trainer.py:196-221 (opt_pose): 'reg' and 'final' pair the PnP rotation with the voted translation."""
    from pose_estimation_b200 import ops_metric
    F, NP = 26, 500
    sps = synth.make_batch_sparse(0x5EED, 2, F)
    b = batch_arrays(sps)
    models = torch.from_numpy(synth.class_models(500))
    t = lambda a: torch.from_numpy(a).to(dev)
    args = [t(b[k]) for k in ("xyz", "mask", "extent", "lfborder", "K4", "cls", "R_gt", "t_gt")]
    sw = pb.EvalSweep(models, torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM), F, synth.IMG_H,
                      synth.IMG_W, 1280, device=dev)
    # every cloud point votes t_gt exactly (offset = t_gt - point): the voted translation is the ground truth
    g = torch.Generator(device=dev).manual_seed(2)
    cloud = torch.randn(F, NP, 3, device=dev, generator=g) * 0.05 + args[7][:, None, :]
    t_res = (args[7][:, None, :] - cloud).permute(0, 2, 1).contiguous()
    sw.step(*args, cloud=cloud, t_res=t_res)
    res = ops_metric.table_to_result(sw.finish(), synth.OBJLIST)
    tot = lambda k: sum(res[k].values())
    assert tot("obj_num") == F
    assert tot("dis_reg_r") == tot("dis_base_r") and tot("succ_reg_r") == tot("succ_base_r")     # same rotation
    assert tot("dis_reg_t") < 1e-5 * F and tot("succ_reg_t") == F                               # voted translation = GT
    assert tot("succ_reg_rt") >= tot("succ_base_rt")
    for k in ("dis_{}_rt", "dis_{}_r", "dis_{}_t", "succ_{}_rt", "succ_{}_r", "succ_{}_t"):
        assert res[k.format("final")] == res[k.format("reg")]                                  # trainer.py:198-199
    # an explicit pred_t gives the same rows as voting it here
    sw2 = pb.EvalSweep(models, torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM), F, synth.IMG_H,
                       synth.IMG_W, 1280, device=dev)
    sw2.step(*args, pred_t=pb.vote_translation(cloud, t_res))
    assert torch.equal(sw2.finish(), sw.table)
    # and matches the per-frame call sites (frame 0: get_pose keys its hypothesis stream on frame index 0)
    metric = pb.Metric(synth.SYM_CLS)
    pred, data = _datas(sps[0], synth.class_models(500))
    R, _ = pb.get_pose({"xyz": pred["xyz"]}, data)
    a, _, _ = pb.cal_dis(metric, R, pb.vote_translation(cloud[:1], t_res[:1]).cpu(), data)
    assert abs(a - float(sw.last_reg[0][0])) <= 1e-5 * max(a, 1e-6) + 1e-7


def test_eval_sweep_free_running_lanes_equal_single_stream(dev):
    """EvalSweep(lanes=2, free_running=True): batches follow each other on each lane without a join; the table after
    finish() and the per-frame outputs after join() equal the single-stream ones, also across reset()."""
    F = 16
    sps = synth.make_batch_sparse(0x5EED, 2, F)
    b = batch_arrays(sps)
    models = torch.from_numpy(synth.class_models(500))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    t = lambda a: torch.from_numpy(a).to(dev)
    args = [t(b[k]) for k in ("xyz", "mask", "extent", "lfborder", "K4", "cls", "R_gt", "t_gt")]
    sw1 = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev)
    for k in range(3):
        out1 = [x.clone() for x in sw1.step(*args, frame0=k * F)]
    want = sw1.finish().clone()
    sw2 = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev, lanes=2, free_running=True)
    for rep in range(2):
        sw2.reset()
        for k in range(3):
            out2 = sw2.step(*args, frame0=k * F)
        sw2.join()
        assert all(torch.equal(a, c) for a, c in zip(out1, out2)), rep
        assert torch.equal(sw2.finish(), want), rep
    # whole batches dealt in turn to three streams
    sw3 = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev, lanes=3, deal="alternate")
    for rep in range(2):
        sw3.reset()
        outs = [sw3.step(*args, frame0=k * F) for k in range(3)]
        sw3.join()
        assert all(torch.equal(a, c) for a, c in zip(out1, outs[2])), rep
        assert torch.equal(sw3.finish(), want), rep


def test_eval_sweep_lanes_inputs_freed_or_overwritten_between_steps(dev):
    """The caller may drop a batch as soon as step() returns (the lanes record the tensors on their streams) and
    may overwrite it once ``last_event`` has been waited on: the table equals the single-stream run's."""
    F, steps = 16, 6
    sps = synth.make_batch_sparse(0x5EED, 2, F * steps)
    models = torch.from_numpy(synth.class_models(500))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    keys = ("xyz", "mask", "extent", "lfborder", "K4", "cls", "R_gt", "t_gt")
    host = [batch_arrays(sps[k * F:(k + 1) * F]) for k in range(steps)]
    sw1 = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev)
    for k in range(steps):
        sw1.step(*[torch.from_numpy(host[k][n]).to(dev) for n in keys], frame0=k * F)
    want = sw1.finish().clone()
    # (a) fresh tensors per step, dropped at once, with allocator churn of the same sizes in between
    sw3 = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev, lanes=3, deal="alternate")
    for k in range(steps):
        args = [torch.from_numpy(host[k][n]).to(dev) for n in keys]
        sw3.step(*args, frame0=k * F)
        del args
        junk = [torch.full((F, 3, synth.IMG_H, synth.IMG_W), 7.0, device=dev), torch.full((F, synth.IMG_H, synth.IMG_W), 1, dtype=torch.uint8, device=dev)]
        del junk
    assert torch.equal(sw3.finish(), want)
    # (b) one set of buffers overwritten in place after waiting for the lane that read them
    bufs = [torch.from_numpy(host[0][n]).to(dev) for n in keys]
    sw3.reset()
    for k in range(steps):
        if k:
            torch.cuda.current_stream(dev).wait_event(sw3.last_event)
            for bt, n in zip(bufs, keys):
                bt.copy_(torch.from_numpy(host[k][n]), non_blocking=False)
        sw3.step(*bufs, frame0=k * F)
    assert torch.equal(sw3.finish(), want)


def _choose_lists(sps, NP):
    """The reference's per-frame lists (batchdataset.py:673-687, 712-716) for sparse frames: choose = flat index of
    every masked pixel in row-major order, x/y_map_choosed = its column / row; ragged -> padded with n_valid."""
    F = len(sps)
    choose = np.zeros((F, NP), np.int64)
    xm, ym = np.zeros((F, NP), np.float32), np.zeros((F, NP), np.float32)
    nv = np.zeros(F, np.int32)
    for f, sp in enumerate(sps):
        n = sp["rows"].shape[0]
        choose[f, :n] = sp["rows"].astype(np.int64) * synth.IMG_W + sp["cols"]
        xm[f, :n], ym[f, :n], nv[f] = sp["cols"], sp["rows"], n
    return choose, xm, ym, nv


def test_gather_choose_forms_equal_mask_compaction(dev):
    """pb_gather_choose: int64 lists with coordinate lists (the reference's interface), int32 lists with the
    coordinates taken from the index, a crop with roi -- all give the mask-first compaction's output bit for bit."""
    from pose_estimation_b200 import ops
    F, cap = 6, 1280
    sps = synth.make_batch_sparse(0x5EED, 2, F)
    b = batch_arrays(sps)
    t = lambda a: torch.from_numpy(a).to(dev)
    want_pts, want_cnt = ops.compact_map(t(b["xyz"]), t(b["mask"]), t(b["extent"]), t(b["lfborder"]), cap)
    choose, xm, ym, nv = _choose_lists(sps, 1216)
    g = ops.ChooseGatherer(F, synth.IMG_H, synth.IMG_W, cap, device=dev)
    forms = [dict(choose=t(choose), xmap=t(xm), ymap=t(ym), n_valid=t(nv)),
             dict(choose=t(choose.astype(np.int32)), n_valid=t(nv)),
             dict(choose=t(choose), n_valid=t(nv))]
    for kw in forms:
        pts, cnt = g(t(b["xyz"]), kw.pop("choose"), t(b["extent"]), t(b["lfborder"]), **kw)
        assert torch.equal(cnt, want_cnt)
        for f in range(F):
            n = int(cnt[f])
            assert torch.equal(pts[f, :, :n], want_pts[f, :, :n]), f
    # pinned host map read in place
    pts, cnt = g(torch.from_numpy(b["xyz"]).pin_memory(), t(choose), t(b["extent"]), t(b["lfborder"]), n_valid=t(nv))
    assert all(torch.equal(pts[f, :, :int(cnt[f])], want_pts[f, :, :int(cnt[f])]) for f in range(F))
    # wrap-padded fixed-length lists (the reference pads with np.pad(..., 'wrap')): NP entries, duplicates included
    NPf = 1300
    wrap = np.stack([np.resize(choose[f, :nv[f]], NPf) for f in range(F)])
    g2 = ops.ChooseGatherer(F, synth.IMG_H, synth.IMG_W, NPf, device=dev)
    pts, cnt = g2(t(b["xyz"]), t(wrap), t(b["extent"]), t(b["lfborder"]))
    assert (cnt == NPf).all()
    for f in range(F):
        n = int(nv[f])
        assert torch.equal(pts[f, :, :n], want_pts[f, :, :n]) and torch.equal(pts[f, :, n:NPf], want_pts[f, :, :NPf - n])


def test_host_eval_sweep_choose_lists_equal_device_mask_path(dev):
    """HostEvalSweep.run_choose (the reference's host-side interface: maps + choose / x_map / y_map lists) from
    pinned host maps and from device-resident maps: the accuracy table of the mask path, bit for bit."""
    from pose_estimation_b200 import sweep
    F, chunk = 32, 8
    sps = synth.make_batch_sparse(0x5EED, 2, F)
    b = batch_arrays(sps)
    models = torch.from_numpy(synth.class_models(500))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    t = lambda a: torch.from_numpy(a).to(dev)
    sw = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev)
    sw.step(t(b["xyz"]), t(b["mask"]), t(b["extent"]), t(b["lfborder"]), t(b["K4"]), t(b["cls"]), t(b["R_gt"]), t(b["t_gt"]))
    want = sw.finish().cpu()
    choose, xm, ym, nv = _choose_lists(sps, 1216)
    pin = lambda a: torch.from_numpy(a).pin_memory()
    small_h = sweep.pack_small(b["extent"], b["lfborder"], b["K4"], b["R_gt"], b["t_gt"], chunk).pin_memory()
    hs = sweep.HostEvalSweep(models, diam, sym, synth.IMG_H, synth.IMG_W, 1280, chunk=chunk, device=dev, zero_copy=True)
    xyz_h = pin(b["xyz"])
    got = hs.run_choose(xyz_h, pin(choose), small_h, pin(b["cls"]), xmap_h=pin(xm), ymap_h=pin(ym), n_valid_h=pin(nv))
    assert torch.equal(got, want) and hs.h2d_bytes == F * 1216 * 16 + F * 4 + small_h.numel() * 4 + F * 4
    hs.reset()
    got = hs.run_choose(t(b["xyz"]), pin(choose.astype(np.int32)), small_h, pin(b["cls"]), n_valid_h=pin(nv))
    assert torch.equal(got, want)
    # the sweep object's own step(choose=...) with lanes
    sw3 = pb.EvalSweep(models, diam, sym, F, synth.IMG_H, synth.IMG_W, 1280, device=dev, lanes=2)
    sw3.step(t(b["xyz"]), None, t(b["extent"]), t(b["lfborder"]), t(b["K4"]), t(b["cls"]), t(b["R_gt"]), t(b["t_gt"]),
             choose=t(choose), xmap=t(xm), ymap=t(ym), n_valid=t(nv))
    assert torch.equal(sw3.finish().cpu(), want)


def test_host_eval_sweep_choose_graph_replay_equals_eager(dev):
    """HostEvalSweep.capture_choose: the pass as one replayed CUDA graph gives run_choose's table, also after the
    pinned buffers have been refilled with another batch (same addresses, new contents)."""
    from pose_estimation_b200 import sweep
    F, chunk = 32, 8
    models = torch.from_numpy(synth.class_models(500))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    hs = sweep.HostEvalSweep(models, diam, sym, synth.IMG_H, synth.IMG_W, 1280, chunk=chunk, device=dev, zero_copy=True)
    bufs, replay = None, None
    for first in (0, 100):
        sps = synth.make_batch_sparse(0x5EED, 2, F, first_frame=first)
        b = batch_arrays(sps)
        choose, xm, ym, nv = _choose_lists(sps, 1216)
        small = sweep.pack_small(b["extent"], b["lfborder"], b["K4"], b["R_gt"], b["t_gt"], chunk)
        new = [torch.from_numpy(b["xyz"]), torch.from_numpy(choose), small, torch.from_numpy(b["cls"]), torch.from_numpy(xm),
               torch.from_numpy(ym), torch.from_numpy(nv)]
        if bufs is None:
            bufs = [t.pin_memory() for t in new]
            replay = hs.capture_choose(bufs[0], bufs[1], bufs[2], bufs[3], xmap_h=bufs[4], ymap_h=bufs[5], n_valid_h=bufs[6])
        else:
            for dst, src in zip(bufs, new):
                dst.copy_(src)
        got = replay().clone()
        hs.reset()
        want = hs.run_choose(bufs[0], bufs[1], bufs[2], bufs[3], xmap_h=bufs[4], ymap_h=bufs[5], n_valid_h=bufs[6]).clone()
        assert torch.equal(got, want) and got[0].sum() == F, first
        assert torch.equal(replay(), want)                                   # replays do not accumulate


def test_pipelined_passes_return_each_pass_table_in_order(dev):
    """sweep.PipelinedPasses: two captured passes (own sweep objects, own pinned buffers holding DIFFERENT batches) kept in
    flight together return, in submission order, exactly the tables the synchronous replays give."""
    from pose_estimation_b200 import sweep
    F, chunk = 32, 8
    models = torch.from_numpy(synth.class_models(500))
    diam, sym = torch.from_numpy(synth.DIAMETER), torch.from_numpy(synth.IS_SYM)
    replays, want = [], []
    for first in (0, 100):
        sps = synth.make_batch_sparse(0x5EED, 2, F, first_frame=first)
        b = batch_arrays(sps)
        choose, xm, ym, nv = _choose_lists(sps, 1216)
        small = sweep.pack_small(b["extent"], b["lfborder"], b["K4"], b["R_gt"], b["t_gt"], chunk)
        bufs = [t.pin_memory() for t in (torch.from_numpy(b["xyz"]), torch.from_numpy(choose), small, torch.from_numpy(b["cls"]),
                                         torch.from_numpy(xm), torch.from_numpy(ym), torch.from_numpy(nv))]
        hs = sweep.HostEvalSweep(models, diam, sym, synth.IMG_H, synth.IMG_W, 1280, chunk=chunk, device=dev, zero_copy=True)
        r = hs.capture_choose(bufs[0], bufs[1], bufs[2], bufs[3], xmap_h=bufs[4], ymap_h=bufs[5], n_valid_h=bufs[6])
        replays.append(r)
        want.append(r().clone())
    assert not torch.equal(want[0], want[1])
    pipe = sweep.PipelinedPasses(replays, device=dev)
    got = [pipe.submit() for _ in range(5)]                    # passes 0 1 0 1 0: None, then tables of passes 0 1 0 1
    assert got[0] is None
    for k, tb in enumerate(got[1:]):
        assert torch.equal(tb, want[k & 1]), k
    rest = pipe.drain()
    assert len(rest) == 1 and torch.equal(rest[0], want[0]) and pipe.drain() == []
    with pytest.raises(ValueError):
        sweep.PipelinedPasses(replays[:1], device=dev)
