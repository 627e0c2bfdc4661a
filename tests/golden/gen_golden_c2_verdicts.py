"""Generate tests/golden/c2_ref_verdicts.npz by RUNNING THE REFERENCE'S OWN get_pose + cal_dis over ALL 1024
frames of the benchmark workload (C2: synth.make_frame_sparse(0x5EED, 2, f), f = 0 .. 1023 -- the frames bench.py
times on rank 0).

  Trainer.get_pose / Trainer.cal_dis   /root/reference/tools/trainer.py:370-438 (compiled from the file's syntax
                                       tree by gen_golden_callsites.py, which this script imports; same stand-ins
                                       for the absent kornia / KeOps packages, stated in its header)

Per frame it stores what the reference's test loop (tools/trainer.py:170-191) computes: R, t, (add, r_dis, t_dis)
and the three verdicts  add < 0.1 * diameter, r_dis < 5 deg, t_dis < 0.05 m.  ``torch.randperm`` (trainer.py:407) is
seeded per frame with 2_000_000 + f; the drawn 256 positions of the first 128 frames are stored so that the CPU
oracle (oracle/ref_callsites.get_pose(pick=...)) can replay them without depending on torch's generator.  ``ok`` is
cv2's success flag, which the reference drops (trainer.py:423): it is obtained by repeating the identical cv2 call
through the oracle and checking that it reproduces the reference's tvec bit for bit.

The reference's result depends on its own unseeded ``randperm`` draw, so the whole pass is run a SECOND time with
other seeds (3_000_000 + f) -> ``dis_b`` / ``verdict_b``: the number of frames on which the reference disagrees
with itself is the yardstick for "pass/fail identical" (tests/test_c2_verdicts.py).

Run here only (the GPU box has no /root/reference):   python tests/golden/gen_golden_c2_verdicts.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import gen_golden_callsites as G  # noqa: E402  (stubs kornia/mmcv/pykeops, loads the reference's modules)
from oracle import ref_callsites as rc  # noqa: E402
from pose_estimation_b200 import synth  # noqa: E402

SEED, CONFIG_ID, N_FRAMES, N_MODEL = 0x5EED, 2, 1024, 500
SEED0 = 2_000_000
SEED0_B = 3_000_000
N_PICKS = 128


def main():
    models = synth.class_models(N_MODEL)
    me = types.SimpleNamespace(metric=G.ref_metric.Metric(list(synth.SYM_CLS)))
    R = np.zeros((N_FRAMES, 3, 3), np.float32)
    t = np.zeros((N_FRAMES, 3), np.float32)
    dis = np.zeros((N_FRAMES, 3), np.float64)
    ok = np.zeros(N_FRAMES, np.uint8)
    dis_b = np.zeros((N_FRAMES, 3), np.float64)
    picks = np.zeros((N_PICKS, 256), np.int16)
    for f in range(N_FRAMES):
        sp = synth.make_frame_sparse(SEED, CONFIG_ID, f)
        fr = synth.reference_frame_dict(sp, models[sp["cls"]])
        pred, data = G.torch_frame(fr)
        torch.manual_seed(SEED0 + f)
        pick = torch.randperm(fr["choose"].shape[0])[:256].numpy()        # what trainer.py:407 will draw
        torch.manual_seed(SEED0 + f)
        Rf, tf = G.Trainer.get_pose(pred, data)                             # the reference's method
        add, r_dis, t_dis = G.Trainer.cal_dis(me, Rf, tf, data)             # the reference's method
        R[f], t[f], dis[f] = Rf[0].numpy(), tf[0].numpy(), (add, r_dis, t_dis)
        # cv2's ok flag (dropped by the reference) from the identical call through the oracle
        Ro, to, oko, _ = rc.get_pose(fr["xyz"], fr["extent"], fr["lfborder"], fr["choose"], fr["x_map_choosed"],
                                     fr["y_map_choosed"], fr["intrinsic"], pick=pick)
        assert oko and np.array_equal(to, t[f]), f"frame {f}: oracle replay differs from the reference run"
        ok[f] = 1
        if f < N_PICKS:
            picks[f] = pick
        torch.manual_seed(SEED0_B + f)                                      # the reference against itself
        Rb, tb = G.Trainer.get_pose(pred, data)
        dis_b[f] = G.Trainer.cal_dis(me, Rb, tb, data)
        if f % 128 == 0:
            print(f"frame {f}: add {add:.5f} r {r_dis:.3f} t {t_dis:.5f}", flush=True)
    cls = (np.arange(N_FRAMES) % synth.N_CLS).astype(np.int32)
    verdict = np.stack([dis[:, 0] < 0.1 * synth.DIAMETER[cls], dis[:, 1] < 5.0, dis[:, 2] < 0.05], axis=1)
    verdict_b = np.stack([dis_b[:, 0] < 0.1 * synth.DIAMETER[cls], dis_b[:, 1] < 5.0, dis_b[:, 2] < 0.05], axis=1)
    print("frames on which two runs of the reference disagree:", int((verdict != verdict_b).any(axis=1).sum()))
    out = os.path.join(HERE, "c2_ref_verdicts.npz")
    np.savez_compressed(out, R=R, t=t, dis=dis, ok=ok, verdict=verdict.astype(np.uint8), picks=picks,
                        dis_b=dis_b, verdict_b=verdict_b.astype(np.uint8), seed0_b=np.int64(SEED0_B),
                        seed0=np.int64(SEED0), n_model=np.int32(N_MODEL), config=np.int32(CONFIG_ID),
                        seed=np.int64(SEED))
    print("wrote", out, os.path.getsize(out), "bytes; success rates add / r / t:", verdict.mean(axis=0).round(4).tolist())


if __name__ == "__main__":
    main()
