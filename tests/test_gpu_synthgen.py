"""The device-side frame generator of config C5 (csrc/synthgen.cu) against its NumPy twin (synth.c5_frame_twin):
identical bytes, and erase() restores the all-zero slab."""
import numpy as np
import pytest
import torch

from pose_estimation_b200 import ops, synth

pytestmark = pytest.mark.gpu


def test_generated_frames_equal_numpy_twin_and_erase(dev):
    seed, frame0, n = 0x5EED, 4090, 12                     # spans a chunk border of the 2048-frame sweep
    P = synth.c5_params(seed, frame0, n)
    gen = ops.FrameGenerator(n, synth.IMG_H, synth.IMG_W, seed, device=dev)
    xyz, mask = gen.write(torch.from_numpy(P["params"]).to(dev), frame0, P["bbox_max"])
    xyz_h, mask_h, area = xyz.cpu().numpy(), mask.cpu().numpy(), gen.n_written.cpu().numpy().copy()
    for i in range(n):
        sp = synth.c5_frame_twin(P["params"][i], seed, frame0 + i)
        want_xyz, want_mask = synth.dense_from_sparse(sp)
        assert area[i] == sp["rows"].shape[0] and 700 < sp["rows"].shape[0] < 1300, i
        assert np.array_equal(mask_h[i], want_mask), i
        assert np.array_equal(xyz_h[i], want_xyz), i
    with pytest.raises(RuntimeError):
        gen.write(torch.from_numpy(P["params"]).to(dev), frame0, P["bbox_max"])
    gen.erase()
    assert int(xyz.abs().sum().item() == 0) and int(mask.sum()) == 0
    # a second chunk into the same slab; frames are a function of (seed, frame index) only
    P2 = synth.c5_params(seed, frame0 + 5, 4)
    assert np.array_equal(P2["params"], P["params"][5:9])
    xyz, mask = gen.write(torch.from_numpy(P2["params"]).to(dev), frame0 + 5, P2["bbox_max"])
    assert np.array_equal(mask[:4].cpu().numpy(), mask_h[5:9]) and np.array_equal(xyz[:4].cpu().numpy(), xyz_h[5:9])
