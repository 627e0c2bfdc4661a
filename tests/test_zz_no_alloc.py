"""pb_compact_map_noalloc (no counterpart in the reference): the
mask-first K0 with its scratch in a file-scope device array -- no workspace, no allocation inside the call.
Called straight through the C ABI and compared bit for bit with pb_compact_map on two calls in a row (different
batches, so contents left in the shared scratch by the first call would show in the second).  Collected last."""
import ctypes as C

import numpy as np
import pytest
import torch

from pose_estimation_b200 import _lib, ops, synth
from tests.util import batch_arrays

pytestmark = pytest.mark.gpu


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _noalloc(lib, xyz, mask, roi, ext, lfb, cap, pts, count, total):
    F, _, H, W = xyz.shape
    return lib.pb_compact_map_noalloc(_p(xyz), _p(mask), _p(roi), _p(ext), _p(lfb), F, H, W, cap, _p(pts), _p(count),
                                      _p(total), C.c_void_p(torch.cuda.current_stream(xyz.device).cuda_stream))


def test_noalloc_entry_equals_workspace_entry_on_two_calls_in_a_row(dev):
    lib = _lib.load()
    assert lib.pb_compact_map_noalloc_max_tiles() == 65536
    cap = 6000
    batches = [[synth.make_frame_sparse(0x5EED, 2, f) for f in range(24)],                                   # sparse masks, 5 tiles per frame
               [synth.make_frame_sparse(0x5EED, 12, f, dense=True, z_range=(0.45, 1.0)) for f in range(9)]]   # dense blobs, cut at cap
    # outputs allocated before the calls; the entry point itself takes no workspace
    outs = [(torch.full((len(b), 5, cap), -7.0, device=dev), torch.full((len(b),), -1, dtype=torch.int32, device=dev),
             torch.full((len(b),), -1, dtype=torch.int32, device=dev)) for b in batches]
    for rep in range(2):                                   # twice through both batches: four calls sharing the scratch
        for b, (pts, count, total) in zip(batches, outs):
            a = batch_arrays(b)
            t = lambda k: torch.from_numpy(a[k]).to(dev)
            xyz, mask, ext, lfb = t("xyz"), t("mask"), t("extent"), t("lfborder")
            roi = torch.arange(2 * len(b), dtype=torch.int32, device=dev).reshape(-1, 2)
            want_pts, want_count, want_total = ops.compact_map(xyz, mask, ext, lfb, cap, roi=roi, want_total=True)
            pts.fill_(-7.0)
            rc = _noalloc(lib, xyz, mask, roi, ext, lfb, cap, pts, count, total)
            assert rc == 0, lib.pb_last_error()
            torch.cuda.synchronize(dev)
            assert torch.equal(count, want_count) and torch.equal(total, want_total)
            for f in range(len(b)):
                n = int(count[f])
                assert n == min(b[f]["rows"].shape[0], cap)
                assert torch.equal(pts[f, :, :n], want_pts[f, :, :n]), (rep, f)     # bit-identical (same kernel)


def test_noalloc_entry_rejects_what_it_cannot_hold(dev):
    lib = _lib.load()
    H, W, cap = 16, 16, 8
    F = 65537                                              # one tile per frame: one more than the static scratch holds
    xyz = torch.zeros((F, 3, H, W), device=dev)
    mask = torch.zeros((F, H, W), dtype=torch.uint8, device=dev)
    ext = torch.ones((F, 3), device=dev); lfb = torch.zeros((F, 3), device=dev)
    pts = torch.zeros((F, 5, cap), device=dev); count = torch.zeros((F,), dtype=torch.int32, device=dev)
    assert _noalloc(lib, xyz, mask, None, ext, lfb, cap, pts, count, None) == -1
    assert b"exceed the static scratch" in lib.pb_last_error()
    assert _noalloc(lib, xyz[:4], None, None, ext[:4], lfb[:4], cap, pts[:4], count[:4], None) == -1     # a mask is required
    assert _noalloc(lib, xyz[:65536], mask[:65536], None, ext[:65536], lfb[:65536], cap, pts[:65536], count[:65536], None) == 0
    torch.cuda.synchronize(dev)
    assert int(count[:65536].abs().sum()) == 0            # the largest batch it holds: all-empty masks -> zero counts
