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
