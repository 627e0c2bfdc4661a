"""ctypes binding of libposeb200.so (the C ABI declared in include/poseb200.h).

There is no CPU fallback: if the shared library is missing or a call fails this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libposeb200.so")

P = C.c_void_p
I = C.c_int
U32 = C.c_uint32
F32 = C.c_float
SZ = C.c_size_t


class RansacParams(C.Structure):
    """pb_ransac_params (include/poseb200.h)."""
    _fields_ = [("n_hyp", C.c_int32), ("thresh_px", C.c_float), ("gn_iters", C.c_int32),
                ("lo_rounds", C.c_int32), ("min_inliers", C.c_int32),
                ("seed_lo", C.c_uint32), ("seed_hi", C.c_uint32)]


PRM = C.POINTER(RansacParams)

# name -> (restype, argtypes); must list every symbol include/poseb200.h declares
SIGNATURES = {
    "pb_last_error": (C.c_char_p, []),
    "pb_abi_version": (I, []),
    "pb_compact_workspace_bytes": (SZ, [I, I, I, I]),
    "pb_compact_map": (I, [P, P, P, P, P, I, I, I, I, P, P, P, P, SZ, P]),
    "pb_compact_map_noalloc_max_tiles": (I, []),
    "pb_compact_map_noalloc": (I, [P, P, P, P, P, I, I, I, I, P, P, P, P]),
    "pb_gather_correspondences": (I, [P, P, P, P, P, P, I, I, I, I, I, P, P, P]),
    "pb_gather_choose": (I, [P, P, I, P, P, P, P, P, P, I, I, I, I, I, P, P, P]),
    "pb_region_decode_points": (I, [P, P, P, P, I, P, P, P, I, I, I, I, I, I, P, P, P]),
    "pb_region_decode_map": (I, [P, P, P, I, I, I, I, I, P, P]),
    "pb_ransac_workspace_bytes": (SZ, [I, I, PRM]),
    "pb_pnp_ransac_pts": (I, [P, P, P, I, I, U32, PRM, P, P, P, P, P, P, P, P, P, P, P, P, SZ, P]),
    "pb_ransac_score_impl": (I, [I, F32]),
    "pb_ransac_tc_stats_offset": (C.c_longlong, [I, I, PRM]),
    "pb_pnp_ransac_map_workspace_bytes": (SZ, [I, I, I, I, PRM]),
    "pb_pnp_ransac_map": (I, [P, P, P, P, P, P, I, I, I, I, U32, PRM, P, P, P, P, P, P, P, SZ, P]),
    "pb_resample_inliers": (I, [P, P, I, I, I, U32, U32, U32, P, P]),
    "pb_knn1": (I, [P, P, I, I, I, I, P, P, P]),
    "pb_knn1_tc_workspace_bytes": (SZ, [I, I, I]),
    "pb_knn_tc_impl": (I, [I]),
    "pb_knn1_tc": (I, [P, P, I, I, I, I, P, P, P, SZ, P]),
    "pb_add_adds": (I, [P, P, I, I, P, P, P, P]),
    "pb_add_adds_tc_workspace_bytes": (SZ, [I, I]),
    "pb_add_adds_tc": (I, [P, P, I, I, P, P, I, P, P, P, SZ, P]),
    "pb_transform_points": (I, [P, I, P, P, I, I, P, P]),
    "pb_pose_metrics": (I, [P, P, P, P, P, P, P, P, I, I, I, P, P, P, P]),
    "pb_accumulate_results": (I, [P, P, P, P, P, P, I, I, I, F32, F32, F32, I, P, P]),
    "pb_pose_loss_fwd": (I, [P, P, P, P, P, I, I, P, P, P, P]),
    "pb_pose_loss_fwd_tc": (I, [P, P, P, P, P, P, I, I, I, P, P, P, P, SZ, P]),
    "pb_pose_loss_bwd": (I, [P, P, P, P, I, I, F32, P, P, P]),
    "pb_vote_translation": (I, [P, P, I, I, P, P]),
    "pb_sim3_ransac_workspace_bytes": (SZ, [I, P]),
    "pb_sim3_ransac": (I, [P, P, P, I, I, U32, P, P, P, P, P, P, P, P, P, P, P, SZ, P]),
    "pb_umeyama_fit": (I, [P, P, P, I, I, P, P, P, P, P]),
    "pb_kabsch_fit": (I, [P, P, P, I, I, P, P, P, P, P, P]),
    "pb_icp_refine_workspace_bytes": (SZ, [I, I, I]),
    "pb_icp_refine": (I, [P, I, P, I, P, P, I, I, I, I, I, I, P, P, SZ, P]),
    "pb_backproject_cloud": (I, [P, P, I, P, P, P, P, F32, I, I, I, I, P, P]),
    "pb_uvd2cloud": (I, [P, P, I, I, P, P]),
    "pb_fps_workspace_bytes": (SZ, [I, I]),
    "pb_fps": (I, [P, P, I, I, I, P, P, SZ, P]),
    "pb_pnp_refine": (I, [P, P, I, P, P, P, P, I, I, I, P, P, P, P]),
    "pb_bpnp_backward_workspace_bytes": (SZ, [I, I]),
    "pb_bpnp_backward": (I, [P, P, I, P, P, P, I, I, P, P, P, P, P, SZ, P]),
    "pb_poss_loss_workspace_bytes": (SZ, [I, I, I]),
    "pb_poss_loss_fwd": (I, [P, P, P, P, P, I, I, I, I, F32, P, P, P, P, SZ, P]),
    "pb_poss_loss_bwd": (I, [P, P, P, P, P, I, I, I, I, F32, P, P, P, P, SZ, P, P, P, P]),
    "pb_synth_workspace_bytes": (SZ, [I]),
    "pb_synth_frames": (I, [P, I, I, I, I, U32, U32, U32, P, P, P, SZ, I, P]),
    "pb_probe_fma": (I, [I, I, I, P, P]),
    "pb_tc_selftest": (I, [P, P, P, P, P]),
    "pb_stage_timing": (I, [I]),
    "pb_stage_timing_read": (I, [P, P]),
}

_lib = None


class PoseB200Error(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load the CUDA library; raise loudly if it has not been built (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise PoseB200Error(
            f"{SO_PATH} is missing: build it with `python -m pose_estimation_b200.build` "
            "(nvcc, sm_100a). pose_estimation_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.pb_abi_version() != 2:
        raise PoseB200Error("libposeb200.so ABI version mismatch")
    _lib = lib
    return lib


def call(name: str, *args):
    """Call an int-returning ABI function and raise on a non-zero return code.  When the last argument is a stream
    made by ``ops._stream`` for a device other than the current one, that device is made current around the call
    (a CUDA launch goes to the calling thread's current device, whatever device its pointers live on)."""
    lib = load()
    idx = getattr(args[-1], "device_index", None) if args else None
    if idx is not None:
        import torch
        if idx != torch.cuda.current_device():
            with torch.cuda.device(idx):
                rc = getattr(lib, name)(*args)
            if rc != 0:
                raise PoseB200Error(f"{name} failed ({rc}): {lib.pb_last_error().decode(errors='replace')}")
            return rc
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise PoseB200Error(f"{name} failed ({rc}): {lib.pb_last_error().decode(errors='replace')}")
    return rc
