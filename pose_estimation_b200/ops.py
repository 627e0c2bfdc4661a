"""Tensor-level wrappers over the C ABI: torch owns device memory and the stream, nothing else.

Every function takes CUDA tensors, allocates its outputs / workspace with torch, and launches the
sm_100a kernels asynchronously on the current stream of the tensors' device.  No host synchronisation
happens here.  CPU tensors are rejected: there is no fallback implementation.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from . import _lib
from ._lib import RansacParams, call


@dataclass
class RansacConfig:
    """Defaults are the reference's hard-coded constants where it has them
    (tools/trainer.py:426: reprojectionError=1) plus this path's own (256 hypotheses)."""
    n_hyp: int = 256
    thresh_px: float = 1.0
    gn_iters: int = 6
    lo_rounds: int = 0
    min_inliers: int = 5
    seed: int = 0x5EED

    def c_struct(self) -> RansacParams:
        return RansacParams(self.n_hyp, self.thresh_px, self.gn_iters, self.lo_rounds, self.min_inliers,
                            self.seed & 0xFFFFFFFF, (self.seed >> 32) & 0xFFFFFFFF)


class _StreamArg(C.c_void_p):
    """The stream argument of an ABI call plus the device it belongs to (``_lib.call`` makes that device current
    for the duration of the call when it is not)."""
    device_index = None


def _stream(ref):
    """Current stream of the device ``ref`` (a CUDA tensor the call writes) lives on -- not of whatever device
    happens to be current: EvalSweep / MapCompactor / PtsRansac accept ``device=`` without a prior set_device."""
    if not ref.is_cuda:
        raise _lib.PoseB200Error("stream reference must be a CUDA tensor (pose_estimation_b200 has no CPU path)")
    s = _StreamArg(torch.cuda.current_stream(ref.device).cuda_stream)
    s.device_index = ref.device.index
    return s


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _chk(t, dtype, name, allow_none=False):
    if t is None:
        if allow_none:
            return None
        raise TypeError(f"{name} is required")
    if not t.is_cuda:
        raise _lib.PoseB200Error(f"{name} must be a CUDA tensor (pose_estimation_b200 has no CPU path)")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.contiguous()


def _ws(nbytes, device):
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)


# ----------------------------------------------------------------------------- K0
def compact_map(xyz, mask, extent, lfborder, cap, roi=None, want_total=False):
    """xyz [F,3,H,W] f32, mask [F,H,W] u8|None, extent/lfborder [F,3] f32 -> pts [F,5,cap], count [F]."""
    xyz = _chk(xyz, torch.float32, "xyz")
    mask = _chk(mask, torch.uint8, "mask", True)
    extent = _chk(extent, torch.float32, "extent")
    lfborder = _chk(lfborder, torch.float32, "lfborder")
    roi = _chk(roi, torch.int32, "roi", True)
    F, _, H, W = xyz.shape
    dev = xyz.device
    pts = torch.empty((F, 5, cap), dtype=torch.float32, device=dev)
    count = torch.empty((F,), dtype=torch.int32, device=dev)
    total = torch.empty((F,), dtype=torch.int32, device=dev) if want_total else None
    nb = _lib.load().pb_compact_workspace_bytes(F, H, W, cap)
    ws = _ws(nb, dev)
    call("pb_compact_map", _ptr(xyz), _ptr(mask), _ptr(roi), _ptr(extent), _ptr(lfborder), F, H, W, cap,
         _ptr(pts), _ptr(count), _ptr(total), _ptr(ws), ws.numel(), _stream(pts))
    return (pts, count, total) if want_total else (pts, count)


def gather_correspondences(xyz, choose, xmap, ymap, extent, lfborder, cap=None):
    xyz = _chk(xyz, torch.float32, "xyz")
    choose = _chk(choose, torch.int64, "choose")
    xmap = _chk(xmap, torch.float32, "xmap")
    ymap = _chk(ymap, torch.float32, "ymap")
    extent = _chk(extent, torch.float32, "extent")
    lfborder = _chk(lfborder, torch.float32, "lfborder")
    F, _, H, W = xyz.shape
    NP = choose.shape[-1]
    cap = cap or NP
    pts = torch.empty((F, 5, cap), dtype=torch.float32, device=xyz.device)
    count = torch.empty((F,), dtype=torch.int32, device=xyz.device)
    call("pb_gather_correspondences", _ptr(xyz), _ptr(choose), _ptr(xmap), _ptr(ymap), _ptr(extent),
         _ptr(lfborder), F, H, W, NP, cap, _ptr(pts), _ptr(count), _stream(pts))
    return pts, count


class ChooseGatherer:
    """Reusable launcher of pb_gather_choose with persistent pts [F,5,cap] / count [F]: the K0 of callers that hold
    the reference's ``choose`` lists (tools/trainer.py:406-421) instead of a mask."""

    launches = 1

    def __init__(self, F, H, W, cap, device="cuda"):
        self.F, self.H, self.W, self.cap = F, H, W, cap
        dev = torch.device(device)
        self.pts = torch.empty((F, 5, cap), dtype=torch.float32, device=dev)
        self.count = torch.empty((F,), dtype=torch.int32, device=dev)

    def __call__(self, xyz, choose, extent, lfborder, xmap=None, ymap=None, n_valid=None, roi=None):
        """xyz [F,3,H,W] f32: CUDA or PINNED host (read in place); choose [F,NP] int64 / int32 CUDA; xmap / ymap
        [F,NP] f32 CUDA or None (pixel coordinates from the index and roi); n_valid [F] int32 CUDA or None."""
        if tuple(xyz.shape) != (self.F, 3, self.H, self.W) or not xyz.is_contiguous() or xyz.dtype != torch.float32:
            raise ValueError("xyz must be a contiguous float32 [F,3,H,W] tensor of the planned shape")
        if not (xyz.is_cuda or xyz.is_pinned()):
            raise _lib.PoseB200Error("xyz must be a CUDA tensor or a pinned host tensor (no CPU path)")
        if choose.dtype not in (torch.int64, torch.int32) or not choose.is_cuda or not choose.is_contiguous() \
                or choose.dim() != 2 or choose.shape[0] != self.F:
            raise ValueError("choose must be a contiguous CUDA int64 / int32 [F,NP] tensor")
        if (xmap is None) != (ymap is None):
            raise ValueError("xmap and ymap go together")
        NP = choose.shape[1]
        call("pb_gather_choose", _ptr(xyz), _ptr(choose), 1 if choose.dtype == torch.int32 else 0, _ptr(xmap), _ptr(ymap),
             _ptr(n_valid), _ptr(roi), _ptr(extent), _ptr(lfborder), self.F, self.H, self.W, NP, self.cap,
             _ptr(self.pts), _ptr(self.count), _stream(self.pts))
        return self.pts, self.count


def _region_mode(mode):
    if mode not in ("hard", "soft"):
        raise ValueError("region mode must be 'hard' (tools/script/eval.py:93-103) or 'soft' (loss_utils.get_xyz)")
    return 1 if mode == "soft" else 0


def region_decode_points(pts, count, xyz_off, region, region_point, extent, lfborder, choose=None, roi=None, mode="hard"):
    """Rewrites X,Y,Z of compacted correspondences in place: X = (xyz_off + base(region)) * extent + lfborder.
    pts [F,5,cap] / count [F] from compact_map (pixel taken from U,V - roi) or gather_correspondences (pass the same
    ``choose`` [F,NP] int64); xyz_off [F,3,H,W], region [F,n,H,W], region_point [F,n,3] f32.  Returns pts."""
    pts = _chk(pts, torch.float32, "pts")
    count = _chk(count, torch.int32, "count")
    xyz_off = _chk(xyz_off, torch.float32, "xyz_off")
    region = _chk(region, torch.float32, "region")
    region_point = _chk(region_point, torch.float32, "region_point")
    extent = _chk(extent, torch.float32, "extent")
    lfborder = _chk(lfborder, torch.float32, "lfborder")
    choose = _chk(choose, torch.int64, "choose", True)
    roi = _chk(roi, torch.int32, "roi", True)
    F, _, H, W = xyz_off.shape
    n = region.shape[1]
    if region.shape != (F, n, H, W) or region_point.shape != (F, n, 3):
        raise ValueError("region must be [F,n,H,W] and region_point [F,n,3]")
    cap = pts.shape[-1]
    NP = 0 if choose is None else choose.shape[-1]
    call("pb_region_decode_points", _ptr(xyz_off), _ptr(region), _ptr(region_point), _ptr(choose), NP, _ptr(roi),
         _ptr(extent), _ptr(lfborder), F, H, W, n, cap, _region_mode(mode), _ptr(pts), _ptr(count), _stream(pts))
    return pts


def region_decode_map(xyz_off, region, region_point, mode="soft"):
    """Dense form: xyz [F,3,H,W] = xyz_off + base(region) at every pixel (lib/network/loss_utils.py:18-28 for
    mode='soft'; the arg-max form of tools/script/eval.py:93-103 for mode='hard').  Forward only."""
    xyz_off = _chk(xyz_off, torch.float32, "xyz_off")
    region = _chk(region, torch.float32, "region")
    region_point = _chk(region_point, torch.float32, "region_point")
    F, _, H, W = xyz_off.shape
    n = region.shape[1]
    if region.shape != (F, n, H, W) or region_point.shape != (F, n, 3):
        raise ValueError("region must be [F,n,H,W] and region_point [F,n,3]")
    out = torch.empty_like(xyz_off)
    call("pb_region_decode_map", _ptr(xyz_off), _ptr(region), _ptr(region_point), F, H, W, n, _region_mode(mode),
         _ptr(out), _stream(out))
    return out


# ----------------------------------------------------------------------------- K1-K4
def pnp_ransac_pts(pts, count, K4, cfg: RansacConfig | None = None, frame0=0, sample_idx=None,
                   want_counts=False, want_bitmap=False, want_minimal=False):
    """pts [F,5,cap] f32, count [F] i32, K4 [F,4] f32 -> dict of device tensors."""
    cfg = cfg or RansacConfig()
    pts = _chk(pts, torch.float32, "pts")
    count = _chk(count, torch.int32, "count")
    K4 = _chk(K4, torch.float32, "K4")
    sample_idx = _chk(sample_idx, torch.int32, "sample_idx", True)
    F, _, cap = pts.shape
    dev = pts.device
    prm = cfg.c_struct()
    out = dict(R=torch.empty((F, 3, 3), dtype=torch.float32, device=dev),
               t=torch.empty((F, 3), dtype=torch.float32, device=dev),
               n_inl=torch.empty((F,), dtype=torch.int32, device=dev),
               best_hyp=torch.empty((F,), dtype=torch.int32, device=dev),
               n_inl_refined=torch.empty((F,), dtype=torch.int32, device=dev),
               status=torch.empty((F,), dtype=torch.int32, device=dev))
    hc = torch.empty((F, cfg.n_hyp), dtype=torch.int32, device=dev) if want_counts else None
    bm = torch.empty((F, (cap + 31) // 32), dtype=torch.int32, device=dev) if want_bitmap else None
    R0 = torch.empty((F, 3, 3), dtype=torch.float32, device=dev) if want_minimal else None
    t0 = torch.empty((F, 3), dtype=torch.float32, device=dev) if want_minimal else None
    nb = _lib.load().pb_ransac_workspace_bytes(F, cap, C.byref(prm))
    ws = _ws(nb, dev)
    call("pb_pnp_ransac_pts", _ptr(pts), _ptr(count), _ptr(K4), F, cap, frame0, C.byref(prm), _ptr(sample_idx),
         _ptr(out["R"]), _ptr(out["t"]), _ptr(out["n_inl"]), _ptr(out["best_hyp"]), _ptr(out["n_inl_refined"]),
         _ptr(out["status"]), _ptr(hc), _ptr(bm), _ptr(R0), _ptr(t0), _ptr(ws), ws.numel(), _stream(out["R"]))
    if want_counts:
        out["hyp_counts"] = hc
    if want_bitmap:
        out["inl_bitmap"] = bm
    if want_minimal:
        out["R0"], out["t0"] = R0, t0
    return out


_SCORE_IMPL = {"auto": 0, "fp32": 1, "tc": 2, "tc_calib": 3, "tc_prof": 4, "mma": 5, "mma_calib": 6}


def set_score_impl(impl="auto", band_scale=1.0):
    """Which kernel scores the hypotheses (process-wide): 'auto', 'fp32' (FP32-pipe kernels), 'tc' (tensor cores whenever
    the shape allows), 'tc_calib' (tensor cores + calibration statistics).  The integers they return are identical."""
    call("pb_ransac_score_impl", _SCORE_IMPL[impl], float(band_scale))


def ransac_tc_stats(ws, F, cap, cfg: RansacConfig | None = None):
    """Statistics the tensor-core scoring kernel left in a pb_pnp_ransac_pts workspace (synchronises), or None."""
    prm = (cfg or RansacConfig()).c_struct()
    off = _lib.load().pb_ransac_tc_stats_offset(F, cap, C.byref(prm))
    if off < 0:
        return None
    raw = ws.view(torch.uint8)[off:off + 128].cpu().numpy()
    import numpy as np
    i = raw.view(np.int32)
    return dict(timeout=int(i[0]), reevaluated_groups=int(i[1]), calib_max_ratio=float(raw.view(np.float32)[2]),
                calib_mismatch=int(i[3]), clocks=[int(x) for x in raw[16:112].view(np.uint64)])


class MapCompactor:
    """Reusable K0 launcher: persistent pts [F,5,cap] / count [F] / workspace."""

    def __init__(self, F, H, W, cap, device="cuda"):
        self.F, self.H, self.W, self.cap = F, H, W, cap
        dev = torch.device(device)
        nb = _lib.load().pb_compact_workspace_bytes(F, H, W, cap)
        self.sparse_ok = (H * W) % 32 == 0
        self.launches_dense = 2 if nb else 1
        self.ws = _ws(nb, dev)
        self.pts = torch.empty((F, 5, cap), dtype=torch.float32, device=dev)
        self.count = torch.empty((F,), dtype=torch.int32, device=dev)

    def launches(self, has_mask=True):
        """Kernels launched per call: the mask-first kernel is one launch; the maskless stream adds a
        merge launch when frames are split into row bands."""
        return 1 if (has_mask and self.sparse_ok) else self.launches_dense

    def __call__(self, xyz, mask, extent, lfborder, roi=None):
        """xyz / mask may be CUDA tensors or PINNED host tensors (mapped under UVA): with a mask the
        mask-first kernel then reads only the mask and the foreground pixels' sectors across PCIe."""
        if tuple(xyz.shape) != (self.F, 3, self.H, self.W) or not xyz.is_contiguous() or xyz.dtype != torch.float32:
            raise ValueError("xyz must be a contiguous float32 [F,3,H,W] tensor of the planned shape")
        for name, t in (("xyz", xyz), ("mask", mask)):
            if t is not None and not (t.is_cuda or t.is_pinned()):
                raise _lib.PoseB200Error(f"{name} must be a CUDA tensor or a pinned host tensor (no CPU path)")
        if mask is not None and (tuple(mask.shape) != (self.F, self.H, self.W) or mask.dtype != torch.uint8
                                 or not mask.is_contiguous()):
            raise ValueError("mask must be a contiguous uint8 [F,H,W] tensor")
        call("pb_compact_map", _ptr(xyz), _ptr(mask), _ptr(roi), _ptr(extent), _ptr(lfborder), self.F, self.H,
             self.W, self.cap, _ptr(self.pts), _ptr(self.count), None, _ptr(self.ws), self.ws.numel(), _stream(self.pts))
        return self.pts, self.count


class PtsRansac:
    """Reusable K1-K4 launcher with persistent outputs and workspace."""

    launches = 3      # hypotheses, scoring, finalize

    def __init__(self, F, cap, cfg: RansacConfig | None = None, device="cuda"):
        self.F, self.cap = F, cap
        self.cfg = cfg or RansacConfig()
        self.prm = self.cfg.c_struct()
        dev = torch.device(device)
        self.ws = _ws(_lib.load().pb_ransac_workspace_bytes(F, cap, C.byref(self.prm)), dev)
        self.R = torch.empty((F, 3, 3), dtype=torch.float32, device=dev)
        self.t = torch.empty((F, 3), dtype=torch.float32, device=dev)
        self.n_inl = torch.empty((F,), dtype=torch.int32, device=dev)
        self.best_hyp = torch.empty((F,), dtype=torch.int32, device=dev)
        self.n_inl_refined = torch.empty((F,), dtype=torch.int32, device=dev)
        self.status = torch.empty((F,), dtype=torch.int32, device=dev)

    def __call__(self, pts, count, K4, frame0=0):
        call("pb_pnp_ransac_pts", _ptr(pts), _ptr(count), _ptr(K4), self.F, self.cap, frame0, C.byref(self.prm),
             None, _ptr(self.R), _ptr(self.t), _ptr(self.n_inl), _ptr(self.best_hyp), _ptr(self.n_inl_refined),
             _ptr(self.status), None, None, None, None, _ptr(self.ws), self.ws.numel(), _stream(self.R))
        return self


class MapRansac:
    """Reusable map -> pose launcher (K0 + K1-K4) with a persistent workspace and outputs."""

    def __init__(self, F, H, W, cap, cfg: RansacConfig | None = None, device="cuda"):
        self.F, self.H, self.W, self.cap = F, H, W, cap
        self.cfg = cfg or RansacConfig()
        self.prm = self.cfg.c_struct()
        dev = torch.device(device)
        nb = _lib.load().pb_pnp_ransac_map_workspace_bytes(F, H, W, cap, C.byref(self.prm))
        self.ws = _ws(nb, dev)
        self.R = torch.empty((F, 3, 3), dtype=torch.float32, device=dev)
        self.t = torch.empty((F, 3), dtype=torch.float32, device=dev)
        self.n_inl = torch.empty((F,), dtype=torch.int32, device=dev)
        self.best_hyp = torch.empty((F,), dtype=torch.int32, device=dev)
        self.n_inl_refined = torch.empty((F,), dtype=torch.int32, device=dev)
        self.status = torch.empty((F,), dtype=torch.int32, device=dev)

    def __call__(self, xyz, mask, extent, lfborder, K4, roi=None, frame0=0):
        call("pb_pnp_ransac_map", _ptr(xyz), _ptr(mask), _ptr(roi), _ptr(extent), _ptr(lfborder), _ptr(K4),
             self.F, self.H, self.W, self.cap, frame0, C.byref(self.prm), _ptr(self.R), _ptr(self.t),
             _ptr(self.n_inl), _ptr(self.best_hyp), _ptr(self.n_inl_refined), _ptr(self.status),
             _ptr(self.ws), self.ws.numel(), _stream(self.R))
        return self


def resample_inliers(inl_bitmap, count, cap, n, seed=0x5EED, frame0=0):
    inl_bitmap = _chk(inl_bitmap, torch.int32, "inl_bitmap")
    count = _chk(count, torch.int32, "count")
    F = count.shape[0]
    out = torch.empty((F, n), dtype=torch.int32, device=count.device)
    call("pb_resample_inliers", _ptr(inl_bitmap), _ptr(count), F, cap, n, frame0, seed & 0xFFFFFFFF,
         (seed >> 32) & 0xFFFFFFFF, _ptr(out), _stream(out))
    return out


class FrameGenerator:
    """Device-side generator of config C5's frames (csrc/synthgen.cu): writes the foreground pixels of a chunk of F
    frames into a zeroed slab (xyz [F,3,H,W] f32, mask [F,H,W] u8 -- owned here) and erases exactly those pixels
    again, so a sweep over many chunks never touches the 4 MB per frame of background.  Measurement infrastructure:
    the per-frame parameters come from synth.c5_params (host, a pure function of (seed, frame index))."""

    def __init__(self, F, H, W, seed, device="cuda"):
        dev = torch.device(device)
        self.F, self.H, self.W, self.seed = F, H, W, int(seed)
        self.xyz = torch.zeros((F, 3, H, W), dtype=torch.float32, device=dev)
        self.mask = torch.zeros((F, H, W), dtype=torch.uint8, device=dev)
        self.ws = torch.zeros((_lib.load().pb_synth_workspace_bytes(F) // 4,), dtype=torch.int32, device=dev)
        self.n_written = self.ws[F:2 * F]        # foreground pixels per frame of the chunk last written
        self.params, self.frame0, self.bbox_max = None, 0, 1

    def _launch(self, erase):
        call("pb_synth_frames", _ptr(self.params), self.params.shape[0], self.H, self.W, self.bbox_max,
             self.seed & 0xFFFFFFFF, (self.seed >> 32) & 0xFFFFFFFF, self.frame0, _ptr(self.xyz), _ptr(self.mask),
             _ptr(self.ws), self.ws.numel() * 4, 1 if erase else 0, _stream(self.xyz))

    def write(self, params, frame0, bbox_max):
        """params [n <= F, 40] f32 CUDA (synth.c5_params(...)['params']), frame0 = global index of its first frame."""
        if self.params is not None:
            raise RuntimeError("erase() the previous chunk first: the slab must be zero outside the new silhouettes")
        self.params, self.frame0, self.bbox_max = _chk(params, torch.float32, "params"), int(frame0), int(bbox_max)
        self._launch(False)
        return self.xyz, self.mask

    def erase(self):
        if self.params is not None:
            self._launch(True)
            self.params = None
