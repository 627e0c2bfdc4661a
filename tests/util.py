"""Shared helpers for the parity tests."""
import numpy as np

from pose_estimation_b200 import synth


def batch_arrays(sps, H=synth.IMG_H, W=synth.IMG_W):
    """Sparse frames -> dense numpy inputs of the map ABI."""
    F = len(sps)
    xyz = np.zeros((F, 3, H, W), np.float32)
    mask = np.zeros((F, H, W), np.uint8)
    for f, s in enumerate(sps):
        xyz[f, :, s["rows"], s["cols"]] = s["xyzn"]
        mask[f, s["rows"], s["cols"]] = 1
    cls = np.array([s["cls"] for s in sps], np.int32)
    return dict(xyz=xyz, mask=mask, cls=cls, extent=synth.EXTENT[cls].astype(np.float32),
                lfborder=synth.LFBORDER[cls].astype(np.float32), K4=np.tile(synth.LINEMOD_K4, (F, 1)),
                R_gt=np.stack([s["R_gt"] for s in sps]), t_gt=np.stack([s["t_gt"] for s in sps]))


def pack_pts(pts_list, cap):
    """List of [5,P] arrays -> [F,5,cap] padded + count."""
    F = len(pts_list)
    out = np.zeros((F, 5, cap), np.float32)
    cnt = np.zeros(F, np.int32)
    for f, p in enumerate(pts_list):
        n = min(p.shape[1], cap)
        out[f, :, :n] = p[:, :n]
        cnt[f] = n
    return out, cnt


def rot_angle(R1, R2):
    # geodesic angle via the chord |R1-R2|_F = 2*sqrt(2)*sin(angle/2): well conditioned near 0
    # (arccos of the trace loses half the digits there)
    d = np.linalg.norm(np.asarray(R1, np.float64) - np.asarray(R2, np.float64))
    return float(2.0 * np.arcsin(min(d / (2.0 * np.sqrt(2.0)), 1.0)))
