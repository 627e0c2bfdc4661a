"""Seeded synthetic LineMOD-shaped inputs for tests and bench.

Input generation only -- not part of the timed hot path.  Shapes and conventions follow the
reference's dataset code (dataset/linemod/batchdataset.py:667-771): a normalised model-coordinate
map ``xyz [3,h,w]`` (0 = background, :688-693), a mask, integer pixel coordinates, per-class
``extent`` / ``lfborder`` / ``diameter`` taken from dataset_config/models_info.yml (mm / 1000,
batchdataset.py:163-165), LineMOD intrinsics (:80-82), ``target = model @ R^T + t`` (:707).

Each frame shows an ellipsoid inscribed in the object's box; a pixel's model coordinate is the
ray/ellipsoid hit, so correspondences are exact before noise/outliers are injected.

DEVIATION from SURVEY.md section 8(d), which asks for a compact disc of ~1000 px (radius 18 px): the mask is a
seeded uniform SUBSET (P ~ U{800..1200}) of the object's silhouette, which spans up to ~211 px.  Reason: a compact
blob of ~1000 px on a smooth surface makes PnP itself ill-conditioned (every pose within ~1 cm reprojects the
whole blob to < 1 px), which says nothing about either implementation -- and a subset of the silhouette is what
the reference's own data pipeline hands to get_pose: ``choose`` is a random subset of the mask's pixels
(dataset/linemod/batchdataset.py:673-687).  Consequences, recorded in DESIGN.md section 2 and BASELINE.md: the
gather touches one 32-byte sector per pixel and plane (scattered pixels) where a compact blob shares sectors 8:1.
``dense=True`` gives the whole silhouette (P = 2 k .. 30 k) for tests and bench_configs.py's dense-mask lines.
"""
from __future__ import annotations

import numpy as np

IMG_H, IMG_W = 480, 640
LINEMOD_K4 = np.array([572.4114, 573.57043, 325.2611, 242.04899], dtype=np.float32)
OBJLIST = [1, 2, 4, 5, 6, 8, 9, 10, 11, 12, 13, 14, 15]     # batchdataset.py:42
SYM_CLS = [7, 8]                                            # indices into OBJLIST (eggbox, glue), :76

# dataset_config/models_info.yml: id -> (diameter, (min_x,min_y,min_z), (size_x,size_y,size_z)) in mm
_MODELS_INFO_MM = {
    1: (102.09865663, (-37.9343, -38.7996, -45.8845), (75.8686, 77.5992, 91.769)),
    2: (247.50624233, (-107.835, -60.9279, -109.705), (215.67, 121.8557, 219.41)),
    4: (172.49224865, (-68.3297, -71.5151, -50.2485), (136.6594, 143.0302, 100.497)),
    5: (201.40358597, (-50.3958, -90.8979, -96.867), (100.7916, 181.7958, 193.734)),
    6: (154.54551808, (-33.5054, -63.8165, -58.7283), (67.0107, 127.633, 117.4566)),
    8: (261.47178102, (-114.738, -37.7357, -104.001), (229.476, 75.4714, 208.002)),
    9: (108.99920102, (-52.2146, -38.7038, -42.8485), (104.4292, 77.4076, 85.697)),
    10: (164.62758848, (-75.0923, -53.5375, -34.6207), (150.1846, 107.075, 69.2414)),
    11: (175.88933422, (-18.3605, -38.933, -86.4079), (36.7211, 77.866, 172.8158)),
    12: (145.54287471, (-50.4439, -54.2485, -45.4), (100.8878, 108.497, 90.8)),
    13: (278.07811733, (-129.113, -59.241, -70.5662), (258.226, 118.4821, 141.1324)),
    14: (282.60129399, (-101.573, -58.8763, -106.558), (203.146, 117.7525, 213.116)),
    15: (212.35825148, (-46.9591, -73.7167, -92.3737), (93.9181, 147.4334, 184.7474)),
}
N_CLS = len(OBJLIST)
DIAMETER = np.array([_MODELS_INFO_MM[o][0] / 1000.0 for o in OBJLIST], dtype=np.float32)
LFBORDER = np.array([_MODELS_INFO_MM[o][1] for o in OBJLIST], dtype=np.float64) / 1000.0
EXTENT = np.array([_MODELS_INFO_MM[o][2] for o in OBJLIST], dtype=np.float64) / 1000.0
IS_SYM = np.array([1 if c in SYM_CLS else 0 for c in range(N_CLS)], dtype=np.uint8)


def class_models(n_points: int, seed: int = 0x5EED) -> np.ndarray:
    """[C, n_points, 3] f32 model clouds, uniform in each class box."""
    rng = np.random.default_rng([seed, 0xC1A55, n_points])
    u = rng.random((N_CLS, n_points, 3))
    return (LFBORDER[:, None, :] + u * EXTENT[:, None, :]).astype(np.float32)


def random_rotation(rng) -> np.ndarray:
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    w, x, y, z = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def make_frame_sparse(seed: int, config_id: int, frame: int, cls: int | None = None,
                      p_range=(800, 1200), outlier_frac=0.25, noise_m=1e-3,
                      K4=LINEMOD_K4, H=IMG_H, W=IMG_W, outlier_min_px=0.0, z_range=(0.6, 1.2), xy_frac=1.0, dense=False):
    """One frame in sparse form.  ``dense=True`` keeps EVERY pixel of the silhouette (a compact blob like a real
    LineMOD mask, P = 2 k .. 30 k at the default distances) instead of the seeded subset of P pixels.
    Returns dict with
       rows,cols [P] int32 (row-major sorted), xyzn [P,3] f32 (normalised map values, never 0),
       cls, R_gt [3,3] f32, t_gt [3] f32, inlier [P] bool (ground truth of the injected outliers)."""
    rng = np.random.default_rng([seed, config_id, frame])
    if cls is None:
        cls = frame % N_CLS
    ext, lfb = EXTENT[cls], LFBORDER[cls]
    fx, fy, cx, cy = [float(v) for v in K4]
    R = random_rotation(rng)
    t = np.array([rng.uniform(-0.15, 0.15) * xy_frac, rng.uniform(-0.1, 0.1) * xy_frac,
                  rng.uniform(z_range[0], z_range[1])])      # close-up frames (C4) shrink z_range / xy_frac
    P = int(rng.integers(p_range[0], p_range[1] + 1))
    uc, vc = fx * t[0] / t[2] + cx, fy * t[1] / t[2] + cy
    smax = 0.49 * float(ext.max())
    rad = int(np.ceil(max(fx, fy) * smax / max(t[2] - smax, 0.2))) + 2     # silhouette bound (px)
    c0, c1 = max(int(uc) - rad, 0), min(int(uc) + rad + 1, W)
    r0, r1 = max(int(vc) - rad, 0), min(int(vc) + rad + 1, H)
    rr, cc = np.meshgrid(np.arange(r0, r1), np.arange(c0, c1), indexing='ij')
    rr, cc = rr.reshape(-1), cc.reshape(-1)
    d_cam = np.stack([(cc - cx) / fx, (rr - cy) / fy, np.ones_like(cc, dtype=np.float64)], axis=1)
    o = -R.T @ t                                   # camera centre in the object frame
    d = d_cam @ R                                  # rows: R^T d_cam
    centre, semi = lfb + 0.5 * ext, 0.49 * ext
    oo, dd = (o - centre) / semi, d / semi
    qa = (dd * dd).sum(1)
    qb = (dd * oo[None, :]).sum(1)
    qc = float(oo @ oo) - 1.0
    disc = qb * qb - qa * qc
    hit = disc >= 0
    tau = np.where(hit, (-qb - np.sqrt(np.where(hit, disc, 0.0))) / qa, 0.0)
    hit &= tau > 0
    rr, cc, tau, d = rr[hit], cc[hit], tau[hit], d[hit]
    # the mask is a seeded uniform subset of the silhouette (all of it if it is smaller than P)
    if rr.shape[0] > P and not dense:
        keep = np.sort(rng.choice(rr.shape[0], size=P, replace=False))   # row-major order kept
        rr, cc, tau, d = rr[keep], cc[keep], tau[keep], d[keep]
    P = rr.shape[0]
    Xo = o[None, :] + tau[:, None] * d             # exact surface points (object frame, metres)
    inl = rng.random(P) >= outlier_frac
    Xo = Xo + rng.normal(scale=noise_m, size=(P, 3)) * inl[:, None]
    Xout = lfb[None, :] + rng.random((P, 3)) * ext[None, :]
    if outlier_min_px > 0:                          # keep injected outliers clearly off their pixel
        for _ in range(64):
            Xc = Xout @ R.T + t
            e = np.hypot(fx * Xc[:, 0] / Xc[:, 2] + cx - cc, fy * Xc[:, 1] / Xc[:, 2] + cy - rr)
            bad = (~inl) & (e < outlier_min_px)
            if not bad.any():
                break
            Xout[bad] = lfb[None, :] + rng.random((int(bad.sum()), 3)) * ext[None, :]
    Xo = np.where(inl[:, None], Xo, Xout)
    xyzn = np.clip((Xo - lfb[None, :]) / ext[None, :], 1e-3, 1.0).astype(np.float32)
    return dict(rows=rr.astype(np.int32), cols=cc.astype(np.int32), xyzn=xyzn, cls=int(cls),
                R_gt=R.astype(np.float32), t_gt=t.astype(np.float32), inlier=inl)


def dense_from_sparse(sp, H=IMG_H, W=IMG_W):
    """-> xyz [3,H,W] f32, mask [H,W] u8."""
    xyz = np.zeros((3, H, W), np.float32)
    mask = np.zeros((H, W), np.uint8)
    xyz[:, sp['rows'], sp['cols']] = sp['xyzn'].T
    mask[sp['rows'], sp['cols']] = 1
    return xyz, mask


def reference_frame_dict(sp, model_points, H=IMG_H, W=IMG_W, K4=LINEMOD_K4):
    """The ``datas`` / ``pred`` entries the reference's get_pose / cal_dis read for one frame
    (tools/trainer.py:383-438, :370-381), built from a sparse synthetic frame."""
    xyz, mask = dense_from_sparse(sp, H, W)
    choose = (sp['rows'].astype(np.int64) * W + sp['cols']).astype(np.int64)
    cls = sp['cls']
    target = model_points @ sp['R_gt'].T + sp['t_gt'][None, :]
    return dict(xyz=xyz, mask=mask, choose=choose,
                x_map_choosed=sp['cols'].astype(np.float32), y_map_choosed=sp['rows'].astype(np.float32),
                extent=EXTENT[cls].copy(), lfborder=LFBORDER[cls].copy(),
                intrinsic=np.asarray(K4, np.float32), cls_id=cls,
                model_points=model_points.astype(np.float32), target=target.astype(np.float32),
                target_r=sp['R_gt'], target_t=sp['t_gt'])


def compact_points(sp):
    """De-normalised correspondences as the SoA [5,P] f32 the kernels use (FP32 fma like K0)."""
    cls = sp['cls']
    ext = EXTENT[cls].astype(np.float32)
    lfb = LFBORDER[cls].astype(np.float32)
    # fma(xyzn, ext, lfb) evaluated in float64 then rounded == single-rounded fmaf for f32 inputs
    obj = (sp['xyzn'].astype(np.float64) * ext.astype(np.float64) + lfb.astype(np.float64)).astype(np.float32)
    return np.stack([obj[:, 0], obj[:, 1], obj[:, 2],
                     sp['cols'].astype(np.float32), sp['rows'].astype(np.float32)], axis=0)


def make_batch_sparse(seed: int, config_id: int, n_frames: int, first_frame: int = 0, **kw):
    return [make_frame_sparse(seed, config_id, first_frame + f, **kw) for f in range(n_frames)]


# ---------------------------------------------------------------------------------------------------------------
# Config C5: frames generated ON THE DEVICE by csrc/synthgen.cu; this is its NumPy twin.
# The per-frame parameters come from the host (cheap: 48 floats per frame, a pure function of (seed, frame index),
# so every sharding of the sweep sees the same frames); the per-pixel work -- ray / ellipsoid hit, keep / outlier /
# noise draws from Philox4x32-10 -- is IEEE float32 in a written order on both sides, so the bytes are identical.
C5_PARAM_STRIDE = 48
C5_TAG = 0x47454E35
C5_PARAM_TAG = 0x50415235


def philox4x32(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10 (Random123): uint32 arrays / scalars in, four uint32 arrays out."""
    c = [np.asarray(v, np.uint64) & np.uint64(0xFFFFFFFF) for v in np.broadcast_arrays(c0, c1, c2, c3)]
    k0, k1 = np.uint64(k0 & 0xFFFFFFFF), np.uint64(k1 & 0xFFFFFFFF)
    M0, M1, MASK = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & MASK, p1 >> np.uint64(32), p1 & MASK
        c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
        k0 = (k0 + np.uint64(0x9E3779B9)) & MASK
        k1 = (k1 + np.uint64(0xBB67AE85)) & MASK
    return [v.astype(np.uint32) for v in c]


def _u01(r):
    return (r >> np.uint32(8)).astype(np.float32) * np.float32(2.0 ** -24)


def c5_params(seed: int, frame0: int, n: int, p_range=(800, 1200), outlier_frac=0.25, noise_m=1e-3,
              K4=LINEMOD_K4, H=IMG_H, W=IMG_W, z_range=(0.6, 1.2)):
    """Per-frame parameters [n,48] f32 of frames frame0 .. frame0+n (layout: csrc/synthgen.cu) plus the
    per-frame arrays the evaluation needs: dict(params, cls, R_gt, t_gt, extent, lfborder, K4, bbox_max)."""
    f = np.arange(frame0, frame0 + n, dtype=np.uint64)
    u = []
    for k in range(3):                                  # 12 uniforms per frame from counters (frame, k)
        u += [(r.astype(np.float64) + 0.5) / 2.0 ** 32 for r in philox4x32(f, k, 0, C5_PARAM_TAG, seed, seed >> 32)]
    u = np.stack(u, axis=1)
    # rotation: quaternion from four normals (Box-Muller), host float64 -- only its float32 rounding is shared
    g = np.sqrt(-2.0 * np.log(u[:, 0:2]))
    q = np.stack([g[:, 0] * np.cos(2 * np.pi * u[:, 2]), g[:, 0] * np.sin(2 * np.pi * u[:, 2]),
                  g[:, 1] * np.cos(2 * np.pi * u[:, 3]), g[:, 1] * np.sin(2 * np.pi * u[:, 3])], axis=1)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, x, y, z = q.T
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                  2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                  2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], axis=1).astype(np.float32)
    t = np.stack([(u[:, 4] * 0.3 - 0.15), (u[:, 5] * 0.2 - 0.1), z_range[0] + u[:, 6] * (z_range[1] - z_range[0])],
                 axis=1).astype(np.float32)
    P = np.floor(p_range[0] + u[:, 7] * (p_range[1] - p_range[0] + 1)).astype(np.float32)
    cls = (np.arange(frame0, frame0 + n) % N_CLS).astype(np.int32)
    ext, lfb = EXTENT[cls].astype(np.float32), LFBORDER[cls].astype(np.float32)
    fx, fy, cx, cy = [np.float32(v) for v in K4]
    R64, t64 = R.astype(np.float64).reshape(n, 3, 3), t.astype(np.float64)
    o = (-np.einsum('nji,nj->ni', R64, t64)).astype(np.float32)
    centre = (lfb.astype(np.float64) + 0.5 * ext.astype(np.float64))
    semi = (0.49 * ext.astype(np.float64)).astype(np.float32)
    oo = ((o.astype(np.float64) - centre) / semi.astype(np.float64)).astype(np.float32)
    qc = ((oo.astype(np.float64) ** 2).sum(axis=1) - 1.0).astype(np.float32)
    smax = 0.49 * EXTENT[cls].max(axis=1)
    rad = np.ceil(max(float(fx), float(fy)) * smax / np.maximum(t64[:, 2] - smax, 0.2)).astype(np.int64) + 2
    uc = (float(fx) * t64[:, 0] / t64[:, 2] + float(cx)).astype(np.int64)
    vc = (float(fy) * t64[:, 1] / t64[:, 2] + float(cy)).astype(np.int64)
    r0, r1 = np.clip(vc - rad, 0, H), np.clip(vc + rad + 1, 0, H)
    c0, c1 = np.clip(uc - rad, 0, W), np.clip(uc + rad + 1, 0, W)
    prm = np.zeros((n, C5_PARAM_STRIDE), np.float32)
    prm[:, 0:9], prm[:, 9:12] = R, t
    prm[:, 12:16] = (fx, fy, cx, cy)
    prm[:, 16:19], prm[:, 19:22], prm[:, 22:25], prm[:, 25:28], prm[:, 28:31], prm[:, 31] = o, oo, semi, lfb, ext, qc
    prm[:, 32], prm[:, 33], prm[:, 34], prm[:, 35] = r0, r1, c0, c1
    prm[:, 36], prm[:, 37], prm[:, 38] = P, outlier_frac, np.float32(noise_m * np.sqrt(3.0))
    # keep probability: P over the ANALYTIC area of the silhouette, the ellipse d^T (D u u^T D - qc D) d >= 0 of the
    # pixel's ray direction d = R^T K^-1 (col, row, 1) (within 0.5 % of the pixel count; no counting pass on the device)
    D = np.zeros((n, 3, 3)); D[:, 0, 0], D[:, 1, 1], D[:, 2, 2] = (1.0 / semi.astype(np.float64) ** 2).T
    uu = o.astype(np.float64) - centre
    Du = np.einsum('nij,nj->ni', D, uu)
    M = np.einsum('ni,nj->nij', Du, Du) - (np.einsum('ni,ni->n', uu, Du) - 1.0)[:, None, None] * D
    Kinv = np.array([[1 / float(fx), 0, -float(cx) / float(fx)], [0, 1 / float(fy), -float(cy) / float(fy)], [0, 0, 1.0]])
    A = np.einsum('nji,jk->nik', R64, Kinv)
    Q = np.einsum('nji,njk,nkl->nil', A, M, A)
    area = np.pi * np.abs(np.linalg.det(Q)) / np.abs(np.linalg.det(Q[:, :2, :2])) ** 1.5
    prm[:, 39] = (P / area).astype(np.float32)
    # tight bounding box of that ellipse from its dual conic adj(Q) (+- 2 px), inside the sphere bound above
    Cs = np.linalg.inv(Q) * np.linalg.det(Q)[:, None, None]
    with np.errstate(invalid='ignore'):
        du = np.sqrt(np.maximum(Cs[:, 0, 2] ** 2 - Cs[:, 0, 0] * Cs[:, 2, 2], 0.0)) / np.abs(Cs[:, 2, 2])
        dv = np.sqrt(np.maximum(Cs[:, 1, 2] ** 2 - Cs[:, 1, 1] * Cs[:, 2, 2], 0.0)) / np.abs(Cs[:, 2, 2])
    uc_, vc_ = Cs[:, 0, 2] / Cs[:, 2, 2], Cs[:, 1, 2] / Cs[:, 2, 2]
    ok = np.isfinite(du) & np.isfinite(dv) & np.isfinite(uc_) & np.isfinite(vc_)
    c0 = np.where(ok, np.maximum(c0, np.floor(uc_ - du).astype(np.int64) - 2), c0)
    c1 = np.where(ok, np.minimum(c1, np.ceil(uc_ + du).astype(np.int64) + 3), c1)
    r0 = np.where(ok, np.maximum(r0, np.floor(vc_ - dv).astype(np.int64) - 2), r0)
    r1 = np.where(ok, np.minimum(r1, np.ceil(vc_ + dv).astype(np.int64) + 3), r1)
    c1, r1 = np.maximum(c1, c0), np.maximum(r1, r0)
    prm[:, 32], prm[:, 33], prm[:, 34], prm[:, 35] = r0, r1, c0, c1
    prm[:, 40], prm[:, 41] = np.float32(1.0) / fx, np.float32(1.0) / fy
    prm[:, 42:45] = np.float32(1.0) / semi
    assert p_range[1] <= 2048, "the generator remembers at most 4096 written pixels per frame"
    return dict(params=prm, cls=cls, R_gt=R.reshape(n, 3, 3), t_gt=t, extent=ext, lfborder=lfb,
                K4=np.tile(np.asarray(K4, np.float32), (n, 1)), bbox_max=int(((r1 - r0) * (c1 - c0)).max()))


def c5_frame_twin(prm, seed: int, frame: int, W=IMG_W):
    """NumPy twin of csrc/synthgen.cu for ONE frame: prm [48] f32 (a row of c5_params), frame = its global index.
    Returns the sparse frame dict (rows, cols [P] int32 row-major, xyzn [P,3] f32, area = silhouette pixels)."""
    p = np.asarray(prm, np.float32)
    f32 = np.float32
    r0, r1, c0, c1 = [int(v) for v in p[32:36]]
    rr, cc = np.meshgrid(np.arange(r0, r1), np.arange(c0, c1), indexing='ij')
    rr, cc = rr.reshape(-1), cc.reshape(-1)
    dcx = (cc.astype(f32) - p[14]) * p[40]
    dcy = (rr.astype(f32) - p[15]) * p[41]
    d = [(p[k] * dcx + p[3 + k] * dcy) + p[6 + k] for k in range(3)]
    ddx, ddy, ddz = d[0] * p[42], d[1] * p[43], d[2] * p[44]
    qa = (ddx * ddx + ddy * ddy) + ddz * ddz
    qb = (ddx * p[19] + ddy * p[20]) + ddz * p[21]
    disc = qb * qb - qa * p[31]
    hit = (disc >= 0) & (qb < 0)
    with np.errstate(invalid='ignore', divide='ignore'):
        tau = (-qb - np.sqrt(np.where(hit, disc, f32(0)))) / qa
    area = int(hit.sum())
    rr, cc, tau = rr[hit], cc[hit], tau[hit]
    d = [v[hit] for v in d]
    pix = (rr * W + cc).astype(np.uint32)
    r = philox4x32(pix, frame, 0, C5_TAG, seed, seed >> 32)
    keep = _u01(r[0]) < p[39]
    rr, cc, tau, pix = rr[keep], cc[keep], tau[keep], pix[keep]
    d = [v[keep] for v in d]
    outl = _u01(r[1][keep]) < p[37]
    Xo = [p[16 + k] + tau * d[k] for k in range(3)]
    qo = philox4x32(pix, frame, 1, C5_TAG, seed, seed >> 32)
    xyzn = np.zeros((pix.shape[0], 3), f32)
    for c in range(3):
        qn = philox4x32(pix, frame, 2 + c, C5_TAG, seed, seed >> 32)
        s4 = ((_u01(qn[0]) + _u01(qn[1])) + (_u01(qn[2]) + _u01(qn[3]))) - f32(2.0)
        X = np.where(outl, p[25 + c] + _u01(qo[c]) * p[28 + c], Xo[c] + p[38] * s4).astype(f32)
        v = (X - p[25 + c]) / p[28 + c]
        xyzn[:, c] = np.minimum(np.maximum(v, f32(1.0e-3)), f32(1.0))
    return dict(rows=rr.astype(np.int32), cols=cc.astype(np.int32), xyzn=xyzn, area=area, inlier=~outl)


def c5_frames(seed: int, frame0: int, n: int, **kw):
    """Sparse frames frame0 .. frame0+n of config C5 in the format of make_frame_sparse (for the CPU legs / tests)."""
    P = c5_params(seed, frame0, n, **kw)
    out = []
    for i in range(n):
        sp = c5_frame_twin(P["params"][i], seed, frame0 + i)
        sp.update(cls=int(P["cls"][i]), R_gt=P["R_gt"][i], t_gt=P["t_gt"][i])
        out.append(sp)
    return out
