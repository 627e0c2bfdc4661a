"""Model / fixture ingestion for the evaluation path.

Host-side readers for the files the reference's dataset class loads before the metric can run, plus the one
numeric step among them (farthest point sampling of the model cloud), which runs on the GPU (``pb_fps``):

* ``ply_vtx``            tools/script/sample_model.py:19-30 (vertex block of an ASCII ``.ply``; binary
                         little-endian files are read too)
* ``load_models_info``   dataset/linemod/batchdataset.py:101-103,163-165 (``models_info.yml``: diameter,
                         min_*, size_* in millimetres -> metres)
* ``sample_model``       tools/script/sample_model.py:78-85 (5000 farthest-point samples of every model,
                         written as ``obj_XX.pkl``)
* ``ModelBank``          batchdataset.py:111-165,696-704: per-class model clouds, extents, left borders,
                         diameters and the symmetric flags as the dense arrays ``EvalSweep`` takes

No numeric fallback lives here: sampling calls the CUDA library and raises without it.
"""
from __future__ import annotations

import os
import pickle
from dataclasses import dataclass

import numpy as np

OBJLIST = [1, 2, 4, 5, 6, 8, 9, 10, 11, 12, 13, 14, 15]      # batchdataset.py:42
SYM_CLS = [7, 8]                                             # indices into OBJLIST (eggbox, glue), batchdataset.py:76

_PLY_TYPES = {"char": "b", "int8": "b", "uchar": "B", "uint8": "B", "short": "h", "int16": "h", "ushort": "H",
              "uint16": "H", "int": "i", "int32": "i", "uint": "I", "uint32": "I", "float": "f", "float32": "f",
              "double": "d", "float64": "d"}


def ply_vtx(path) -> np.ndarray:
    """Vertex coordinates [N,3] float32 of a ``.ply`` model (tools/script/sample_model.py:19-30 reads the first
    three columns of an ASCII vertex block; the header is parsed properly here, so property order, comments
    and binary little-endian bodies are handled as well)."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a ply file")
        fmt, n_vtx, props, in_vertex, seen_vertex = None, None, [], False, False
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: no end_header")
            tok = line.decode("ascii", "replace").split()
            if not tok:
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_vertex = tok[1] == "vertex"
                if in_vertex:
                    n_vtx, seen_vertex = int(tok[2]), True
                elif not seen_vertex:
                    raise ValueError(f"{path}: an element precedes the vertex block")
            elif tok[0] == "property" and in_vertex:
                if tok[1] == "list":
                    raise ValueError(f"{path}: list property in the vertex block")
                props.append((tok[-1], tok[1]))
            elif tok[0] == "end_header":
                break
        if n_vtx is None or fmt is None:
            raise ValueError(f"{path}: header lacks format / element vertex")
        names = [p[0] for p in props]
        try:
            cols = [names.index(c) for c in ("x", "y", "z")]
        except ValueError:
            cols = [0, 1, 2]                                  # the reference takes the first three columns
        if fmt == "ascii":
            pts = np.empty((n_vtx, 3), dtype=np.float32)
            for i in range(n_vtx):
                t = f.readline().split()
                pts[i] = [np.float32(t[c]) for c in cols]
            return pts
        if fmt != "binary_little_endian":
            raise ValueError(f"{path}: unsupported ply format {fmt}")
        dt = np.dtype([(n, "<" + np.dtype(_PLY_TYPES[t]).str[1:]) for n, t in props])
        body = np.frombuffer(f.read(n_vtx * dt.itemsize), dtype=dt, count=n_vtx)
        return np.stack([body[names[c]].astype(np.float32) for c in cols], axis=1)


def load_models_info(path) -> dict:
    """``models_info.yml`` -> {obj_id: dict(diameter, lfborder[3], extent[3])} in METRES
    (batchdataset.py:101-103 reads it, :163-165 divides by 1000)."""
    import yaml
    with open(path, "r") as f:
        info = yaml.load(f, Loader=yaml.FullLoader)
    out = {}
    for k, v in info.items():
        out[int(k)] = dict(diameter=float(v["diameter"]) / 1000.0,
                           lfborder=np.asarray([v["min_x"], v["min_y"], v["min_z"]], np.float64) / 1000.0,
                           extent=np.asarray([v["size_x"], v["size_y"], v["size_z"]], np.float64) / 1000.0)
    return out


def sample_model(points, n_samples=5000, device="cuda") -> np.ndarray:
    """tools/script/sample_model.py:33-46,80-82: the ``n_samples`` farthest-point samples of a model cloud, in
    selection order (first sample = point 0), computed by ``pb_fps``.  points [N,3] -> [n_samples,3]."""
    import torch
    from . import ops_sim3
    pts = np.ascontiguousarray(np.asarray(points, dtype=np.float32))
    idx = ops_sim3.farthest_point_sampling(torch.from_numpy(pts).to(device), int(n_samples))
    return pts[idx.cpu().numpy()]


def sample_models(root, objlist=OBJLIST, n_samples=5000, device="cuda"):
    """``main()`` of tools/script/sample_model.py:78-85: ``models/obj_XX.ply`` -> ``models/obj_XX.pkl``."""
    for obj_id in objlist:
        mdl = ply_vtx(os.path.join(root, "models", f"obj_{obj_id:02d}.ply"))
        with open(os.path.join(root, "models", f"obj_{obj_id:02d}.pkl"), "wb") as fp:
            pickle.dump(sample_model(mdl, n_samples, device), fp)


@dataclass
class ModelBank:
    """Per-class arrays in ``objlist`` order: model [C,N,3] f32 (metres), extent / lfborder [C,3] f64,
    diameter [C] f32, is_sym [C] u8 -- what ``EvalSweep`` and ``pb_pose_metrics`` take."""
    objlist: list
    model: np.ndarray
    extent: np.ndarray
    lfborder: np.ndarray
    diameter: np.ndarray
    is_sym: np.ndarray

    @classmethod
    def from_arrays(cls, clouds_m: dict, info: dict, objlist=OBJLIST, sym_cls=SYM_CLS, num_points=2600, seed=0):
        """clouds_m: {obj_id: [Ni,3] points in metres}.  Every class is cut to ``num_points`` points by a seeded
        choice without replacement (batchdataset.py:696-700 deletes a random.sample of the surplus indices);
        a class with fewer points is wrap-padded."""
        rng = np.random.default_rng(seed)
        models = []
        for o in objlist:
            p = np.asarray(clouds_m[o], dtype=np.float32)
            if p.shape[0] >= num_points:
                keep = np.sort(rng.choice(p.shape[0], num_points, replace=False))
            else:
                keep = np.resize(np.arange(p.shape[0]), num_points)
            models.append(p[keep])
        return cls(list(objlist), np.stack(models),
                   np.stack([info[o]["extent"] for o in objlist]), np.stack([info[o]["lfborder"] for o in objlist]),
                   np.asarray([info[o]["diameter"] for o in objlist], np.float32),
                   np.asarray([1 if c in sym_cls else 0 for c in range(len(objlist))], np.uint8))

    @classmethod
    def from_directory(cls, root, info_path, objlist=OBJLIST, sym_cls=SYM_CLS, num_points=2600, seed=0,
                       n_samples=5000, device="cuda"):
        """batchdataset.py:111-165: ``models/obj_XX.pkl`` (millimetres) per class; a missing pickle is made from
        the ``.ply`` by ``sample_model`` first."""
        info = load_models_info(info_path)
        clouds = {}
        for o in objlist:
            pkl = os.path.join(root, "models", f"obj_{o:02d}.pkl")
            if os.path.exists(pkl):
                with open(pkl, "rb") as f:
                    mdl = np.asarray(pickle.load(f), dtype=np.float32)
            else:
                mdl = sample_model(ply_vtx(os.path.join(root, "models", f"obj_{o:02d}.ply")), n_samples, device)
            clouds[o] = mdl / 1000.0
        return cls.from_arrays(clouds, info, objlist, sym_cls, num_points, seed)
