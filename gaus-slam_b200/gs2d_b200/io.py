"""On-disk formats of the map, so that maps optimised with this rasterizer load in the
reference's evaluation / visualisation scripts and vice versa.

  gaussians.ply   Gaussians.save_ply / load_ply (scene/Gaussians.py:435-511): one binary little-endian `vertex`
                  element of float32 properties in the order
                      x y z nx ny nz opacity scale_0 [scale_1] rot_0 rot_1 rot_2 rot_3 r g b
                  holding the RAW optimiser parameters (logit opacity, log scales, unnormalised quaternion, raw rgb);
                  normals are written as zeros.  The reference writes through plyfile's PlyData([el]).write(path),
                  i.e. "format binary_little_endian 1.0", no comments, `property float <name>` lines.
  w2cs.npz.npy    save_scence (scene/__init__.py:19-28): np.save of the stacked [N,4,4] float32 world-to-camera
                  matrices to the path 'w2cs.npz' (np.save appends '.npy'; load_scence reads 'w2cs.npz.npy').

The reference fills its structured array through `list(map(tuple, attributes))` (a Python loop over surfels); here
the [P,K] matrix is packed on the device with one concatenation, crosses PCIe once and is written with one
`tofile`.  plyfile is not installed in this image, so the reader/writer are written against the PLY specification
and plyfile's documented header; the rows written and the arrays read back are pinned by tests/golden/ply_*.npz
(the reference's own save_ply / load_ply run against a recording stand-in for plyfile), the header text is not.
"""
from __future__ import annotations

import os

import numpy as np


def ply_attributes(scale_cols: int = 2):
    """construct_list_of_attributes (scene/Gaussians.py:435-450), use_sh = False."""
    return (["x", "y", "z", "nx", "ny", "nz", "opacity"] + [f"scale_{i}" for i in range(scale_cols)] +
            [f"rot_{i}" for i in range(4)] + list("rgb"))


def _np(t):
    return t.detach().cpu().numpy() if hasattr(t, "detach") else np.asarray(t)


def save_ply(path, xyz, opacity, scaling, rotation, rgb):
    """Raw parameters [P,3], [P,1], [P,S], [P,4], [P,3] (torch tensors on any device, or arrays)."""
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    if hasattr(xyz, "detach"):
        import torch
        with torch.no_grad():
            cols = [xyz, torch.zeros_like(xyz), opacity, scaling, rotation, rgb]
            table = torch.cat([c.detach().to(torch.float32) for c in cols], dim=1).cpu().numpy()
    else:
        xyz = np.asarray(xyz, np.float32)
        table = np.concatenate([np.asarray(c, np.float32) for c in (xyz, np.zeros_like(xyz), opacity, scaling, rotation, rgb)], axis=1)
    names = ply_attributes(int(np.shape(scaling)[1]))
    if table.shape[1] != len(names):
        raise ValueError(f"expected {len(names)} columns, got {table.shape[1]}")
    header = "ply\nformat binary_little_endian 1.0\n" + f"element vertex {table.shape[0]}\n" + \
        "".join(f"property float {n}\n" for n in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        np.ascontiguousarray(table, dtype="<f4").tofile(f)


_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2", "ushort": "u2",
              "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4", "float": "f4", "float32": "f4",
              "double": "f8", "float64": "f8"}


def read_ply_vertices(path):
    """The first element of a PLY file as a numpy structured array (binary little/big endian or ascii; scalar
    properties only -- what Gaussians.load_ply reads through plydata.elements[0])."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("not a PLY file")
        fmt, count, props, seen_element = None, None, [], 0
        while True:
            line = f.readline()
            if not line:
                raise ValueError("PLY header without end_header")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] in ("comment", "obj_info"):
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                seen_element += 1
                if seen_element == 1:
                    count = int(tok[2])
            elif tok[0] == "property" and seen_element == 1:
                if tok[1] == "list":
                    raise ValueError("list properties are not part of the map format")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt == "ascii":
            rows = np.loadtxt(f, max_rows=count, ndmin=2)
            out = np.empty(count, dtype=[(n, "<" + t) for n, t in props])
            for k, (n, _) in enumerate(props):
                out[n] = rows[:, k]
            return out
        order = "<" if fmt == "binary_little_endian" else ">"
        return np.fromfile(f, dtype=[(n, order + t) for n, t in props], count=count)


def load_ply(path):
    """-> dict of float32 arrays xyz [P,3], opacity [P,1], scaling [P,S], rotation [P,4], rgb [P,3]: the selection
    and ordering rules of Gaussians.load_ply (scene/Gaussians.py:467-509; scale_* / rot* sorted by their index)."""
    v = read_ply_vertices(path)
    names = v.dtype.names
    scale_names = sorted([n for n in names if n.startswith("scale_")], key=lambda x: int(x.split("_")[-1]))
    rot_names = sorted([n for n in names if n.startswith("rot")], key=lambda x: int(x.split("_")[-1]))
    f32 = np.float32
    return dict(xyz=np.stack([v["x"], v["y"], v["z"]], axis=1).astype(f32),
                opacity=np.asarray(v["opacity"], f32)[..., np.newaxis],
                scaling=np.stack([v[n] for n in scale_names], axis=1).astype(f32),
                rotation=np.stack([v[n] for n in rot_names], axis=1).astype(f32),
                rgb=np.stack([v["r"], v["g"], v["b"]], axis=1).astype(f32))


def save_w2cs(directory, w2cs, name="w2cs.npz"):
    """np.save(os.path.join(path, 'w2cs.npz'), stacked) -> the file 'w2cs.npz.npy' (scene/__init__.py:19-28)."""
    os.makedirs(directory, exist_ok=True)
    arr = np.stack([_np(w) for w in w2cs], axis=0) if not isinstance(w2cs, np.ndarray) else w2cs
    np.save(os.path.join(directory, name), arr)
    return os.path.join(directory, name + ".npy")


def load_w2cs(directory, name="w2cs.npz"):
    """load_scence reads 'w2cs.npz.npy' (scene/__init__.py:31-45)."""
    return np.load(os.path.join(directory, name + ".npy"))
