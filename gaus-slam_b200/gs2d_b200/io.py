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


# ---- LocalMap queue payload (scene/Frame.py:202-235, slam/Frontend.py:201-204) -----------------
# The reference's frontend hands a finished submap to the backend as `copy.deepcopy(LocalMap)` through an mp.Queue:
# a pickle of Python objects whose bulk is `local_map_params = Gaussians.extract_params()` (scene/Gaussians.py:349-360:
# the five RAW parameter tensors xyz, opacity, scaling, rotation, rgb, detached) plus one 4x4 pose per frame.  When
# frontend and backend sit on different GPUs (or processes without CUDA IPC) what has to cross is exactly that, so the
# wire format is one flat little-endian buffer: a 64-byte header, the [P, 3+1+S+4+3] float32 parameter table in the
# column order of gaussians.ply without the (all-zero) normals, the [F,4,4] float32 world-to-camera estimates of the
# submap's frames and their int64 time indices.  Packing is one concatenation on the device and one device-to-host
# copy; the receiver gets the dict `Gaussians.create_params` / `add_params` take (scene/Gaussians.py:362-392).
# Not carried: what the backend recomputes or does not read from the payload (gt images are already in host memory on
# the reference's side of the queue, the NetVLAD descriptor is out of scope).
LOCAL_MAP_MAGIC = b"GS2DLMAP"
LOCAL_MAP_VERSION = 1
_LM_HEADER = np.dtype([("magic", "S8"), ("version", "<u4"), ("lmid", "<i4"), ("tracking_ok", "<u4"), ("P", "<u8"),
                       ("scale_cols", "<u4"), ("frames", "<u4"), ("reserved", "V28")])
assert _LM_HEADER.itemsize == 64
LOCAL_MAP_KEYS = ("xyz", "opacity", "scaling", "rotation", "rgb")      # Gaussians.extract_params, use_sh = False


def pack_local_map(lmid, params, frame_w2cs, frame_time_idx=None, tracking_ok=True):
    """-> numpy uint8 buffer.  params: the dict of Gaussians.extract_params() (torch tensors on any device, or arrays);
    frame_w2cs: [F,4,4] (or a list of 4x4) estimated world-to-camera matrices of the submap's frames."""
    missing = [k for k in LOCAL_MAP_KEYS if k not in params]
    if missing:
        raise KeyError(f"local map parameters lack {missing} (f_dc / f_rest maps are not carried: use_sh is False in "
                       "every shipped config)")
    cols = [params[k] for k in LOCAL_MAP_KEYS]
    if hasattr(cols[0], "detach"):
        import torch
        with torch.no_grad():
            table = torch.cat([c.detach().to(torch.float32).reshape(c.shape[0], c.shape[1]) for c in cols], dim=1).cpu().numpy()
    else:
        table = np.concatenate([np.asarray(c, np.float32).reshape(np.shape(c)[0], np.shape(c)[1]) for c in cols], axis=1)
    S = int(np.shape(params["scaling"])[1])
    if table.shape[1] != 3 + 1 + S + 4 + 3:
        raise ValueError(f"expected {3 + 1 + S + 4 + 3} parameter columns, got {table.shape[1]}")
    w2c = np.stack([_np(w) for w in frame_w2cs], axis=0).astype("<f4") if not isinstance(frame_w2cs, np.ndarray) \
        else frame_w2cs.astype("<f4")
    if w2c.ndim != 3 or w2c.shape[1:] != (4, 4):
        raise ValueError("frame_w2cs must be [F,4,4]")
    t_idx = np.arange(w2c.shape[0], dtype="<i8") if frame_time_idx is None else np.asarray(frame_time_idx, "<i8")
    if t_idx.shape != (w2c.shape[0],):
        raise ValueError("one time index per frame")
    hdr = np.zeros(1, _LM_HEADER)
    hdr["magic"], hdr["version"], hdr["lmid"], hdr["tracking_ok"] = LOCAL_MAP_MAGIC, LOCAL_MAP_VERSION, int(lmid), int(bool(tracking_ok))
    hdr["P"], hdr["scale_cols"], hdr["frames"] = table.shape[0], S, w2c.shape[0]
    return np.concatenate([hdr.view(np.uint8).ravel(), np.ascontiguousarray(table, "<f4").view(np.uint8).ravel(),
                           np.ascontiguousarray(w2c).view(np.uint8).ravel(), t_idx.view(np.uint8).ravel()])


def unpack_local_map(buf, device=None):
    """-> dict(lmid, tracking_ok, params={xyz, opacity, scaling, rotation, rgb}, frame_w2cs [F,4,4], frame_time_idx [F]).
    With `device` the five parameter tensors are torch tensors there (one host-to-device copy of the table, then views:
    the columns of one submap stay one allocation); without it they are numpy arrays."""
    buf = np.frombuffer(buf, np.uint8) if not isinstance(buf, np.ndarray) else buf.view(np.uint8).ravel()
    if buf.size < 64:
        raise ValueError("not a local-map buffer (shorter than its header)")
    hdr = buf[:64].view(_LM_HEADER)[0]
    if bytes(hdr["magic"]) != LOCAL_MAP_MAGIC or int(hdr["version"]) != LOCAL_MAP_VERSION:
        raise ValueError("not a local-map buffer of this version")
    P, S, F = int(hdr["P"]), int(hdr["scale_cols"]), int(hdr["frames"])
    K = 3 + 1 + S + 4 + 3
    need = 64 + 4 * P * K + 64 * F + 8 * F
    if buf.size != need:
        raise ValueError(f"local-map buffer has {buf.size} bytes, its header describes {need}")
    o = 64
    table = buf[o:o + 4 * P * K].view("<f4").reshape(P, K); o += 4 * P * K
    w2c = buf[o:o + 64 * F].view("<f4").reshape(F, 4, 4).copy(); o += 64 * F
    t_idx = buf[o:o + 8 * F].view("<i8").copy()
    edges = np.cumsum([0, 3, 1, S, 4, 3])
    if device is not None:
        import torch
        t = torch.from_numpy(np.ascontiguousarray(table)).to(device)
        params = {k: t[:, edges[i]:edges[i + 1]] for i, k in enumerate(LOCAL_MAP_KEYS)}
    else:
        params = {k: np.ascontiguousarray(table[:, edges[i]:edges[i + 1]]) for i, k in enumerate(LOCAL_MAP_KEYS)}
    return dict(lmid=int(hdr["lmid"]), tracking_ok=bool(hdr["tracking_ok"]), params=params, frame_w2cs=w2c,
                frame_time_idx=t_idx)
