"""Map formats: gaussians.ply in the reference's attribute order and w2cs.npz.npy.
plyfile is not installed here, so the file is checked byte for byte against the layout plyfile's
PlyData([el]).write produces for the reference's structured array (known-answer header + packed float32 rows), and
read back with the selection rules of Gaussians.load_ply."""
import os

import numpy as np
import pytest
import torch

from gs2d_b200 import io as gio


def _params(P, S, seed):
    rng = np.random.default_rng(seed)
    return dict(xyz=rng.normal(0, 2, (P, 3)).astype(np.float32), opacity=rng.normal(0, 1, (P, 1)).astype(np.float32),
                scaling=rng.normal(-4, 1, (P, S)).astype(np.float32), rotation=rng.normal(0, 1, (P, 4)).astype(np.float32),
                rgb=rng.uniform(0, 1, (P, 3)).astype(np.float32))


def test_attribute_order_matches_the_reference():
    # scene/Gaussians.py:435-450 with use_sh False
    assert gio.ply_attributes(2) == ["x", "y", "z", "nx", "ny", "nz", "opacity", "scale_0", "scale_1", "rot_0", "rot_1",
                                     "rot_2", "rot_3", "r", "g", "b"]
    assert gio.ply_attributes(1)[6:9] == ["opacity", "scale_0", "rot_0"]


@pytest.mark.parametrize("P,S,as_torch", [(5, 2, True), (1000, 2, False), (17, 1, True), (0, 2, False)])
def test_ply_bytes_and_round_trip(tmp_path, P, S, as_torch):
    p = _params(P, S, P + S)
    args = {k: (torch.tensor(v) if as_torch else v) for k, v in p.items()}
    path = os.path.join(tmp_path, "sub", "gaussians.ply")
    gio.save_ply(path, **args)
    raw = open(path, "rb").read()
    names = gio.ply_attributes(S)
    header = ("ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P +
              "".join("property float %s\n" % n for n in names) + "end_header\n").encode()
    assert raw.startswith(header)
    body = np.frombuffer(raw[len(header):], dtype="<f4").reshape(P, len(names))
    want = np.concatenate([p["xyz"], np.zeros_like(p["xyz"]), p["opacity"], p["scaling"], p["rotation"], p["rgb"]], axis=1)
    assert body.tobytes() == want.astype("<f4").tobytes()
    back = gio.load_ply(path)
    for k in p:
        assert back[k].dtype == np.float32 and back[k].shape == p[k].shape and np.array_equal(back[k], p[k])


def test_reader_accepts_ascii_big_endian_and_shuffled_properties(tmp_path):
    p = _params(6, 2, 3)
    names = ["rot_3", "x", "y", "z", "opacity", "scale_1", "scale_0", "rot_0", "rot_1", "rot_2", "b", "g", "r", "extra"]
    cols = dict(x=p["xyz"][:, 0], y=p["xyz"][:, 1], z=p["xyz"][:, 2], opacity=p["opacity"][:, 0], scale_0=p["scaling"][:, 0],
                scale_1=p["scaling"][:, 1], rot_0=p["rotation"][:, 0], rot_1=p["rotation"][:, 1], rot_2=p["rotation"][:, 2],
                rot_3=p["rotation"][:, 3], r=p["rgb"][:, 0], g=p["rgb"][:, 1], b=p["rgb"][:, 2], extra=np.arange(6, dtype=np.float32))
    table = np.stack([cols[n] for n in names], axis=1)
    hdr = "ply\nformat %s 1.0\ncomment made by a test\nelement vertex 6\n" + "".join("property float %s\n" % n for n in names) + \
        "element face 0\nproperty list uchar int vertex_indices\nend_header\n"
    a = os.path.join(tmp_path, "a.ply")
    with open(a, "w") as f:
        f.write(hdr % "ascii")
        for row in table:
            f.write(" ".join(repr(float(x)) for x in row) + "\n")
    b = os.path.join(tmp_path, "b.ply")
    with open(b, "wb") as f:
        f.write((hdr % "binary_big_endian").encode())
        f.write(table.astype(">f4").tobytes())
    for path in (a, b):
        back = gio.load_ply(path)
        for k in p:
            assert np.array_equal(back[k], p[k]), (path, k)


def test_w2cs_file_name_and_content(tmp_path):
    w = [torch.eye(4) * (i + 1) for i in range(3)]
    path = gio.save_w2cs(str(tmp_path), w)
    assert os.path.basename(path) == "w2cs.npz.npy" and os.path.exists(path)      # scene/__init__.py:33
    back = gio.load_w2cs(str(tmp_path))
    assert back.shape == (3, 4, 4) and back.dtype == np.float32 and np.array_equal(back[2], 3 * np.eye(4, dtype=np.float32))


# ---- golden: what the reference's own save_ply hands to plyfile / load_ply makes of a file (make_ply_golden.py) ----
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GROUPS = ("xyz", "opacity", "scaling", "rotation", "rgb")
_PLY_NAME = {"<f4": "float", "<f8": "double"}


def _golden(name):
    path = os.path.join(GOLDEN, f"ply_{name}.npz")
    assert os.path.exists(path), f"{path} missing (tests/golden/make_ply_golden.py)"
    return dict(np.load(path))


@pytest.mark.parametrize("name", ["aniso", "iso", "empty"])
def test_save_ply_writes_the_rows_the_reference_hands_to_plyfile(tmp_path, name):
    g = _golden(name)
    path = os.path.join(tmp_path, "gaussians.ply")
    gio.save_ply(path, **{k: torch.tensor(g["in_" + k]) for k in GROUPS})
    raw = open(path, "rb").read()
    end = raw.index(b"end_header\n") + len(b"end_header\n")
    props = [line.split() for line in raw[:end].decode().splitlines() if line.startswith("property")]
    assert [p[2] for p in props] == [str(n) for n in g["ref_names"]]                  # names, in the reference's order
    assert [p[1] for p in props] == [_PLY_NAME[str(f)] for f in g["ref_formats"]]     # every property is float32
    assert raw[end:] == g["ref_rows"].tobytes()                                       # the reference's structured array
    assert f"element vertex {g['in_xyz'].shape[0]}\n".encode() in raw[:end]


@pytest.mark.parametrize("name", ["aniso", "iso"])
def test_load_ply_selects_what_the_reference_load_ply_selects(tmp_path, name):
    """A file with shuffled, extra and float64 properties: same arrays as Gaussians.load_ply builds from it."""
    g = _golden(name)
    names, formats = [str(n) for n in g["file_names"]], [str(f) for f in g["file_formats"]]
    P = g["in_xyz"].shape[0]
    path = os.path.join(tmp_path, "other_tool.ply")
    with open(path, "wb") as f:
        f.write(("ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P +
                 "".join("property %s %s\n" % (_PLY_NAME[t], n) for n, t in zip(names, formats)) + "end_header\n").encode())
        f.write(g["file_rows"].tobytes())
    back = gio.load_ply(path)
    for k in GROUPS:
        want = g["ref_load_" + k]
        assert back[k].dtype == want.dtype == np.float32 and back[k].shape == want.shape, k
        assert np.array_equal(back[k], want), k
        assert np.array_equal(back[k], g["in_" + k]), k


# ---- LocalMap queue payload ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("P,S,as_torch", [(0, 2, False), (1, 2, True), (257, 1, True), (1000, 2, False)])
def test_local_map_payload_round_trip(P, S, as_torch):
    """What crosses the frontend -> backend queue (scene/Frame.py:202-235): the five raw parameter tensors of
    Gaussians.extract_params (scene/Gaussians.py:349-360) + one pose per frame, bit for bit through the wire format."""
    from gs2d_b200 import io as gio
    rng = np.random.default_rng(P + S)
    params = dict(xyz=rng.normal(size=(P, 3)), opacity=rng.normal(size=(P, 1)), scaling=rng.normal(size=(P, S)),
                  rotation=rng.normal(size=(P, 4)), rgb=rng.uniform(size=(P, 3)))
    params = {k: v.astype(np.float32) for k, v in params.items()}
    assert tuple(params) == gio.LOCAL_MAP_KEYS                      # the keys create_params / add_params read (:362-392)
    w2cs = rng.normal(size=(5, 4, 4)).astype(np.float32)
    tix = np.array([3, 10, 17, 24, 31])
    src = {k: torch.from_numpy(v) for k, v in params.items()} if as_torch else params
    buf = gio.pack_local_map(7, src, list(w2cs) if as_torch else w2cs, tix, tracking_ok=False)
    assert buf.dtype == np.uint8 and buf.size == 64 + 4 * P * (11 + S) + 64 * 5 + 8 * 5
    out = gio.unpack_local_map(buf.tobytes())
    assert out["lmid"] == 7 and out["tracking_ok"] is False
    assert np.array_equal(out["frame_w2cs"], w2cs) and np.array_equal(out["frame_time_idx"], tix)
    for k in gio.LOCAL_MAP_KEYS:
        assert out["params"][k].dtype == np.float32 and np.array_equal(out["params"][k], params[k]), k
    t = gio.unpack_local_map(buf, device="cpu")["params"]
    for k in gio.LOCAL_MAP_KEYS:
        assert torch.equal(t[k], torch.from_numpy(params[k])), k


def test_local_map_payload_rejects_what_it_cannot_carry():
    from gs2d_b200 import io as gio
    p = dict(xyz=np.zeros((2, 3), np.float32), opacity=np.zeros((2, 1), np.float32), scaling=np.zeros((2, 2), np.float32),
             rotation=np.zeros((2, 4), np.float32), rgb=np.zeros((2, 3), np.float32))
    buf = gio.pack_local_map(0, p, np.zeros((1, 4, 4), np.float32))
    with pytest.raises(ValueError):
        gio.unpack_local_map(buf[:-1])
    bad = buf.copy(); bad[0] ^= 1
    with pytest.raises(ValueError):
        gio.unpack_local_map(bad)
    with pytest.raises(KeyError):
        gio.pack_local_map(0, {k: v for k, v in p.items() if k != "rgb"}, np.zeros((1, 4, 4), np.float32))
    with pytest.raises(ValueError):
        gio.pack_local_map(0, p, np.zeros((1, 3, 4), np.float32))
