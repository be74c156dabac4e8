"""Golden vectors for the map file format from the reference's OWN save_ply / load_ply.

Run in the authoring container (needs /root/reference; CPU only):  python tests/golden/make_ply_golden.py

`scene/Gaussians.py` is loaded UNMODIFIED (as in make_optim_golden.py).  plyfile is not installed in this image, so it
is replaced by a RECORDING stand-in that keeps exactly what the reference hands to it and serves exactly what the
reference asks of it:

    save_ply (scene/Gaussians.py:452-464)   PlyElement.describe(elements, 'vertex') -> the structured array `elements`
                                            (property names, order, dtype, the row bytes) is recorded.
    load_ply (scene/Gaussians.py:466-511)   PlyData.read(path).elements[0] serves columns by name and `.properties`
                                            (objects with `.name`, in file order) from a given structured array; the
                                            selection / sorting / stacking into _xyz, _opacity, _scaling, _rotation, _rgb is
                                            the reference's code (its `device="cuda"` is redirected to the CPU).

What stays restated from plyfile's documentation (and unpinned) is only the header text plyfile writes around those
rows: "ply / format binary_little_endian 1.0 / element vertex N / property float <name> ... / end_header".

Output: tests/golden/ply_<case>.npz with in_<group> (raw parameters), ref_names (property names in order),
ref_rows (uint8: the bytes of the reference's structured array), and for the load direction file_names, file_rows (a
vertex table with shuffled and extra properties) and ref_load_<group> (what the reference's load_ply makes of it).
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = "/root/reference"
GROUPS = ("xyz", "opacity", "scaling", "rotation", "rgb")


class Recorder:
    described = None          # (structured array, element name) of the last PlyElement.describe
    served = None             # structured array PlyData.read hands out


class PlyElement:
    def __init__(self, data, name):
        self.data, self.name = data, name
        self.properties = [types.SimpleNamespace(name=n) for n in data.dtype.names]

    @staticmethod
    def describe(data, name):
        Recorder.described = (data.copy(), name)
        return PlyElement(data, name)

    def __getitem__(self, key):
        return self.data[key]


class PlyData:
    def __init__(self, elements):
        self.elements = list(elements)

    def write(self, path):                     # the bytes plyfile would write are not needed for the record
        pass

    @staticmethod
    def read(path):
        return PlyData([PlyElement(Recorder.served, "vertex")])


def import_reference_gaussians():
    def stub(name, **attrs):
        m = types.ModuleType(name)
        sys.modules[name] = m
        for k, v in attrs.items():
            setattr(m, k, v)
        return m
    stub("pytorch3d")
    sys.modules["pytorch3d"].transforms = stub("pytorch3d.transforms")
    stub("plyfile", PlyData=PlyData, PlyElement=PlyElement)
    stub("simple_knn")
    stub("simple_knn._C", distCUDA2=None)
    sys.path.insert(0, REFERENCE)
    spec = importlib.util.spec_from_file_location("reference_gaussians", os.path.join(REFERENCE, "scene", "Gaussians.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)

    class TorchOnCpu:
        """The module's `torch`, with tensor(..., device='cuda') redirected to the CPU (scene/Gaussians.py:504-510)."""
        def __getattr__(self, name):
            return getattr(torch, name)

        @staticmethod
        def tensor(*a, **kw):
            kw.pop("device", None)
            return torch.tensor(*a, **kw)
    mod.torch = TorchOnCpu()
    return mod.Gaussians


def raw_parameters(P, scale_cols, seed):
    rng = np.random.default_rng(seed)
    return dict(xyz=rng.normal(0, 2, (P, 3)).astype(np.float32), opacity=rng.normal(0, 1.5, (P, 1)).astype(np.float32),
                scaling=rng.normal(-4, 0.5, (P, scale_cols)).astype(np.float32),
                rotation=rng.normal(0, 1, (P, 4)).astype(np.float32), rgb=rng.uniform(0, 1, (P, 3)).astype(np.float32))


def new_gaussians(Gaussians, distribution):
    return Gaussians(dict(render=dict(method="2dgs"), gaussians=dict(gaussian_distribution=distribution)))


def shuffled_file(raw, seed):
    """A vertex table another tool might write: properties in a different order (scale_1 may precede scale_0),
    extra properties, one of them float64."""
    rng = np.random.default_rng(seed)
    P = raw["xyz"].shape[0]
    cols = {"x": raw["xyz"][:, 0], "y": raw["xyz"][:, 1], "z": raw["xyz"][:, 2], "opacity": raw["opacity"][:, 0],
            "r": raw["rgb"][:, 0], "g": raw["rgb"][:, 1], "b": raw["rgb"][:, 2],
            "confidence": rng.uniform(size=P), "nx": np.zeros(P, np.float32)}
    for i in range(raw["scaling"].shape[1]):
        cols[f"scale_{i}"] = raw["scaling"][:, i]
    for i in range(4):
        cols[f"rot_{i}"] = raw["rotation"][:, i]
    order = list(cols)
    rng.shuffle(order)
    dtype = [(n, "<f8" if n == "confidence" else "<f4") for n in order]
    table = np.empty(P, dtype=dtype)
    for n in order:
        table[n] = cols[n]
    return table


CASES = dict(aniso=(53, 2, "anisotropic", 0), iso=(17, 1, "isotropic", 1), empty=(0, 2, "anisotropic", 2))


def main():
    Gaussians = import_reference_gaussians()
    for name, (P, cols, dist, seed) in CASES.items():
        raw = raw_parameters(P, cols, seed)
        out = {"in_" + k: raw[k] for k in GROUPS}
        gs = new_gaussians(Gaussians, dist)
        for k in GROUPS:
            setattr(gs, "_" + k, torch.nn.Parameter(torch.tensor(raw[k])))
        gs.save_ply(os.path.join(HERE, "gaussians.ply"))        # the directory exists; the stand-in writes nothing
        elements, element_name = Recorder.described
        assert element_name == "vertex"
        out["ref_names"] = np.array(elements.dtype.names)
        out["ref_formats"] = np.array([elements.dtype[n].str for n in elements.dtype.names])
        out["ref_rows"] = np.frombuffer(elements.tobytes(), np.uint8)
        if P:
            Recorder.served = shuffled_file(raw, seed + 100)
            back = new_gaussians(Gaussians, dist).load_ply("unused")
            out["file_names"] = np.array(Recorder.served.dtype.names)
            out["file_formats"] = np.array([Recorder.served.dtype[n].str for n in Recorder.served.dtype.names])
            out["file_rows"] = np.frombuffer(Recorder.served.tobytes(), np.uint8)
            for k in GROUPS:
                out["ref_load_" + k] = getattr(back, "_" + k).detach().numpy()
        np.savez_compressed(os.path.join(HERE, f"ply_{name}.npz"), **out)
        print(name, list(out["ref_names"]), out["ref_rows"].size, "bytes")


if __name__ == "__main__":
    main()
