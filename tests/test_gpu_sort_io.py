"""On a GPU: the hand-written onesweep radix sort at rasterizer sizes (the north star names it; it replaces
cub::DeviceRadixSort::SortPairs, CR/rasterizer_impl.cu:309-314) against torch.sort(stable=True), with its throughput
printed beside CUB's on the same box; and the device-side packing of the map formats (io.save_ply / load_ply)."""
import ctypes as C
import json
import os

import numpy as np
import pytest
import torch


def _keys(n, end_bit, mode, seed, dev):
    g = torch.Generator(device="cpu").manual_seed(seed)
    if mode == "dups":        # heavy duplicates: 97 distinct keys -> stability is what orders the values
        keys = torch.randint(0, 97, (n,), generator=g, dtype=torch.int64) << (end_bit - 8)
    elif mode == "tiles":     # rasterizer-shaped: tile id above 32 depth bits, depths clustered
        tiles = torch.randint(0, 8030 if end_bit == 45 else 3225, (n,), generator=g, dtype=torch.int64)
        depth = (torch.rand(n, generator=g) * 3 + 0.5).to(torch.float32).view(torch.int32).to(torch.int64)
        keys = (tiles << 32) | depth
    else:
        keys = torch.randint(0, 1 << 62, (n,), generator=g, dtype=torch.int64) & ((1 << end_bit) - 1)
    vals = torch.arange(n, dtype=torch.int32)
    return keys.to(dev), vals.to(dev)


def _sort(lib, L, keys, vals, end_bit, scratch=None):
    n = keys.numel()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    if scratch is None:
        scratch = torch.empty((lib.gs2d_sort_scratch_bytes(n),), dtype=torch.uint8, device=keys.device)
    L.check(lib.gs2d_sort_pairs(n, keys.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(), end_bit,
                                scratch.data_ptr(), None), "gs2d_sort_pairs")
    return ko, vo


@pytest.mark.gpu
@pytest.mark.parametrize("n,end_bit,mode", [(1_000_003, 44, "tiles"), (1_000_003, 45, "dups"), (16_000_000, 44, "random"),
                                            (16_000_000, 45, "tiles"), (16_000_000, 44, "dups"), (2049, 44, "random"),
                                            (1, 45, "random")])
def test_onesweep_equals_the_stable_sort(n, end_bit, mode):
    from gs2d_b200 import _lib as L
    lib = L.load()
    dev = torch.device("cuda:0")
    keys, vals = _keys(n, end_bit, mode, n % 1000 + end_bit, dev)
    ko, vo = _sort(lib, L, keys, vals, end_bit)
    torch.cuda.synchronize()
    sk, order = torch.sort(keys, stable=True)
    assert torch.equal(ko, sk), "keys differ from the stable sort"
    assert torch.equal(vo, vals[order]), "values differ from the stable sort (stability)"
    # twice in a row on the same scratch (the look-back status words must be reset by the call itself)
    scratch = torch.empty((lib.gs2d_sort_scratch_bytes(n),), dtype=torch.uint8, device=dev)
    for _ in range(2):
        ko2, vo2 = _sort(lib, L, keys, vals, end_bit, scratch)
    torch.cuda.synchronize()
    assert torch.equal(ko2, sk) and torch.equal(vo2, vals[order])


@pytest.mark.gpu
def test_onesweep_throughput_beside_cub(capsys):
    """Not a pass/fail bar on speed (a busy box would make it flaky): measures, prints and records GB/s against the
    24 n bytes a pass moves (8 + 4 read, 8 + 4 written), with torch.sort (CUB's radix sort) timed on the same keys."""
    from gs2d_b200 import _lib as L
    lib = L.load()
    dev = torch.device("cuda:0")
    out = {}
    for n, end_bit in ((1_000_000, 44), (16_000_000, 44), (16_000_000, 45)):
        keys, vals = _keys(n, end_bit, "tiles", 7, dev)
        scratch = torch.empty((lib.gs2d_sort_scratch_bytes(n),), dtype=torch.uint8, device=dev)
        ko, vo = torch.empty_like(keys), torch.empty_like(vals)

        def ours():
            L.check(lib.gs2d_sort_pairs(n, keys.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(), end_bit,
                                        scratch.data_ptr(), None), "gs2d_sort_pairs")

        def cub():
            torch.sort(keys, stable=True)

        res = {}
        for name, fn in (("onesweep", ours), ("torch_sort_cub", cub)):
            for _ in range(3):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(10):
                fn()
            e1.record()
            torch.cuda.synchronize()
            res[name + "_ms"] = e0.elapsed_time(e1) / 10
        passes = (end_bit + 7) // 8
        res["passes"] = passes
        res["onesweep_gbs_per_pass"] = 24.0 * n * passes / (res["onesweep_ms"] * 1e-3) / 1e9
        out[f"n={n},bits={end_bit}"] = res
    os.makedirs("gpurun_out", exist_ok=True)
    with open(os.path.join("gpurun_out", "sort_throughput.json"), "w") as f:
        json.dump(out, f, indent=1)
    with capsys.disabled():
        print("\nonesweep vs torch.sort:", json.dumps(out))
    assert all(v["onesweep_ms"] > 0 for v in out.values())


@pytest.mark.gpu
def test_ply_round_trip_from_device_tensors(tmp_path):
    from gs2d_b200 import io as gio
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(5)
    P = 100_003
    p = dict(xyz=torch.randn(P, 3, generator=g), opacity=torch.randn(P, 1, generator=g),
             scaling=torch.randn(P, 2, generator=g) - 4, rotation=torch.randn(P, 4, generator=g),
             rgb=torch.rand(P, 3, generator=g))
    path = os.path.join(tmp_path, "gaussians.ply")
    gio.save_ply(path, **{k: v.to(dev) for k, v in p.items()})
    path_cpu = os.path.join(tmp_path, "gaussians_cpu.ply")
    gio.save_ply(path_cpu, **p)
    assert open(path, "rb").read() == open(path_cpu, "rb").read(), "device-side packing differs from the host path"
    back = gio.load_ply(path)
    for k in p:
        got = back[k] if isinstance(back, dict) else getattr(back, k)
        assert np.array_equal(np.asarray(got.cpu() if hasattr(got, "cpu") else got, dtype=np.float32), p[k].numpy()), k
