"""CPU, world_size 2 over gloo: the view-sharded mapping plumbing (gs2d_b200/mapping.py).  Per-view gradients
come from the CPU oracle here (no GPU in this container); the all-reduced flat buffer must equal the sum of
the per-view gradients computed in one process, and the view assignment must partition the window."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import cases  # noqa: F401
from gs2d_b200 import mapping, scene

P, W, H, V = 600, 96, 64, 4


def _scene():
    intr = dict(W=W, H=H, fx=48.0, fy=48.0, cx=(W - 1) / 2, cy=(H - 1) / 2)
    return intr, scene.room_scene(P, intr, seed=21), scene.view_sweep(V, intr, seed=21)


def _view_grads(sc, cam, seed):
    import cpu as O
    gc, go = scene.image_grads(W, H, seed=seed)
    bg = np.zeros(3, np.float32)
    st = O.forward(sc["means3D"], sc["opacities"], cam.viewmatrix, cam.projmatrix, W, H, bg, scales=sc["scales"],
                   rotations=sc["rotations"], colors=sc["colors"], campos=cam.campos)
    g = O.backward(st, sc["means3D"], cam.viewmatrix, cam.projmatrix, W, H, cam.tanfovx, cam.tanfovy, bg, gc, go,
                   scales=sc["scales"], rotations=sc["rotations"], campos=cam.campos)
    return dict(means3D=torch.from_numpy(g["means3D"]), opacities=torch.from_numpy(g["opacity"]),
                scales=torch.from_numpy(g["scales"]), rotations=torch.from_numpy(g["rotations"]),
                colors=torch.from_numpy(g["colors"]))


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        _, sc, cams = _scene()
        flat = torch.zeros(P * mapping.FLAT_WIDTH)
        for step in range(2):
            v = mapping.view_of(step, rank, world, V)
            mapping.allreduce_map_gradient(mapping.pack_grads(_view_grads(sc, cams[v], v), flat))
            if rank == 0:
                torch.save(flat.clone(), os.path.join(out_dir, f"flat_{step}.pt"))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_view_assignment_partitions_the_window():
    for world in (1, 2, 4, 8):
        seen = []
        for step in range(8 // world):
            vs = mapping.views_of_step(step, world, 8)
            assert len(set(vs)) == world
            seen += vs
        assert sorted(seen) == list(range(8))
        # ... and over the passes every rank gets to every view (also when the world size divides the window)
        for r in range(world):
            assert {mapping.view_of(s, r, world, 8) for s in range(64)} == set(range(8))


def test_flat_layout_roundtrip():
    g = {k: torch.randn(P, w) for k, w in mapping.PARAM_WIDTH.items()}
    flat = torch.empty(P * mapping.FLAT_WIDTH)
    back = mapping.unpack_grads(mapping.pack_grads(g, flat), P)
    assert all(torch.equal(back[k], g[k]) for k in g)
    assert [b - a for _, a, b in mapping.flat_sections(P)] == [3 * P, P, 2 * P, 4 * P, 3 * P]


@pytest.mark.timeout(300)
def test_allreduce_of_view_sharded_gradients_gloo(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    _, sc, cams = _scene()
    for step in range(2):
        want = torch.zeros(P * mapping.FLAT_WIDTH)
        tmp = torch.empty_like(want)
        for r in range(world):
            v = mapping.view_of(step, r, world, V)
            want += mapping.pack_grads(_view_grads(sc, cams[v], v), tmp)
        got = torch.load(os.path.join(tmp_path, f"flat_{step}.pt"))
        # the oracle's OpenMP gradient sums have no fixed order: allow fp32 re-association noise, nothing more
        scale = float(want.abs().max())
        assert scale > 0
        assert torch.allclose(got, want, rtol=1e-4, atol=1e-5 * scale)


def test_pack_is_a_noop_when_gradients_alias_the_flat_buffer():
    flat = torch.zeros(P * mapping.FLAT_WIDTH)
    views = mapping.unpack_grads(flat, P)
    for k, v in views.items():
        v.copy_(torch.randn_like(v))
    assert mapping.grads_alias_flat(views, flat)
    before = flat.clone()
    assert mapping.pack_grads(views, flat) is flat and torch.equal(flat, before)
    other = {k: v.clone() for k, v in views.items()}
    assert not mapping.grads_alias_flat(other, flat)
