"""Multi-GPU parity on hardware (needs >= 2 GPUs in one box; skipped otherwise): the view-sharded step's summed map
gradient -- through NCCL, through libgs2d's in-switch all-reduce and through the FUSED backward (gs2d_backward_reduce) --
against the gathered sum and against the sum of the per-view gradients of the compiled reference (scripts/fused_check.py)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.gpu
def test_summed_gradients_of_two_ranks_on_hardware(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one box")
    out = os.path.join(tmp_path, "check.json")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "scripts", "fused_check.py"), "--surfels", "200000",
           "--iters", "3", "--out", out]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    d = json.load(open(out))
    assert d["world"] == 2 and not d["fails"]
    assert d["checks"]["nccl_vs_gathered_sum"] <= 1e-5
    if d["checks"]["product_sum_vs_reference_sum_worst_group"] is not None:
        assert d["checks"]["product_sum_vs_reference_sum_worst_group"] <= 1e-3
    if d["symmetric_memory"]:
        assert d["checks"]["fused_backward_vs_nccl"] <= 1e-4 and d["checks"]["multimem_allreduce_vs_nccl"] <= 1e-4
