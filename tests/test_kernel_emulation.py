"""CPU: the SOURCE TEXT of the map-compaction kernels (gaus-slam_b200/csrc/compact.cu) executed on the host under
tests/emu/cuda_emu.h (one std::thread per CUDA thread, barriers and warp intrinsics modelled) against a sequential
compaction.  This checks the kernels' indexing, scan and ranking logic in the container that has no GPU; it says
nothing about performance or hardware memory semantics -- the GPU tests (tests/test_zz_map_compaction.py) do."""
import os
import re
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
EMU = os.path.join(HERE, "emu")


def _kernel_text(path, first, last):
    """The text of `path` from the line containing `first` up to (not including) the line containing `last`."""
    src = open(path).read()
    a, b = src.index(first), src.index(last)
    assert 0 <= a < b
    return src[src.rfind("\n", 0, a) + 1:src.rfind("\n", 0, b) + 1]


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++ (C++20)")
def test_compaction_kernels_under_host_emulation(tmp_path):
    text = _kernel_text(os.path.join(ROOT, "gaus-slam_b200", "csrc", "compact.cu"),
                        "constexpr int CMP_THREADS", "inline int compact_blocks")
    assert len(re.findall(r"__global__", text)) == 3        # count, scan, move: the whole device side
    harness = open(os.path.join(EMU, "compact_harness.cpp.in")).read().replace("//@KERNELS@", text)
    cpp, exe = os.path.join(tmp_path, "compact_emu.cpp"), os.path.join(tmp_path, "compact_emu")
    with open(cpp, "w") as f:
        f.write(harness)
    r = subprocess.run(["g++", "-std=c++20", "-O1", "-pthread", "-I", EMU, cpp, "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "all cases agree" in r.stdout, (r.stdout[-4000:], r.stderr[-2000:])
