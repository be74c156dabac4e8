#!/usr/bin/env python
"""Per-kernel SASS evidence of the Blackwell-native mechanisms in libgs2d.so (no GPU needed):

    python scripts/sass_extract.py > profiles/r2_sass_extract.txt

Counts, per kernel function: UBLKCP (cp.async.bulk: TMA 1-D bulk copy), SYNCS.* (mbarrier), CCTL.E.RML2 (discard.global.L2),
LDGMC / STMC-like multicast forms (multimem.ld_reduce / multimem.st), REDG...STRONG.SYS (multimem.red on a multicast
address, gs2d_backward_reduce), MUFU.* (XU pipe), and -- as a negative check -- UTC*MMA / HMMA (none: this path has no
dense contraction)."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "gaus-slam_b200", "lib", "libgs2d.so")
PAT = [("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS\."), ("CCTL.E.RML2", r"CCTL\.E\.RML2"), ("LDGMC", r"\bLDGMC"),
       ("ST.*MC / STGMC", r"\bST\w*MC|\bSTGMC|MULTIMEM"), ("REDG.STRONG.SYS", r"\bREDG\.E\.ADD\.F32\S*STRONG\.SYS"),
       ("ATOMG / REDG (global atomics, all)", r"\b(ATOMG|REDG|RED)\.E"), ("MUFU", r"\bMUFU\."), ("tensor-core MMA", r"UTC\w*MMA|\bHMMA|\bIMMA")]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} -- instruction counts per kernel (sm_100a)")
    name, counts, total = None, None, 0
    rows = []

    def flush():
        if name is not None:
            rows.append((name, total, dict(counts)))
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            flush()
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = name.replace("(anonymous namespace)::", "").replace("void ", "")
            name = re.sub(r"\(.*", "", name)
            counts, total = {k: 0 for k, _ in PAT}, 0
            continue
        if name is None or not re.search(r"/\*[0-9a-f]{4}\*/", line):
            continue
        total += 1
        for k, p in PAT:
            if re.search(p, line):
                counts[k] += 1
    flush()
    hdr = ["kernel", "SASS instr"] + [k for k, _ in PAT]
    print(" | ".join(hdr))
    tot = {k: 0 for k, _ in PAT}
    for n, t, c in sorted(rows, key=lambda r: (r[0], r[1])):
        print(" | ".join([n[:70], str(t)] + [str(c[k]) for k, _ in PAT]))
        for k in tot:
            tot[k] += c[k]
    print(" | ".join(["TOTAL", str(sum(t for _, t, _ in rows))] + [str(tot[k]) for k, _ in PAT]))


if __name__ == "__main__":
    sys.exit(main())
