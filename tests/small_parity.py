"""Diagnostic (GPU): product vs compiled reference on every small case of tests/cases.py, printed per field.
The pass/fail versions are tests/test_gpu_parity.py; this prints details for debugging."""
from __future__ import annotations

import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import cases  # noqa: E402
import compare  # noqa: E402
import gpu_run  # noqa: E402


def main():
    bad = 0
    for name in cases.CASES:
        case = cases.CASES[name]()
        ref = gpu_run.run_ref(case)
        ours = gpu_run.run_ours(case, debug=True)
        ours2 = gpu_run.run_ours(case, debug=False)
        f = compare.check_binning(ours, ref, exact_float_state=True)
        f += compare.check_images(ours, ref, exact=True)
        f += compare.check_grads(ours, ref)
        f += ["nodebug: " + x for x in compare.check_images(ours2, ours, exact=True)]
        nc = int((ours["n_contrib"] != ref["n_contrib"]).sum())
        if nc:
            idx = np.argwhere(ours["n_contrib"] != ref["n_contrib"])[:5]
            f.append(f"n_contrib: {nc} differ, samples " +
                     str([(tuple(i), int(ours['n_contrib'][tuple(i)]), int(ref['n_contrib'][tuple(i)])) for i in idx]))
        gr = {k: compare.grad_rel(ours["grad_" + k], ref["grad_" + k]) for k in compare.GRADS
              if ref["grad_" + k].size}
        f += [] if int(ours["footprint_audit"][2]) == 0 else [f"footprint audit: {ours['footprint_audit']}"]
        print(f"  footprint audit (exact hits, mask bits, missed): {ours['footprint_audit']}")
        print(f"[{name}] I={int(ref['num_rendered'])} visible={int((ref['radii'] > 0).sum())} "
              f"{'OK' if not f else 'FAIL'}  grad rel: " + " ".join(f"{k}={v:.1e}" for k, v in gr.items()))
        for x in f:
            print("    ", x)
        bad += bool(f)
    print("cases failing:", bad)


if __name__ == "__main__":
    main()
