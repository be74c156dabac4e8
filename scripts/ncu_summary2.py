#!/usr/bin/env python
"""ncu raw CSV (ncu -i x.ncu-rep --page raw --csv) -> profiles/r2_ncu_summary.json + a readable table.

    python scripts/ncu_summary2.py gpurun_out/raw_all.csv [commit] > profiles/r2_ncu_full_all_kernels.txt

Per kernel (first capture of each name): duration, DRAM bytes read + written, L2 bytes, issue-slot use, lanes per
instruction, FP32 (fma) / XU / LSU / ALU pipe utilisation (% of the pipe's peak), shared-memory bank-conflict share, occupancy,
registers, and the warp stall reasons per issued instruction.  The JSON carries the hash of csrc/ (bench.py drops the
numbers when the sources have changed since)."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NAMES = {"composite_forward_kernel": "composite_forward", "composite_backward_pixels_kernel": "composite_backward",
         "composite_backward_kernel": "composite_backward", "tile_sort_kernel": "tile_sort_gather_footprint",
         "preprocess_forward_kernel": "preprocess_forward", "preprocess_backward_kernel": "preprocess_backward",
         "scatter_kernel": "scatter_pairs", "scan_tiles_kernel": "scan_tiles"}
M = {"us": "gpu__time_duration.sum", "dram_rd": "dram__bytes_read.sum", "dram_wr": "dram__bytes_write.sum",
     "l2_sectors": "lts__t_sectors.sum", "inst": "smsp__inst_executed.sum",
     "lanes": "smsp__thread_inst_executed_per_inst_executed.ratio",
     "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active",
     "warps": "sm__warps_active.avg.pct_of_peak_sustained_active", "regs": "launch__registers_per_thread",
     "fma": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "xu": "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
     "lsu": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "alu": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
     "bank": "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "wave": "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"}
STALL = "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio"
STALLS = ["barrier", "long_scoreboard", "short_scoreboard", "wait", "mio_throttle", "lg_throttle", "math_pipe_throttle",
          "branch_resolving", "not_selected", "no_instruction", "dispatch_stall"]


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0, "us": 1.0, "ms": 1e3, "ns": 1e-3}
    out = {}
    for r in rows[2:]:
        kn = r[col["Kernel Name"]]
        name = next((v for k, v in NAMES.items() if k in kn), None)
        if name is None or name in out:
            continue
        g = lambda key: (num(r[col[key]]) or 0.0) * scale.get(units[col[key]], 1.0) if key in col else None   # noqa: E731
        v = {k: g(m) for k, m in M.items()}
        e = {"duration_us": v["us"], "dram_bytes": (v["dram_rd"] or 0) + (v["dram_wr"] or 0), "dram_read_bytes": v["dram_rd"],
             "dram_write_bytes": v["dram_wr"], "l2_bytes": (v["l2_sectors"] or 0) * 32.0, "inst_executed": v["inst"],
             "active_lanes_per_inst": v["lanes"], "issue_active_pct": v["issue"], "warps_active_pct": v["warps"],
             "registers": v["regs"],
             "pipe_pct_of_peak": {k: v[k] for k in ("fma", "xu", "lsu", "alu")},
             "smem_bank_conflict_share": (v["bank"] / v["wave"] if v["wave"] else None),
             "stall_per_issue": {s: g(STALL % s) for s in STALLS}}
        out[name] = e
    import bench
    commit = sys.argv[2] if len(sys.argv) > 2 else None
    with open(os.path.join(ROOT, "profiles", "r2_ncu_summary.json"), "w") as f:
        json.dump({"commit": commit, "csrc_sha256": bench.csrc_sha256(), "source": os.path.basename(sys.argv[1]),
                   "kernels": out}, f, indent=1)
    for n, e in out.items():
        print(f"== {n}: {e['duration_us']:.1f} us (under ncu, cold cache), {e['registers']:.0f} regs, warps active {e['warps_active_pct']:.1f} %")
        print(f"   DRAM {(e['dram_read_bytes'] or 0) / 1e6:.1f} MB read + {(e['dram_write_bytes'] or 0) / 1e6:.1f} MB written; L2 {(e['l2_bytes'] or 0) / 1e6:.1f} MB")
        ps = e["pipe_pct_of_peak"]
        print(f"   {e['inst_executed'] / 1e6:.1f} M warp instructions, {e['active_lanes_per_inst']:.1f} lanes / instruction, issue slots "
              f"{e['issue_active_pct']:.1f} %; pipe use (% of each pipe's peak): fma {ps['fma'] or 0:.0f}, alu {ps['alu'] or 0:.0f}, "
              f"lsu {ps['lsu'] or 0:.0f}, xu {ps['xu'] or 0:.1f}; shared-memory bank-conflict share "
              f"{100 * (e['smem_bank_conflict_share'] or 0):.0f} %")
        st = sorted(((k, x) for k, x in e["stall_per_issue"].items() if x), key=lambda kv: -kv[1])[:5]
        print("   stalls (warps per issued instruction): " + ", ".join(f"{k} {x:.2f}" for k, x in st))


if __name__ == "__main__":
    main()
