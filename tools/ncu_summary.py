#!/usr/bin/env python
"""Print the handful of ncu metrics we track per kernel from a .ncu-rep (via `ncu --page raw --csv`)."""
import csv
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__grid_size", "grid"),
    ("smsp__warps_active.avg.per_cycle_active", "warps active / SMSP"),
    ("smsp__warps_eligible.avg.per_cycle_active", "warps eligible / SMSP"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64 pipe %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu pipe %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu pipe %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts"),
    ("sm__cycles_elapsed.max", "cycles"),
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        print("== %s" % r[idx["Kernel Name"]][:90])
        for key, label in WANT:
            if key in idx:
                print("   %-26s %s %s" % (label, r[idx[key]], units[idx[key]]))
        # warp-state samples (smsp__pcsamp_warps_issue_stalled_<reason>): share of all samples per reason
        stalls = []
        for h, i in idx.items():
            if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued"):
                try:
                    stalls.append((float(r[i].replace(",", "")), h[len("smsp__pcsamp_warps_issue_stalled_"):]))
                except ValueError:
                    pass
        tot = sum(v for v, _ in stalls) or 1.0
        stalls.sort(reverse=True)
        print("   top stalls: " + ", ".join("%s %.0f%%" % (n, 100 * v / tot) for v, n in stalls[:7]))
        ops = []
        for key, label in (("smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "dfma"), ("smsp__sass_thread_inst_executed_op_dmul_pred_on.sum", "dmul"),
                           ("smsp__sass_thread_inst_executed_op_dadd_pred_on.sum", "dadd"), ("sm__sass_thread_inst_executed_op_dfma_pred_on.sum", "dfma"),
                           ("sm__sass_thread_inst_executed_op_dmul_pred_on.sum", "dmul"), ("sm__sass_thread_inst_executed_op_dadd_pred_on.sum", "dadd")):
            if key in idx and r[idx[key]]:
                ops.append("%s %s" % (label, r[idx[key]]))
        if ops:
            print("   fp64 thread instructions: " + ", ".join(ops))


if __name__ == "__main__":
    main()
