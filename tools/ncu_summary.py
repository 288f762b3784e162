#!/usr/bin/env python
"""Condense an .ncu-rep (ncu --set full) into the few per-launch numbers quoted in DESIGN.md / profiles/.

    python tools/ncu_summary.py gpurun_out/prof_x.ncu-rep > profiles/rN_x.md

One row per profiled launch: duration, DRAM bytes read/written and achieved DRAM bandwidth, pipe utilisation
(FP64, DMMA tensor sub-pipe), L1/L2 throughput, occupancy, registers, top warp-stall reasons."""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("dram__bytes.sum.per_second", "dram_bw"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_%"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1_%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_%"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64_pipe_%"),
    ("sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active", "dmma_pipe_%"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_dynamic", "dyn_smem"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_bank_conflicts"),
    ("smsp__inst_executed.sum", "warp_insts"),
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {k: i for i, k in enumerate(hdr)}
    print("# %s\n" % rep.split("/")[-1])
    print("`ncu -i %s --page raw --csv`, %d launch(es)\n" % (rep.split("/")[-1], len(data)))
    for li, r in enumerate(data):
        print("## launch %d: `%s`\n" % (li, r[col["Kernel Name"]][:90]))
        print("| metric | value | unit |\n|---|---|---|")
        for k, short in KEYS:
            if k in col and r[col[k]] != "":
                print("| %s (`%s`) | %s | %s |" % (short, k, r[col[k]], units[col[k]]))
        stalls = []
        for k, i in col.items():
            if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio"):
                try:
                    stalls.append((float(r[i]), k))
                except ValueError:
                    pass
        stalls.sort(reverse=True)
        if stalls:
            print("\nwarp stalls (average warps stalled per issue-active cycle, top 7):\n")
            for v, k in stalls[:7]:
                name = k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]
                print("* %s: %.2f" % (name, v))
        print()


if __name__ == "__main__":
    main()
