#!/usr/bin/env python
"""Key metrics per kernel launch from an .ncu-rep (reads `ncu -i rep --page raw --csv`). usage: ncu_summary.py rep [...]"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "dur"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("sm__inst_executed.sum", "inst"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "bankconf"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"),
    ("sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
    ("sm__inst_executed_pipe_tensor_op_gen5_mma.sum", "tc5"),
    ("launch__registers_per_thread", "regs"),
    ("smsp__cycles_active.avg", "cyc"),
]
STALL = "smsp__average_warp"  # prefix of per-reason latency metrics (ncu >= 2024)


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print(f"== {path}")
    for r in rows[2:]:
        name = r[idx["Kernel Name"]].split("(")[0]
        grid, block = r[idx["Grid Size"]], r[idx["Block Size"]]
        print(f"-- {name} grid={grid} block={block}")
        for k, short in KEYS:
            if k in idx:
                print(f"   {short:9s} {r[idx[k]]:>18s} {units[idx[k]]}")
        st = []
        for h, i in idx.items():
            if "warp_issue_stalled" in h and h.endswith("_per_warp_active.pct"):
                try:
                    st.append((float(r[i].replace(",", "")), h.split("warp_issue_stalled_")[1].replace("_per_warp_active.pct", "")))
                except ValueError:
                    pass
        st.sort(reverse=True)
        print("   stalls:", ", ".join(f"{n}={v:.0f}" for v, n in st[:6]))


if __name__ == "__main__":
    for p in sys.argv[1:]:
        main(p)
