#!/usr/bin/env python
"""Raw-page CSV of an `ncu --set full` capture -> a small markdown table: one row per launch with the counters the design
discussion uses (duration, DRAM bytes, L2->SM bytes, tensor-pipe activity, issue stalls, registers, shared-memory conflicts)."""
import csv
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
COLS = [
    ("us", "gpu__time_duration.sum"), ("grid", "launch__grid_size"), ("regs", "launch__registers_per_thread"),
    ("DRAM rd MB", "dram__bytes_read.sum"), ("DRAM wr MB", "dram__bytes_write.sum"), ("L2->SM MB", "l1tex__m_xbar2l1tex_read_bytes.sum"),
    ("L2 hit %", "lts__t_sector_hit_rate.pct"), ("DRAM % of peak", "dram__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("L2 % of peak", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("tensor pipe % (active cycles)", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    ("SM active cycles", "sm__cycles_active.avg"), ("elapsed cycles", "sm__cycles_elapsed.max"),
    ("stall long_scoreboard", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
    ("stall barrier", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"),
    ("smem st conflicts", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum"),
    ("warp insts", "smsp__inst_executed.sum"),
]


def main():
    for path in sys.argv[1:]:
        rows = list(csv.reader(open(path)))
        hdr, units, data = rows[0], rows[1], rows[2:]
        col = {h: i for i, h in enumerate(hdr)}
        print(f"### {path}\n")
        print("| kernel | " + " | ".join(c[0] for c in COLS) + " |")
        print("|---|" + "---|" * len(COLS))
        for r in data:
            name = r[col["Kernel Name"]].replace("void ", "").replace("unnamed>::", "").split("(")[0][:70]
            vals = []
            for label, key in COLS:
                if key not in col or r[col[key]] == "":
                    vals.append("")
                    continue
                v = float(r[col[key]])
                u = units[col[key]]
                if u in UNIT:
                    v = v * UNIT[u] / 1e6
                vals.append(f"{v:.1f}" if abs(v) < 1e5 else f"{v:.3g}")
            print(f"| `{name}` | " + " | ".join(vals) + " |")
        print()


if __name__ == "__main__":
    main()
