"""Summarise an .ncu-rep (ncu --set full) into a small markdown table: one row per profiled launch.

    python tools/ncu_summary.py gpurun_out/r1_step_20cm.ncu-rep > profiles/r1_step_20cm_ncu.md
"""
import csv
import io
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "time_us", 1e-3),
    ("dram__bytes_read.sum", "dram_rd_MB", 1e-6),
    ("dram__bytes_write.sum", "dram_wr_MB", 1e-6),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%", 1),
    ("lts__t_sector_hit_rate.pct", "l2_hit_%", 1),
    ("l1tex__t_sector_hit_rate.pct", "l1_hit_%", 1),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ_%", 1),
    ("launch__registers_per_thread", "regs", 1),
    ("launch__grid_size", "grid", 1),
    ("launch__block_size", "block", 1),
    ("smsp__inst_executed.sum", "inst_M", 1e-6),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_%", 1),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_thr_%", 1),
    ("l1tex__m_l1tex2xbar_write_sectors_mem_global_op_atom.sum", "atom_sect_K", 1e-3),
    ("l1tex__m_l1tex2xbar_write_sectors_mem_global_op_red.sum", "red_sect_K", 1e-3),
    ("smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "st_long_sb", 1),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long_sb", 1),
    ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg_thr", 1),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "st_barrier", 1),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "st_short_sb", 1),
    ("smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "st_membar", 1),
]


def to_bytes(val, unit):
    unit = unit.strip()
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
             "ns": 1, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "nsecond": 1, "second": 1e9}
    return val * scale.get(unit, 1)


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    header, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(header)}
    names = []
    seen = set()
    for m, short, _ in METRICS:
        if m in col and short not in seen:
            names.append((m, short))
            seen.add(short)
    print("| kernel | " + " | ".join(s for _, s in names) + " |")
    print("|---|" + "---|" * len(names))
    for r in data:
        k = r[col["Kernel Name"]].split("(")[0]
        cells = []
        for m, short in names:
            raw = r[col[m]].replace(",", "")
            try:
                v = float(raw)
            except ValueError:
                cells.append(raw)
                continue
            u = units[col[m]]
            if short.endswith("_MB"):
                v = to_bytes(v, u) * 1e-6
            elif short == "time_us":
                v = to_bytes(v, u) * 1e-3
            elif short == "inst_M":
                v *= 1e-6
            elif short.endswith("_K"):
                v *= 1e-3
            cells.append(f"{v:.1f}" if abs(v) < 1e4 else f"{v:.0f}")
        print(f"| {k} | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
