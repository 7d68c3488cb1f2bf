"""Summarise an .ncu-rep (read with `ncu -i ... --page raw --csv`) into a small table: one line per captured launch."""
import csv
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64%"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
    ("l1tex__t_sector_hit_rate.pct", "l1hit%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("local_load", "lld"),
]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    head, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(head)}
    for r in rows[2:]:
        name = r[col["Kernel Name"]].split("(")[0][-48:]
        parts = [name]
        for k, short in KEYS:
            if k in col:
                parts.append(f"{short}={r[col[k]]}{units[col[k]] if short in ('time', 'dram_rd', 'dram_wr') else ''}")
        print("  ".join(parts))


if __name__ == "__main__":
    main(sys.argv[1])
