"""Summarise one kernel of an .ncu-rep into the JSON files kept under profiles/ (reads `ncu -i REP --page raw --csv`).

usage: python tools/ncu_summary.py REP OUT.json "note" [row]
"""
import csv
import json
import subprocess
import sys

KEEP = (
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
)


def main():
    rep, out, note = sys.argv[1], sys.argv[2], sys.argv[3]
    row = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    head, units, vals = rows[0], rows[1], rows[2 + row]
    col = {n: i for i, n in enumerate(head)}
    metrics = {}
    for n in KEEP:
        if n in col and vals[col[n]] != "":
            metrics[n] = {"value": float(vals[col[n]].replace(",", "")), "unit": units[col[n]]}
    stalls = []
    for n, i in col.items():
        if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio") and vals[i] != "":
            stalls.append((float(vals[i]), n[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
    stalls.sort(reverse=True)
    doc = {"kernel": vals[col["Kernel Name"]], "report": rep, "note": note, "metrics": metrics,
           "top_stalls_warps_per_issue": {k: v for v, k in stalls[:6]}}
    json.dump(doc, open(out, "w"), indent=1)
    print(json.dumps(doc, indent=1)[:1500])


if __name__ == "__main__":
    main()
