"""Selected metrics of the first launch matching a kernel-name substring in an .ncu-rep -> JSON (for profiles/)."""
import csv
import json
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum"]
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}


def main(rep, needle, out, note=""):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    h, u = rows[0], rows[1]
    ki = h.index("Kernel Name")
    row = next(r for r in rows[2:] if needle in r[ki])
    m = {}
    for k in WANT:
        if k in h:
            i = h.index(k)
            try:
                m[k] = {"value": float(row[i].replace(",", "")), "unit": u[i]}
            except ValueError:
                pass

    def nbytes(k):
        return m[k]["value"] * UNIT.get(m[k]["unit"], 1.0)

    rec = {"kernel": row[ki].split("(")[0], "report": rep, "note": note,
           "dram_read_bytes": nbytes("dram__bytes_read.sum"), "dram_write_bytes": nbytes("dram__bytes_write.sum"),
           "dram_bytes_per_launch": nbytes("dram__bytes_read.sum") + nbytes("dram__bytes_write.sum"), "metrics": m}
    with open(out, "w") as f:
        json.dump(rec, f, indent=1)
    print(out, rec["kernel"], rec["dram_bytes_per_launch"] / 1e9, "GB", m["gpu__time_duration.sum"])


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
