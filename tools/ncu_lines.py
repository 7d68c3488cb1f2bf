"""Aggregate the warp-stall samples / executed instructions of an .ncu-rep by CUDA source line (needs -lineinfo and
--import-source on):  python tools/ncu_lines.py report.ncu-rep [top_n] [kernel-name regex] [launch index among matches]"""
import csv
import subprocess
import sys


def main(path, top=40, kernel=None, skip=0):
    cmd = ["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"]
    if kernel:
        cmd += ["-k", "regex:" + kernel, "--launch-skip", str(skip), "--launch-count", "1"]
    out = subprocess.run(cmd, capture_output=True, text=True).stdout
    cur, agg = None, {}
    for r in csv.reader(out.splitlines()):
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if len(r) > 7 and r[0] not in ("", "Line No"):
            try:
                key = (cur, int(r[0]), r[1][:100])
                s, n = int(r[6]), int(r[7])
            except ValueError:
                continue
            a = agg.setdefault(key, [0, 0])
            a[0] += s
            a[1] += n
    tot = sum(v[0] for v in agg.values()) or 1
    toti = sum(v[1] for v in agg.values()) or 1
    print("samples", tot, "warp instructions", toti)
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{k[0][:20]:20s} {k[1]:4d} {100 * v[0] / tot:5.1f}% {100 * v[1] / toti:5.1f}%  {k[2]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40, sys.argv[3] if len(sys.argv) > 3 else None,
         int(sys.argv[4]) if len(sys.argv) > 4 else 0)
