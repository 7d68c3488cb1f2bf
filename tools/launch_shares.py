"""Per-kernel totals of an ncu launch list (`--metrics gpu__time_duration.sum --csv`): python tools/launch_shares.py f.csv"""
import csv
import sys
from collections import defaultdict


def main(path):
    tot, cnt = defaultdict(float), defaultdict(int)
    with open(path) as f:
        rows = [r for r in csv.reader(f) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    for r in rows[1:]:
        name = r[ki].split("(")[0].replace("void ", "").replace("g2v::", "")
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
        tot[name] += v
        cnt[name] += 1
    all_ms = sum(tot.values())
    print(f"{'kernel':60s} {'launches':>8s} {'total ms':>10s} {'share':>7s}")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        print(f"{k[:60]:60s} {cnt[k]:8d} {v:10.3f} {100 * v / all_ms:6.2f}%")
    print(f"{'sum':60s} {sum(cnt.values()):8d} {all_ms:10.3f}")


if __name__ == "__main__":
    main(sys.argv[1])
