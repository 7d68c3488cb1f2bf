"""Occurrences of selected SASS mnemonics per kernel of the built library (profiles/rN_sass_histogram.txt):
python tools/sass_histogram.py gtsfm_b200/libgtsfm_b200.so > profiles/r2_sass_histogram.txt"""
import collections
import re
import subprocess
import sys

MNEMONICS = ["UTCHMMA", "LDTM", "UTMALDG", "UTCBAR", "SYNCS", "IMMA", "FFMA2", "FMNMX3", "DFMA", "POPC", "REDUX", "CREDUX", "MUFU", "HFMA2"]


def main(lib):
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    head = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    counts, kernel = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            kernel = re.sub(r"^void ", "", name).split("(")[0].replace("g2v::", "")
            counts[kernel] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)", line)
        if m and kernel:
            op = m.group(1)
            for mn in MNEMONICS:
                if op == mn:
                    counts[kernel][mn] += 1
    print(f"# cuobjdump -sass {lib} (sm_100a, built from commit {head}): occurrences of selected mnemonics per kernel")
    print("# UTCHMMA = tcgen05.mma kind::f16, LDTM = tcgen05.ld (TMEM), UTMALDG = TMA tensor load, IMMA = mma.sync u8, FFMA2 = fma.rn.f32x2,")
    print("# CREDUX = redux.sync.min.f32 (the column fold of match_screen_kernel)")
    print("kernel".ljust(46) + "".join(m.rjust(9) for m in MNEMONICS))
    for k in sorted(counts):
        print(k[:45].ljust(46) + "".join(str(counts[k][m]).rjust(9) for m in MNEMONICS))


if __name__ == "__main__":
    main(sys.argv[1])
