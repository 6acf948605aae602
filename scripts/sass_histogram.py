"""profiles/r2_sass_histogram.txt: per-kernel SASS instruction histogram of kernels_tensor.o (cuobjdump -sass, sm_100a).
usage: python scripts/sass_histogram.py > profiles/r2_sass_histogram.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "alink_b200", "csrc", "kernels_tensor.o")
KEYS = ("UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "SYNCS", "FFMA2", "FMNMX3", "FMNMX", "BAR", "LDS", "STS", "LDG", "STG", "ATOMG", "RED")
DETAIL = ("UTC", "UTMA", "LDTM", "SYNCS", "FFMA2", "FMNMX")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", OBJ], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["cu++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            short = re.sub(r"\((bool|int)\)", "", name)          # template arguments print as (bool)0
            cur = per.setdefault(short.split("(")[0], collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
    print("# SASS instruction histogram of alink_b200/csrc/kernels_tensor.o (cuobjdump -sass, sm_100a), scripts/sass_histogram.py")
    print("# columns: kernel | total instructions | selected mnemonics (prefix match: UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld,")
    print("#          UTMALDG = TMA tensor load, UTCBAR = tcgen05.commit); second line: the exact tensor / TMA / barrier / min-max forms")
    print()
    for name, c in per.items():
        total = sum(c.values())
        sel = " ".join("%s=%d" % (k, sum(v for op, v in c.items() if op == k or op.startswith(k + "."))) for k in KEYS)
        print("%-72s %5d  %s" % (name, total, sel))
        det = " ".join("%s=%d" % (op, v) for op, v in sorted(c.items()) if op.startswith(DETAIL))
        if det:
            print("    " + det)


if __name__ == "__main__":
    main()
