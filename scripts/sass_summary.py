"""Per-kernel SASS evidence for profiles/: counts of the Blackwell-specific mnemonics in libveles_b200.so
(UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG/UTMAREDG = TMA, multimem = LDGMC/ST.MC/REDG..., SYNCS =
mbarrier).  Run here (no GPU needed): python scripts/sass_summary.py > profiles/sass_summary.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "veles_b200", "lib", "libveles_b200.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "UBLKCP", "UGETNEXTWORKID", "SYNCS",
        "HMMA", "MULTIMEM", "LDG", "STG", "RED", "ATOM", "LDS", "STS", "MUFU", "BAR", "UCGABAR", "MEMBAR"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, check=True).stdout.decode(errors="replace")
    counts = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        counts[cur]["_total"] += 1
        base = op.split(".")[0]
        if ".MC" in op or "MULTIMEM" in op or base in ("LDGMC", "STGMC"):
            counts[cur]["MULTIMEM"] += 1
        for k in KEYS:
            if base == k:
                counts[cur][k] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(counts).encode(), stdout=subprocess.PIPE).stdout.decode().splitlines()
    print("# SASS summary of veles_b200/lib/libveles_b200.so (sm_100a)\n")
    print("`cuobjdump -sass`, instruction counts per kernel; tcgen05.mma = UTC*MMA, tcgen05.ld = LDTM, "
          "TMA = UTMALDG/UTMASTG/UTMAREDG, mbarrier = SYNCS, multimem = *.MC*.\n")
    cols = [k for k in KEYS if any(c[k] for c in counts.values())]
    print("| kernel | instrs | " + " | ".join(cols) + " |")
    print("|---|---|" + "---|" * len(cols))
    for (name, c), dn in zip(counts.items(), demangle):
        short = re.sub(r"\(.*", "", dn)[:90]
        print("| `%s` | %d | %s |" % (short, c["_total"], " | ".join(str(c[k]) if c[k] else "" for k in cols)))


if __name__ == "__main__":
    main()
