#!/usr/bin/env python
"""Per-kernel SASS evidence of the built libvo_b200.so -> profiles/sass_summary_rNN.txt.

`cuobjdump -sass` + `cuobjdump -res-usage` on the in-tree library (sm_100a); for every kernel the
counts of the instructions that prove which hardware path it uses:
  UTCHMMA / UTCIMMA / UTCQMMA  tcgen05.mma kind::f16 / kind::i8 / kind::f8f6f4
  UTCOMMA                      tcgen05.mma kind::mxf4 .block_scale (K1f; its first source is tmem[..]:
                               the row operand is read from TMEM)
  LDTM / STTM                  tcgen05.ld / tcgen05.st  (TMEM <-> registers)
  FMNMX3                       3-input fp32 max (K1f's whole ranking)
  UTCBAR                       tcgen05.commit
  UBLKCP                       cp.async.bulk (1-D bulk copy; no tensor-map TMA is used: UTMALDG = 0)
  SYNCS                        mbarrier operations
  POPC, LOP3, VIMNMX*, REDUX/CREDUX, DFMA/DMUL/DADD (fp64 geometry), SHFL
Run on the build container (no GPU needed):  python tools/sass_summary.py [round-tag]
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "visual_odometry_b200", "csrc", "libvo_b200.so")
PATTERNS = ["UTCHMMA", "UTCIMMA", "UTCQMMA", "UTCOMMA", "LDTM", "STTM", "FMNMX3", "UTCBAR", "UBLKCP", "UTMALDG", "SYNCS", "POPC",
            "LOP3", "VIMNMX", "REDUX", "SHFL", "DFMA", "DMUL", "DADD", "MUFU", "IMAD", "BAR.SYNC", "ATOM", "RED."]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout
    return dict(zip(names, out.splitlines()))


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    regs = {}
    cur = None
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        m = re.search(r"REG:(\d+).*?SHARED:(\d+)", line)
        if m and cur:
            regs[cur] = (int(m.group(1)), int(m.group(2)))
    counts = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        counts[cur]["_total"] += 1
        for p in PATTERNS:
            if op.startswith(p) or (p.endswith(".") and op.startswith(p[:-1] + ".")) or (p == "BAR.SYNC" and op.startswith("BAR.SYNC")):
                counts[cur][p] += 1
    names = demangle(list(counts))
    lines = ["# SASS summary of visual_odometry_b200/csrc/libvo_b200.so (sm_100a), cuobjdump -sass / -res-usage",
             "# kernel | registers | static smem | SASS instructions | counted mnemonics (non-zero only)"]
    for k, c in counts.items():
        r = regs.get(k, (None, None))
        parts = [f"{p}={c[p]}" for p in PATTERNS if c[p]]
        short = re.sub(r"\(.*", "", names.get(k, k))
        lines.append(f"{short} | regs={r[0]} | smem={r[1]} | n={c['_total']} | " + " ".join(parts))
    out = os.path.join(ROOT, "profiles", f"sass_summary_{tag}.txt")
    open(out, "w").write("\n".join(lines) + "\n")
    print(out, len(counts), "kernels")


if __name__ == "__main__":
    main()
