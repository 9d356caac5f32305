#!/usr/bin/env python
"""Per-kernel SASS instruction histogram of librgp_sm100.so (cuobjdump -sass), restricted to the mnemonics that prove the
Blackwell data path (B200_PROFILING.md): UTCHMMA (tcgen05.mma), UTMALDG (TMA tensor load, .MULTICAST / .IM2COL forms),
LDTM / STTM (tcgen05.ld / st), UTCBAR (tcgen05.commit), UTCCP, SYNCS (mbarrier), plus totals.
Usage: python tools/sass_histogram.py > profiles/r02_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "recurrent_gaze_prediction_b200", "librgp_sm100.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCOMMA", "UTMALDG", "UTMASTG", "UTMAPF", "LDTM", "STTM", "UTCBAR", "UTCCP", "UTCATOMSWS",
        "SYNCS", "HMMA", "FFMA", "DFMA", "DADD", "DMUL", "MUFU", "LDG", "STG", "RED", "ATOMG", "LDS", "STS", "BAR"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", line)
        if m and cur:
            kernels[cur]["_total"] += 1
            op = m.group(1)
            kernels[cur][op] += 1
            if op in ("UTMALDG", "UTCHMMA", "UTCBAR", "LDG", "STG"):
                kernels[cur][op + m.group(2)] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), stdout=subprocess.PIPE, text=True).stdout.splitlines()
    print("# cuobjdump -sass %s : per-kernel instruction histogram (selected mnemonics)" % os.path.relpath(LIB, ROOT))
    tot = collections.Counter()
    for (name, c), dn in zip(kernels.items(), demangle):
        short = re.sub(r"\(.*", "", dn.replace("(anonymous namespace)::", ""))
        sel = [(k, c[k]) for k in KEYS if c[k]]
        forms = sorted((k, v) for k, v in c.items() if "." in k and k.split(".")[0] in ("UTMALDG", "UTCHMMA", "UTCBAR"))
        print("%-70s total %6d  %s" % (short[:70], c["_total"], " ".join("%s=%d" % kv for kv in sel)))
        if forms:
            print("%-70s        forms: %s" % ("", " ".join("%s=%d" % kv for kv in forms)))
        for k in KEYS:
            tot[k] += c[k]
    print("# library totals: " + " ".join("%s=%d" % (k, tot[k]) for k in KEYS if tot[k]))


if __name__ == "__main__":
    main()
