#!/usr/bin/env python
"""cuobjdump -sass of the in-tree libegoscaler_b200.so -> per-kernel histogram of the Blackwell-specific mnemonics
(UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA load/store, UTCBAR = tcgen05.commit, SYNCS = mbarrier,
HMMA = mma.sync) plus the instruction count.  Usage: python tools/sass_histogram.py > profiles/rNN_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "egoscaler_b200", "libegoscaler_b200.so")
KEYS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "UTCCP", "SYNCS", "HMMA", "MUFU.EX2", "LDG", "STG", "RED", "ATOMG"]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
    cur, hist, total = None, collections.OrderedDict(), {}
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            hist[cur] = collections.Counter()
            total[cur] = 0
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            op = m.group(1)
            total[cur] += 1
            for k in KEYS:
                if op == k or op.startswith(k + ".") or (k == "MUFU.EX2" and op.startswith("MUFU.EX2")):
                    hist[cur][k] += 1
    print(f"# SASS mnemonic histogram of {os.path.relpath(SO, ROOT)} (sm_100a), cuobjdump -sass; columns: instructions | " + " ".join(KEYS))
    for fn, h in hist.items():
        name = re.sub(r"\(.*", "", demangle(fn))
        name = name.replace("esb::(anonymous namespace)::", "esb::").replace("void ", "")
        print(f"{name:<70s} {total[fn]:6d} | " + " ".join(f"{k}={h[k]}" for k in KEYS if h[k]))


if __name__ == "__main__":
    main()
