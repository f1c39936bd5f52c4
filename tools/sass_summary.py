#!/usr/bin/env python
"""Mnemonic counts per kernel of the built library (cuobjdump -sass): which kernels use the 5th-generation tensor cores
(UTC*MMA = tcgen05.mma), tensor memory (LDTM / STTM = tcgen05.ld / st, UTCBAR = tcgen05.commit, UTCATOMSWS = alloc), TMA
(UTMALDG / UTMASTG / UBLKCP), multicast stores (multimem -> *.MMEM / REDG...), peer-visible release/acquire flags, and
how many instructions each kernel has.  Runs on a CPU-only box.

usage: sass_summary.py [lib.so] > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "layer-wise-aaai20_b200", "_native", "liblwc_b200.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "SYNCS", "HMMA",
        "MULTIMEM", "MEMBAR", "ATOM", "RED", "LDG", "STG", "LDS", "STS", "SHFL", "BAR", "LDL", "STL"]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kern, counts, total = None, collections.OrderedDict(), {}
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            kern = m.group(1)
            counts[kern] = collections.Counter()
            total[kern] = 0
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P[0-9T]+\s+)?([A-Z0-9_.]+)", line)
        if m and kern:
            op = m.group(1)
            total[kern] += 1
            if ".STRONG.SYS" in op and (op.startswith("ST") or op.startswith("LD")):
                counts[kern]["SYS-scope " + op.split(".")[0]] += 1      # peer-visible flags / payload, multimem stores
            for k in KEYS:
                if op.startswith(k) or (k == "MULTIMEM" and "MMEM" in op) or ("." + k) in ("." + op.split(".")[0]):
                    counts[kern][k if not op.startswith("UTMALDG") else op.split(".")[0] + "." + (op.split(".")[1] if "." in op else "")] += 1
                    break
    demangle = subprocess.run(["c++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
    print(f"# {os.path.basename(LIB)}: SASS mnemonic counts per kernel (sm_100a)")
    for name, pretty in zip(counts, demangle):
        c = counts[name]
        tags = []
        if any(k.startswith("UTCHMMA") or k.startswith("UTCQMMA") for k in c):
            tags.append("tcgen05.mma")
        if "LDTM" in c or "STTM" in c:
            tags.append("TMEM")
        if any(k.startswith("UTMALDG") for k in c) or "UTMASTG" in c or "UBLKCP" in c:
            tags.append("TMA")
        if "MULTIMEM" in c:
            tags.append("multimem")
        print(f"\n{pretty[:150]}\n   instructions {total[name]:6d}   [{', '.join(tags) or 'CUDA cores'}]")
        print("   " + "  ".join(f"{k}={v}" for k, v in sorted(c.items())))


if __name__ == "__main__":
    main()
