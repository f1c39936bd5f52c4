#!/usr/bin/env python
"""Top stall sites of one kernel from an ncu report:  ncu -i rep --page source --csv --print-source cuda|sass ... | this.

    python tools/ncu_source_top.py file.csv [N]
Prints the N source lines / instructions with the most warp-stall samples and, for each, its two dominant stall reasons.
"""
import csv
import sys


def main():
    path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if "Source" in r and any(c.startswith("stall_") for c in r))
    hdr = rows[hi]
    src = hdr.index("Source")
    samp = hdr.index("Warp Stall Sampling (All Samples)")
    stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    line_col = hdr.index("#") if "#" in hdr else None
    body = []
    for r in rows[hi + 1:]:
        if r and r[0] == "Kernel Name":          # next kernel of the report
            break
        if len(r) == len(hdr) and r != hdr:
            body.append(r)
    total = sum(float(r[samp] or 0) for r in body) or 1.0
    print(rows[0][1] if len(rows[0]) > 1 else "")
    print(f"total samples {total:.0f}")
    agg = {}
    for i, h in stall_cols:
        agg[h] = sum(float(r[i] or 0) for r in body)
    print("by reason: " + ", ".join(f"{h[6:]} {100 * v / total:.1f}%" for h, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    order = {id(r): k for k, r in enumerate(body)}
    ranked = sorted(body, key=lambda r: -float(r[samp] or 0))
    for r in ranked[:top]:
        s = float(r[samp] or 0)
        reasons = sorted(((float(r[i] or 0), h[6:]) for i, h in stall_cols), reverse=True)[:2]
        tag = (r[line_col] + ": ") if line_col is not None else f"[{order[id(r)]:5d}] "
        print(f"{100 * s / total:5.1f}%  {tag}{r[src].strip()[:110]:110s}  " + ", ".join(f"{h} {v:.0f}" for v, h in reasons if v))


if __name__ == "__main__":
    main()
