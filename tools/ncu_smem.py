#!/usr/bin/env python
"""Shared-memory wavefronts per source line of one kernel in an .ncu-rep (actual vs ideal: the excess is bank conflicts).

    python tools/ncu_smem.py prof.ncu-rep <kernel-regex> [top]
"""
import csv
import io
import subprocess
import sys


def main():
    rep, rx = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name",
                          f"regex:{rx}", "--launch-count", "1"], capture_output=True, text=True).stdout
    hdr, agg = None, {}
    for r in csv.reader(io.StringIO(out)):
        if r and r[0] == "Line No":
            hdr = r
            wi, ii = hdr.index("L1 Wavefronts Shared"), hdr.index("L1 Wavefronts Shared Ideal")
        elif hdr and r and r[0].isdigit() and len(r) > wi:
            try:
                w, i = float(r[wi] or 0), float(r[ii] or 0)
            except ValueError:
                continue
            if w > 0:
                a = agg.setdefault((r[0], r[1].strip()[:100]), [0.0, 0.0])
                a[0] += w
                a[1] += i
    tw = sum(a[0] for a in agg.values()) or 1
    ti = sum(a[1] for a in agg.values())
    print(f"shared wavefronts {tw:,.0f}  ideal {ti:,.0f}  ({tw / max(ti, 1):.2f}x)")
    for (ln, src), (w, i) in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        print(f"{w / tw * 100:5.1f}%  x{w / max(i, 1):4.1f}  {ln:>4s}  {src}")


if __name__ == "__main__":
    main()
