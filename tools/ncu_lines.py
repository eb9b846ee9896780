#!/usr/bin/env python
"""Top source lines of one kernel in an .ncu-rep by executed warp instructions and stall samples.

    python tools/ncu_lines.py prof.ncu-rep <kernel-regex> [launch-skip] [top]
(reads `ncu --page source --print-source cuda,sass --csv`; needs -lineinfo at compile time)
"""
import csv
import io
import subprocess
import sys


def main():
    rep, rx = sys.argv[1], sys.argv[2]
    skip = sys.argv[3] if len(sys.argv) > 3 else "0"
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name",
                          f"regex:{rx}", "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    data, fname, hdr = [], "", None
    for r in rows:
        if not r:
            continue
        if r[0] in ("File Name", "File Path"):
            fname = r[1].split("/")[-1]
        elif r[0] in ("Kernel Name", "Function Name"):
            print("kernel:", r[1])
        elif r[0] == "Line No":
            hdr = r
            ii, wi = hdr.index("Instructions Executed"), hdr.index("# Samples")
            stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        elif hdr and r[0].isdigit() and len(r) > ii and r[ii].isdigit():
            st = sorted(((int(r[i]) if r[i].isdigit() else 0, h) for i, h in stall_cols), reverse=True)[:2]
            data.append((int(r[ii]), int(r[wi]) if r[wi].isdigit() else 0, fname, r[0], r[1].strip(), st))
    tot, totw = sum(d[0] for d in data) or 1, sum(d[1] for d in data) or 1
    print(f"total warp-instructions {tot:,}  stall samples {totw:,}")
    for n, w, f, ln, s, st in sorted(data, reverse=True)[:top]:
        sts = " ".join(f"{h[6:]}={c}" for c, h in st if c)
        print(f"{n / tot * 100:5.1f}% inst {w / totw * 100:5.1f}% smp  {f}:{ln:>4s}  {s[:90]:90s} {sts}")


if __name__ == "__main__":
    main()
