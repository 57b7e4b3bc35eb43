"""Prints the per-CUDA-source-line totals of an `ncu --page source --print-source cuda,sass --csv` dump.
    ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:K > dump.csv
    python tools/ncu_lines.py dump.csv [min_pct]
Development aid: shows where a kernel's executed instructions and stall samples go."""
import csv
import sys


def main():
    path = sys.argv[1]
    min_pct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    hdr, lines, fname = None, [], ""
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and r[0].isdigit():
            ie, isamp = hdr.index("Instructions Executed"), hdr.index("# Samples")
            try:
                lines.append((fname, int(r[0]), r[1].strip()[:105], float(r[ie] or 0), float(r[isamp] or 0)))
            except ValueError:
                pass
    ti = sum(l[3] for l in lines) or 1.0
    ts = sum(l[4] for l in lines) or 1.0
    print("total warp instructions %.0f, stall samples %.0f" % (ti, ts))
    for f, n, src, i, s in lines:
        if 100 * i / ti >= min_pct or 100 * s / ts >= min_pct:
            print("%5.1f%% inst %5.1f%% smp  %s:%d | %s" % (100 * i / ti, 100 * s / ts, f, n, src))


if __name__ == "__main__":
    main()
