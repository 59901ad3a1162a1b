#!/usr/bin/env python
"""Per CUDA source line: stall samples and executed warp instructions, from
`ncu -i REP --page source --csv --print-source cuda,sass [-k ...]` on stdin/file.
python profiles/ncu_lines.py FILE.csv [top_n]"""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    hdr, fname, out = None, "", []
    for r in rows:
        if r and r[0] == "File Path":
            fname = r[1].split("/")[-1]
            continue
        if r and r[0] == "Line No":
            hdr = r
            i_n, i_i = hdr.index("# Samples"), hdr.index("Instructions Executed")
            continue
        if hdr is None or len(r) <= i_i or r[0] == "":
            continue
        try:
            out.append((int(r[i_n]), int(r[i_i]), fname, r[0], r[1].strip()[:90]))
        except ValueError:
            pass
    ts, ti = sum(o[0] for o in out) or 1, sum(o[1] for o in out) or 1
    print("samples", ts, "warp instructions", ti)
    for o in sorted(out, reverse=True)[:top]:
        print("%5.1f%% smp %5.1f%% inst  %s:%s  %s" % (100 * o[0] / ts, 100 * o[1] / ti, o[2], o[3], o[4]))


if __name__ == "__main__":
    main()
