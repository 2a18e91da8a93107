#!/usr/bin/env python
"""Per-source-line instruction counts and stall samples of one kernel from an .ncu-rep
(needs -lineinfo and --import-source on).  usage: src_hotspots.py report.ncu-rep kernel-regex [top]"""
import csv
import io
import subprocess
import sys


def main(rep, kern, top=40):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass,cuda", "--csv",
                          "--kernel-name", f"regex:{kern}"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    agg = {}
    hdr = None
    fname = ""
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            fname = r[1].split("/")[-1]
        if len(r) > 8 and r[0] == "Line No":
            hdr = r
            ie = hdr.index("Instructions Executed")
            sm = hdr.index("# Samples")
            continue
        if hdr and len(r) == len(hdr) and r[2] == "-":      # a source line row (SASS rows carry an address)
            try:
                n, s = int(r[ie]), int(r[sm])
            except ValueError:
                continue
            key = (fname, int(r[0]))
            a = agg.setdefault(key, [0, 0, r[1].strip()])
            a[0] += n
            a[1] += s
    tot = sum(v[0] for v in agg.values()) or 1
    tots = sum(v[1] for v in agg.values()) or 1
    print(f"# {rep} {kern}: {tot} warp instructions, {tots} samples")
    for (f, ln), (n, s, src) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{n:10d} {100 * n / tot:5.1f}%  samples {100 * s / tots:5.1f}%  {f}:{ln}: {src[:100]}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40)
