#!/usr/bin/env python
"""Opcode histogram per kernel of libgs2d.so from `cuobjdump -sass` (no GPU needed).

    python profiles/sass_summary.py [lib.so] > profiles/sass_summary.txt

Per kernel: instruction count and the opcodes that say what the kernel is built from --
packed fp32 (FFMA2 / FMUL2 / FADD2), cp.async (LDGSTS), TMA / bulk copies (UBLKCP / UTMALDG),
reductions (RED / REDG / ATOMG), programmatic dependent launch (ACQBULK / PREEXIT), 16-byte
global accesses (LDG.E.128 / STG.E.128), MUFU by kind, barriers, shuffles, multimem."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gsvc_b200", "lib", "libgs2d.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = lambda s: subprocess.run(["c++filt", s], capture_output=True, text=True).stdout.strip()

WATCH = ["FFMA2", "FMUL2", "FADD2", "FFMA", "FMUL", "FADD", "MUFU.EX2", "MUFU.RCP", "MUFU.RSQ", "MUFU.SQRT", "MUFU.LG2",
         "MUFU.TANH", "LDGSTS", "UBLKCP", "UTMALDG", "SYNCS", "LDG.E.128", "LDG.E.64", "LDG.E.U8", "LDG.E", "STG.E.128",
         "STG.E", "LDS.128", "LDS", "STS.128", "STS", "RED", "REDG", "ATOMG", "ATOM", "ATOMS", "ACQBULK", "PREEXIT",
         "BAR.SYNC", "SHFL", "VOTE", "MATCH", "MULTIMEM", "LDGMC", "REDUX", "HMMA", "UTCHMMA"]

kern = None
hist = collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = m.group(1)
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and kern:
        op = m.group(1)
        hist[kern]["_total"] += 1
        hist[kern][op] += 1

print(f"# {os.path.relpath(lib, ROOT)}: SASS opcode summary per kernel (cuobjdump -sass; sm_100a)")
for k, h in hist.items():
    name = re.sub(r"\(.*", "", demangle(k))
    tot = h.pop("_total", 0)
    picks = []
    for w in WATCH:
        c = sum(v for op, v in h.items() if op == w or op.startswith(w + "."))
        if c:
            picks.append(f"{w} {c}")
    print(f"{name}: {tot} instructions | " + ", ".join(picks))
