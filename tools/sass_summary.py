#!/usr/bin/env python
"""Opcode histogram per kernel of libkvx.so (cuobjdump -sass), the mnemonics that show which hardware path a kernel
uses: UBLKCP (1-d bulk copy, TMA unit), UTMASTG/UTMALDG (tensor-map TMA), SYNCS (mbarrier), LDS/STS/LDG/STG (thread
copies), PREEXIT/ACQBULK (programmatic dependent launch), MEMBAR/ATOM/RED (completion).   python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "dynamo-k8s-llm-inference_b200", "lib", "libkvx.so")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
WATCH = ("UBLKCP", "UTMASTG", "UTMALDG", "UTMAPF", "SYNCS", "LDS", "STS", "LDG", "STG", "LD.", "ST.", "PREEXIT", "ACQBULK", "MEMBAR", "ATOM", "RED",
         "UTCMMA", "UTCHMMA", "HMMA", "TCGEN", "NANOSLEEP", "CCTL", "ERRBAR", "BAR.")
kernels = collections.OrderedDict()
cur, arch = None, None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
        kernels[cur] = collections.Counter()
        continue
    m = re.search(r"arch = (sm_\w+)", line)
    if m:
        arch = m.group(1)
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        kernels[cur][m.group(1)] += 1
print(f"library: {os.path.relpath(LIB, ROOT)}   cubin arch: {arch}   kernels: {len(kernels)}")
print("(counts of SASS instructions per kernel; only the mnemonics that identify the data path are listed, then the total)\n")
for name, ops in kernels.items():
    short = re.sub(r"\(.*", "", name)
    total = sum(ops.values())
    groups = collections.OrderedDict()
    for op, c in sorted(ops.items()):
        for w in WATCH:
            if op.startswith(w):
                groups[op] = c
                break
    print(f"{short}\n    total {total}: " + ", ".join(f"{k} x{v}" for k, v in groups.items()) + "\n")
