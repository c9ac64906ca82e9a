#!/usr/bin/env python
"""Compare two tools/overhead.py outputs point by point (best of the repeated runs in each file)."""
import json
import sys


def load(path):
    best = {}
    for line in open(path):
        d = json.loads(line)
        key = (d["exp"], d.get("case", ""), d.get("n"), d.get("pool_blocks"))
        if key not in best or d["us"] < best[key]:
            best[key] = d["us"]
    return best


a, b = load(sys.argv[1]), load(sys.argv[2])
for k in a:
    if k in b:
        print(f"{str(k):70s} {a[k]:9.2f} {b[k]:9.2f}  {100 * (b[k] / a[k] - 1):+6.2f}%")
