"""`-p`: render plots from a results directory (the reference runs `benchmarks.utils.plot --data-dir`,
run-benchmarks.sh:70-72).  No plotting library is assumed: the charts are hand-written SVG."""
from __future__ import annotations

import json
import os
from typing import List, Tuple


def _svg_scatter(points: List[Tuple[float, float]], title: str, xlabel: str, ylabel: str, hline=None) -> str:
    W, H, L, B, T, R = 720, 420, 70, 50, 40, 20
    xs, ys = [p[0] for p in points] or [0, 1], [p[1] for p in points] or [0, 1]
    x0, x1 = min(xs), max(xs) if max(xs) > min(xs) else min(xs) + 1
    y0, y1 = 0.0, max(ys + ([hline[0]] if hline else [])) * 1.08 or 1.0
    sx = lambda x: L + (x - x0) / (x1 - x0) * (W - L - R)
    sy = lambda y: H - B - (y - y0) / (y1 - y0) * (H - B - T)
    out = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{W}" height="{H}" font-family="sans-serif" font-size="12">',
           f'<rect width="{W}" height="{H}" fill="white"/>', f'<text x="{W / 2}" y="20" text-anchor="middle" font-size="14">{title}</text>',
           f'<line x1="{L}" y1="{H - B}" x2="{W - R}" y2="{H - B}" stroke="black"/>', f'<line x1="{L}" y1="{T}" x2="{L}" y2="{H - B}" stroke="black"/>',
           f'<text x="{W / 2}" y="{H - 10}" text-anchor="middle">{xlabel}</text>',
           f'<text x="15" y="{H / 2}" text-anchor="middle" transform="rotate(-90 15 {H / 2})">{ylabel}</text>']
    for i in range(5):
        yv = y0 + (y1 - y0) * i / 4
        out.append(f'<text x="{L - 6}" y="{sy(yv) + 4}" text-anchor="end">{yv:.3g}</text>')
        out.append(f'<line x1="{L}" y1="{sy(yv)}" x2="{W - R}" y2="{sy(yv)}" stroke="#ddd"/>')
        xv = x0 + (x1 - x0) * i / 4
        out.append(f'<text x="{sx(xv)}" y="{H - B + 16}" text-anchor="middle">{xv:.4g}</text>')
    if hline:
        out.append(f'<line x1="{L}" y1="{sy(hline[0])}" x2="{W - R}" y2="{sy(hline[0])}" stroke="red" stroke-dasharray="6 4"/>')
        out.append(f'<text x="{W - R}" y="{sy(hline[0]) - 4}" text-anchor="end" fill="red">{hline[1]}</text>')
    for x, y in points:
        out.append(f'<circle cx="{sx(x):.1f}" cy="{sy(y):.1f}" r="3" fill="#1f77b4" fill-opacity="0.7"/>')
    out.append("</svg>")
    return "\n".join(out)


def plot_dir(data_dir: str) -> List[str]:
    """One throughput chart and one added-TTFT chart per results.json found under data_dir."""
    written = []
    for root, _, files in os.walk(data_dir):
        if "results.json" not in files:
            continue
        res = json.load(open(os.path.join(root, "results.json")))
        if "decode_ranks" not in res:  # an HTTP run (kvx/httpload.py): time to first token against the prompt length
            pts = [(q["prompt_tokens"], 1e3 * q["ttft_s"]) for q in res.get("requests", []) if q.get("prompt_tokens") and "ttft_s" in q]
            if pts:
                path = os.path.join(root, "time_to_first_token_ms.svg")
                with open(path, "w") as f:
                    f.write(_svg_scatter(pts, f"{res['summary'].get('endpoint', 'endpoint')}: time to first token",
                                         "prompt tokens", "time to first token (ms)"))
                written.append(path)
            continue
        reqs = [q for r in res["decode_ranks"] for q in r["requests"]]
        thr = [(q["isl"], q["throughput_GBps"]) for q in reqs]
        ttft = [(q["isl"], q["time_to_first_token_added_ms"]) for q in reqs]
        for name, pts, yl, hl in (("throughput_GBps.svg", thr, "KV transfer GB/s per request", (900.0, "NVLink 5 per direction (nominal)")),
                                  ("added_ttft_ms.svg", ttft, "added TTFT (ms)", None)):
            path = os.path.join(root, name)
            with open(path, "w") as f:
                f.write(_svg_scatter(pts, f"{res['summary']['topology']['name']}: {yl}", "ISL (tokens)", yl, hl))
            written.append(path)
    return written
