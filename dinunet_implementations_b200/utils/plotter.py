"""Progress plots of train/validation logs.

The image has no matplotlib, so plots are written as stand-alone SVG (no dependency);
if matplotlib happens to be importable it is used to write a PNG next to the SVG.
"""
from __future__ import annotations

import os


def _svg_polyline(series, width=640, height=360, pad=40):
    vals = [v for s in series.values() for v in s if v is not None]
    if not vals:
        return ""
    lo, hi = min(vals), max(vals)
    if hi - lo < 1e-12:
        hi = lo + 1.0
    n = max(max(len(s) for s in series.values()), 2)
    colours = ["#1f77b4", "#d62728", "#2ca02c", "#9467bd", "#ff7f0e", "#8c564b"]
    parts = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{width}" height="{height}">',
             f'<rect width="{width}" height="{height}" fill="white"/>',
             f'<line x1="{pad}" y1="{height - pad}" x2="{width - pad}" y2="{height - pad}" stroke="black"/>',
             f'<line x1="{pad}" y1="{pad}" x2="{pad}" y2="{height - pad}" stroke="black"/>',
             f'<text x="4" y="{pad}" font-size="10">{hi:.4g}</text>',
             f'<text x="4" y="{height - pad}" font-size="10">{lo:.4g}</text>']
    for i, (name, s) in enumerate(series.items()):
        pts = []
        for j, v in enumerate(s):
            x = pad + (width - 2 * pad) * j / (n - 1)
            y = height - pad - (height - 2 * pad) * (v - lo) / (hi - lo)
            pts.append(f"{x:.1f},{y:.1f}")
        c = colours[i % len(colours)]
        parts.append(f'<polyline fill="none" stroke="{c}" stroke-width="1.5" points="{" ".join(pts)}"/>')
        parts.append(f'<text x="{width - pad - 150}" y="{pad + 14 * i}" font-size="11" fill="{c}">{name}</text>')
    parts.append("</svg>")
    return "\n".join(parts)


def plot_progress(cache: dict, log_dir: str, name: str = "progress", keys=("train_log", "validation_log"),
                  header: str | None = None):
    """One SVG per log key, one polyline per column of the log (columns named by header)."""
    os.makedirs(log_dir, exist_ok=True)
    header = header or cache.get("log_header", "loss")
    cols = [c for grp in str(header).split("|") for c in grp.split(",")]
    written = []
    for k in keys:
        rows = cache.get(k) or []
        if not rows:
            continue
        width = len(rows[0]) if isinstance(rows[0], (list, tuple)) else 1
        series = {}
        for ci in range(width):
            nm = cols[ci] if ci < len(cols) else f"col{ci}"
            series[f"{k}:{nm}"] = [float(r[ci]) if isinstance(r, (list, tuple)) else float(r) for r in rows]
        svg = _svg_polyline(series)
        if svg:
            p = os.path.join(log_dir, f"{name}_{k}.svg")
            with open(p, "w") as f:
                f.write(svg)
            written.append(p)
        try:  # optional
            import matplotlib
            matplotlib.use("Agg")
            import matplotlib.pyplot as plt
            plt.figure()
            for nm, s in series.items():
                plt.plot(s, label=nm)
            plt.legend()
            plt.savefig(os.path.join(log_dir, f"{name}_{k}.png"))
            plt.close()
        except Exception:
            pass
    return written
