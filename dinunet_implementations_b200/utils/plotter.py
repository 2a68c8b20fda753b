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


def _quartiles(vals):
    v = sorted(vals)
    n = len(v)

    def q(f):
        if n == 1:
            return v[0]
        x = f * (n - 1)
        lo = int(x)
        hi = min(lo + 1, n - 1)
        return v[lo] + (v[hi] - v[lo]) * (x - lo)
    return v[0], q(0.25), q(0.5), q(0.75), v[-1]


def svg_boxplot(groups: dict, title: str = "", width=640, height=360, pad=50):
    """Box-and-whisker plot (min / quartiles / max, every point drawn) of ``{label: [values]}`` as stand-alone SVG —
    the per-fold accuracy / F1 comparison the reference draws with matplotlib (``NB.ipynb:96-110,156``)."""
    groups = {k: [float(x) for x in v if x is not None] for k, v in groups.items()}
    groups = {k: v for k, v in groups.items() if v}
    if not groups:
        return ""
    allv = [x for v in groups.values() for x in v]
    lo, hi = min(allv), max(allv)
    if hi - lo < 1e-12:
        lo, hi = lo - 0.5, hi + 0.5
    span = hi - lo
    lo, hi = lo - 0.05 * span, hi + 0.05 * span

    def y(v):
        return height - pad - (height - 2 * pad) * (v - lo) / (hi - lo)
    parts = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{width}" height="{height}">',
             f'<rect width="{width}" height="{height}" fill="white"/>',
             f'<text x="{width / 2:.0f}" y="18" font-size="13" text-anchor="middle">{title}</text>',
             f'<line x1="{pad}" y1="{height - pad}" x2="{width - pad / 2}" y2="{height - pad}" stroke="black"/>',
             f'<line x1="{pad}" y1="{pad / 2}" x2="{pad}" y2="{height - pad}" stroke="black"/>']
    for t in range(5):
        v = lo + (hi - lo) * t / 4
        parts.append(f'<text x="{pad - 4}" y="{y(v) + 3:.1f}" font-size="10" text-anchor="end">{v:.3g}</text>')
        parts.append(f'<line x1="{pad}" y1="{y(v):.1f}" x2="{width - pad / 2}" y2="{y(v):.1f}" stroke="#dddddd"/>')
    n = len(groups)
    slot = (width - 1.5 * pad) / n
    bw = min(60.0, slot * 0.5)
    colours = ["#1f77b4", "#d62728", "#2ca02c", "#9467bd", "#ff7f0e", "#8c564b"]
    for i, (name, vals) in enumerate(groups.items()):
        cx = pad + slot * (i + 0.5)
        mn, q1, med, q3, mx = _quartiles(vals)
        c = colours[i % len(colours)]
        parts.append(f'<line x1="{cx:.1f}" y1="{y(mn):.1f}" x2="{cx:.1f}" y2="{y(mx):.1f}" stroke="{c}"/>')
        parts.append(f'<rect x="{cx - bw / 2:.1f}" y="{y(q3):.1f}" width="{bw:.1f}" height="{max(y(q1) - y(q3), 1):.1f}" '
                     f'fill="{c}" fill-opacity="0.25" stroke="{c}"/>')
        parts.append(f'<line x1="{cx - bw / 2:.1f}" y1="{y(med):.1f}" x2="{cx + bw / 2:.1f}" y2="{y(med):.1f}" stroke="{c}" stroke-width="2"/>')
        for k, v in enumerate(vals):
            parts.append(f'<circle cx="{cx + (k % 5 - 2) * 3:.1f}" cy="{y(v):.1f}" r="2" fill="{c}"/>')
        parts.append(f'<text x="{cx:.1f}" y="{height - pad + 16}" font-size="11" text-anchor="middle">{name}</text>')
    parts.append("</svg>")
    return "\n".join(parts)
