"""Summarise finished runs — the job the reference's result notebooks do by hand.

``nnlogs.ipynb`` opens the ``*.zip`` of global results of one run per aggregation engine, reads
``fold_<k>/logs.json`` from each and prints ``agg_engine``, ``test_metrics`` and the summed
``time_spent_on_computation`` / last ``cumulative_total_duration`` side by side
(``nnlogs.ipynb:68-92``); ``NB.ipynb`` reads ``fold_<k>/test_metrics.csv`` over ten folds
(accuracy / F1 columns, ``NB.ipynb:96-110``) and ``best_val_epoch`` (``:49,112``).

    python -m dinunet_implementations_b200.tools.nnlogs RUN [RUN ...] [--json]

``--boxplot PREFIX`` additionally writes ``PREFIX_<metric>.svg`` box plots over the folds of every run (one box per
run: e.g. from-scratch vs pre-trained, ``NB.ipynb:96-110,156,189``) and prints the mean best-validation / early-stop
epoch per run (``NB.ipynb:200``).

``RUN`` is any of: a results zip, a node output directory (``…/output/remote/simulatorRun``), a
task directory holding ``fold_<k>/``, or a work directory (searched for the remote's output).
One table row per (run, fold), then mean ± std over folds per run.
"""
from __future__ import annotations

import argparse
import csv
import io
import json
import os
import statistics
import zipfile


def _fold_key(name):
    try:
        return int(str(name).split("_")[-1])
    except ValueError:
        return 1 << 30


def _read_dir(task_dir):
    folds = {}
    for d in sorted(os.listdir(task_dir), key=_fold_key):
        p = os.path.join(task_dir, d)
        if not (d.startswith("fold_") and os.path.isdir(p)):
            continue
        rec = {}
        lj = os.path.join(p, "logs.json")
        if os.path.exists(lj):
            with open(lj) as f:
                rec["logs"] = json.load(f)
        cj = os.path.join(p, "test_metrics.csv")
        if os.path.exists(cj):
            with open(cj, newline="") as f:
                rec["csv"] = list(csv.reader(f))
        if rec:
            folds[d] = rec
    return folds


def _read_zip(path):
    folds = {}
    with zipfile.ZipFile(path) as z:
        for n in z.namelist():
            parts = n.split("/")
            if len(parts) < 2 or not parts[-2].startswith("fold_"):
                continue
            rec = folds.setdefault(parts[-2], {})
            if parts[-1] == "logs.json":
                rec["logs"] = json.loads(z.read(n))
            elif parts[-1] == "test_metrics.csv":
                rec["csv"] = list(csv.reader(io.StringIO(z.read(n).decode())))
    return dict(sorted(folds.items(), key=lambda kv: _fold_key(kv[0])))


def _find_task_dirs(root):
    """Directories that directly contain ``fold_<k>`` sub-directories, remote's output preferred."""
    hits = []
    for base, dirs, _ in os.walk(root):
        if any(d.startswith("fold_") for d in dirs):
            hits.append(base)
            dirs[:] = []
    remote = [h for h in hits if f"{os.sep}remote{os.sep}" in h + os.sep]
    return remote or hits


def load_run(path):
    """``{fold_name: {'logs': dict, 'csv': rows}}`` for one run."""
    if os.path.isfile(path) and zipfile.is_zipfile(path):
        return _read_zip(path)
    if os.path.isdir(path):
        tds = _find_task_dirs(path)
        if tds:
            return _read_dir(tds[0])
        zips = [os.path.join(path, f) for f in os.listdir(path) if f.endswith(".zip")]
        if zips:
            return _read_zip(zips[0])
    raise FileNotFoundError(f"no fold_<k>/logs.json under {path!r}")


def summarise(path):
    rows = []
    for fold, rec in load_run(path).items():
        log = rec.get("logs", {})
        tm = (log.get("test_metrics") or [[None, None]])[0]
        row = {"run": path, "fold": fold, "agg_engine": log.get("agg_engine"), "test_loss": tm[0],
               "test_score": tm[1] if len(tm) > 1 else None, "best_val_epoch": log.get("best_val_epoch"),
               "compute_s": sum(log.get("time_spent_on_computation") or []) or None,
               "total_s": (log.get("cumulative_total_duration") or [None])[-1]}
        c = rec.get("csv")
        if c and len(c) >= 2:
            hdr = [h.lower() for h in c[0]]
            for nm in ("accuracy", "f1"):
                if nm in hdr:
                    try:
                        row[nm] = float(c[1][hdr.index(nm)])
                    except ValueError:
                        pass
        rows.append(row)
    return rows


def _fmt(v):
    if v is None:
        return "-"
    if isinstance(v, float):
        return f"{v:.4f}" if abs(v) < 100 else f"{v:.1f}"
    return str(v)


def render(all_rows):
    cols = ["agg_engine", "fold", "test_loss", "test_score", "accuracy", "f1", "best_val_epoch", "compute_s", "total_s"]
    lines = ["  ".join(f"{c:>14s}" for c in cols)]
    by_run = {}
    for r in all_rows:
        by_run.setdefault(r["run"], []).append(r)
        lines.append("  ".join(f"{_fmt(r.get(c)):>14s}" for c in cols))
    lines.append("")
    for run, rs in by_run.items():
        parts = []
        for c in ("test_loss", "test_score", "accuracy", "f1", "total_s"):
            vals = [r[c] for r in rs if isinstance(r.get(c), (int, float))]
            if vals:
                sd = statistics.pstdev(vals) if len(vals) > 1 else 0.0
                parts.append(f"{c} {statistics.mean(vals):.4f}±{sd:.4f}")
        lines.append(f"{rs[0].get('agg_engine')} [{len(rs)} fold(s)] {run}: " + ", ".join(parts))
    return "\n".join(lines)


def boxplots(all_rows, prefix, labels=None, metrics=("accuracy", "f1", "test_score", "best_val_epoch")):
    """One SVG per metric, one box per run over its folds.  Returns the written paths and the per-run mean
    best-validation epoch (the scratch-vs-pretrain early-stop comparison of ``NB.ipynb:200``)."""
    from ..utils.plotter import svg_boxplot
    runs = []
    for r in all_rows:
        if r["run"] not in runs:
            runs.append(r["run"])
    names = {run: (labels[i] if labels and i < len(labels) else
                   (os.path.basename(os.path.normpath(run)) or f"run{i}")) for i, run in enumerate(runs)}
    written = []
    os.makedirs(os.path.dirname(os.path.abspath(prefix)) or ".", exist_ok=True)
    for m in metrics:
        groups = {names[run]: [r[m] for r in all_rows if r["run"] == run and isinstance(r.get(m), (int, float))] for run in runs}
        svg = svg_boxplot(groups, title=f"{m} over folds")
        if svg:
            p = f"{prefix}_{m}.svg"
            with open(p, "w") as f:
                f.write(svg)
            written.append(p)
    epochs = {}
    for run in runs:
        ev = [r["best_val_epoch"] for r in all_rows if r["run"] == run and isinstance(r.get("best_val_epoch"), (int, float))]
        if ev:
            epochs[names[run]] = statistics.mean(ev)
    return written, epochs


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("runs", nargs="+")
    ap.add_argument("--json", action="store_true")
    ap.add_argument("--boxplot", metavar="PREFIX", help="write PREFIX_<metric>.svg box plots over the folds of every run")
    ap.add_argument("--labels", default=None, help="comma-separated run labels for the plots (e.g. scratch,pretrained)")
    a = ap.parse_args(argv)
    rows = [r for p in a.runs for r in summarise(p)]
    print(json.dumps(rows) if a.json else render(rows))
    if a.boxplot:
        written, epochs = boxplots(rows, a.boxplot, a.labels.split(",") if a.labels else None)
        print("box plots:", ", ".join(written))
        print("mean best-validation epoch per run:", ", ".join(f"{k} {v:.1f}" for k, v in epochs.items()))
    return rows


if __name__ == "__main__":
    main()
