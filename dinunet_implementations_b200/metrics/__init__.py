"""Metrics that aggregate across sites as *sufficient statistics*.

The reference's trainers call ``self.new_metrics().add(pred_or_score, labels)`` and
``self.new_averages().add(value, n)`` every iteration (``comps/fs/__init__.py:58-61``,
``comps/icalstm/__init__.py:64-68``) and the aggregator rebuilds the same objects from
each site's serialised payload (``remote.py:29,32`` passes the Trainer class for exactly
that purpose).  Every class here therefore has:

* ``add(...)``        – fold one batch in (accepts tensors on any device, numpy, lists)
* ``accumulate(o)``   – merge another instance of the same class
* ``serialize()``     – JSON-safe payload
* ``reduce_sites(l)`` – merge a list of serialised payloads (what the remote calls)
* ``get()``           – list of rounded scalars in ``log_header`` order
* ``extract(name)``   – a single named score (``monitor_metric``)

Device tensors are kept on the device and only synchronised when a value is actually
read (``get``/``serialize``), so ``val.add(loss, n)`` with a device scalar does not force
the per-iteration ``.item()`` sync the reference pays.
"""
from .averages import COINNAverages
from .classification import Prf1a, ConfusionMatrix, AUCROCMetrics, COINNMetrics, roc_auc

__all__ = [
    "COINNAverages", "COINNMetrics", "Prf1a", "ConfusionMatrix", "AUCROCMetrics", "roc_auc",
]
