"""Minimal stand-in for the ``coinstac`` Python runtime package (``entry.py:1,5``).

``coinstac.start(local_fn, remote_fn)`` is the only symbol the reference uses.  The real
package serves the two callbacks to the COINSTAC pipeline; it is not installable here
(SURVEY §0.2), so this shim offers two ways to drive them:

* ``DINUNET_SIMULATE=<data_path>`` (+ optional ``DINUNET_WORK_DIR``, ``DINUNET_SITES``):
  run the in-repo simulator on that simulator-layout directory, with ``local_fn`` shared by
  all sites in this process (fine for callbacks that keep no module-global state; the
  reference's ``local.py`` does, so for it use ``dinunet_implementations_b200.simulator``
  which loads one module instance per site).
* otherwise: a JSON-lines server on stdin/stdout — each line
  ``{"mode": "local"|"remote", "data": {"input": …, "state": …}}`` is answered with one
  line holding the callback's return value.  This is the seam a real COINSTAC container
  adapter plugs into.
"""
from __future__ import annotations

import json
import os
import sys


def start(local_fn, remote_fn):
    data_path = os.environ.get("DINUNET_SIMULATE")
    if data_path:
        from dinunet_implementations_b200.simulator import Simulator
        from dinunet_implementations_b200.utils import unwrap_inputspec
        with open(os.path.join(data_path, "inputspec.json")) as f:
            spec = json.load(f)
        n = int(os.environ.get("DINUNET_SITES", len(spec)))
        work = os.environ.get("DINUNET_WORK_DIR", os.path.join(os.getcwd(), "test"))
        inputs = [unwrap_inputspec(spec[i % len(spec)]) for i in range(n)]
        in_dirs = [os.path.join(data_path, "input", f"local{i}", "simulatorRun") for i in range(n)]
        sim = Simulator([local_fn] * n, remote_fn, inputs, work, in_dirs, verbose=True)
        return sim.run()
    from dinunet_implementations_b200.utils import clean_json
    for line in sys.stdin:
        line = line.strip()
        if not line:
            continue
        try:
            msg = json.loads(line)
            fn = remote_fn if msg.get("mode") == "remote" else local_fn
            res = fn(msg["data"])
            sys.stdout.write(json.dumps(clean_json(res)) + "\n")
        except Exception as e:  # report, keep serving
            sys.stdout.write(json.dumps({"error": repr(e)}) + "\n")
        sys.stdout.flush()
