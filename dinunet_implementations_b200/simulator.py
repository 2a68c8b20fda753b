"""In-repo stand-in for ``coinstac-simulator`` (SURVEY §4 "fake backend *is* the simulator").

Runs a COINSTAC computation (any object with ``local.run(data)`` / ``remote.run(data)``
callbacks — in particular the reference's own ``local.py`` / ``remote.py``, unmodified)
for N sites + 1 aggregator on one machine, relaying the per-round JSON and the files in the
nodes' transfer directories exactly the way the runtime does:

* site ``transferDirectory``  →  aggregator ``baseDirectory/<site_id>/``
* aggregator ``transferDirectory``  →  every site's ``baseDirectory``

Input layout (``datasets/test_fsl``): ``<data>/inputspec.json`` (one ``{"k": {"value": v}}``
object per site) and ``<data>/input/local<i>/simulatorRun``.  Because the reference tree is
read-only, all mutable directories (a link-farm copy of each site's input dir, output,
cache, transfer) live under ``work_dir``; outputs follow the reference layout
``<work>/output/local<i>/simulatorRun`` and ``<work>/output/remote/simulatorRun``
(``NB.ipynb:33,851,885``).

Each site gets its *own instance* of the ``local`` module (unique module name), because
the reference keeps per-site state in module globals (``local.py:15-16``).
"""
from __future__ import annotations

import importlib.util
import json
import os
import shutil
import sys
import time

from .utils import unwrap_inputspec, clean_json, logger


def _load_module(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def load_computation(comp_dir, num_sites, tag=None):
    """Import ``<comp_dir>/local.py`` once per site and ``remote.py`` once."""
    comp_dir = os.path.abspath(comp_dir)
    if comp_dir not in sys.path:
        sys.path.insert(0, comp_dir)
    tag = tag or f"sim{int(time.time() * 1e6) % 10 ** 9}"
    locals_ = [_load_module(os.path.join(comp_dir, "local.py"), f"_{tag}_local{i}") for i in range(num_sites)]
    remote = _load_module(os.path.join(comp_dir, "remote.py"), f"_{tag}_remote")
    return [m.run for m in locals_], remote.run, locals_, remote


def _link_tree(src, dst):
    os.makedirs(dst, exist_ok=True)
    if not os.path.isdir(src):
        return
    for name in os.listdir(src):
        s, d = os.path.join(src, name), os.path.join(dst, name)
        if os.path.lexists(d):
            continue
        try:
            os.symlink(s, d)
        except OSError:
            if os.path.isdir(s):
                shutil.copytree(s, d)
            else:
                shutil.copy(s, d)


def _move_all(src_dir, dst_dir):
    if not os.path.isdir(src_dir):
        return
    os.makedirs(dst_dir, exist_ok=True)
    for name in os.listdir(src_dir):
        s, d = os.path.join(src_dir, name), os.path.join(dst_dir, name)
        if os.path.lexists(d):
            os.remove(d)
        shutil.move(s, d)


def _copy_all(src_dir, dst_dirs, remove=True):
    if not os.path.isdir(src_dir):
        return
    for name in os.listdir(src_dir):
        s = os.path.join(src_dir, name)
        for dd in dst_dirs:
            os.makedirs(dd, exist_ok=True)
            d = os.path.join(dd, name)
            if os.path.lexists(d):
                os.remove(d)
            shutil.copy(s, d)
        if remove:
            os.remove(s)


class Simulator:
    def __init__(self, local_runs, remote_run, site_inputs, work_dir, site_input_dirs=None, verbose=False):
        self.local_runs = list(local_runs)
        self.remote_run = remote_run
        self.n = len(self.local_runs)
        self.work_dir = os.path.abspath(work_dir)
        self.verbose = verbose
        self.site_ids = [f"local{i}" for i in range(self.n)]
        self.site_inputs = [dict(x) for x in site_inputs]
        self.states = []
        for i, sid in enumerate(self.site_ids):
            base = os.path.join(self.work_dir, "input", sid, "simulatorRun")
            if site_input_dirs and site_input_dirs[i]:
                _link_tree(site_input_dirs[i], base)
            else:
                os.makedirs(base, exist_ok=True)
            self.states.append(self._state(sid, base))
        self.remote_state = self._state("remote", os.path.join(self.work_dir, "input", "remote", "simulatorRun"))
        self.rounds = 0
        self.history = []

    def _state(self, node, base):
        st = {"clientId": node, "baseDirectory": base, "iteration": 0}
        for k, sub in (("outputDirectory", "output"), ("cacheDirectory", "cache"), ("transferDirectory", "transfer")):
            st[k] = os.path.join(self.work_dir, sub, node, "simulatorRun")
            os.makedirs(st[k], exist_ok=True)
        os.makedirs(base, exist_ok=True)
        return st

    @classmethod
    def from_data_path(cls, comp_dir, data_path, work_dir, num_sites=None, overrides=None, verbose=False):
        """Reference-style launch: computation dir + simulator data dir."""
        comp_dir, data_path, work_dir = (os.path.abspath(p) for p in (comp_dir, data_path, work_dir))   # inputs are sym-linked
        with open(os.path.join(data_path, "inputspec.json")) as f:
            spec = json.load(f)
        spec = spec if isinstance(spec, list) else [spec]
        n = int(num_sites or len(spec))
        inputs = []
        for i in range(n):
            inp = unwrap_inputspec(spec[i % len(spec)])
            inp.update(overrides or {})
            inputs.append(inp)
        in_dirs = [os.path.join(data_path, "input", f"local{i}", "simulatorRun") for i in range(n)]
        lruns, rrun, _, _ = load_computation(comp_dir, n)
        return cls(lruns, rrun, inputs, work_dir, in_dirs, verbose)

    def step(self):
        """One COINSTAC pipeline iteration: all sites, then the aggregator."""
        self.rounds += 1
        outs = {}
        for i, sid in enumerate(self.site_ids):
            st = dict(self.states[i], iteration=self.rounds)
            res = self.local_runs[i]({"input": self.site_inputs[i], "state": st})
            outs[sid] = json.loads(json.dumps(clean_json(res["output"])))  # wire = JSON
            _move_all(st["transferDirectory"], os.path.join(self.remote_state["baseDirectory"], sid))
        rst = dict(self.remote_state, iteration=self.rounds)
        rres = self.remote_run({"input": outs, "state": rst})
        rout = json.loads(json.dumps(clean_json(rres["output"])))
        _copy_all(rst["transferDirectory"], [s["baseDirectory"] for s in self.states])
        self.site_inputs = [dict(rout) for _ in self.site_ids]
        self.history.append({"round": self.rounds, "phase": rout.get("phase"),
                             "modes": rout.get("global_modes")})
        if self.verbose:
            logger.info(f"round {self.rounds}: phase={rout.get('phase')} modes={rout.get('global_modes')}")
        return bool(rres.get("success")), rout

    def run(self, max_rounds=100000):
        done, rout = False, {}
        while not done and self.rounds < max_rounds:
            done, rout = self.step()
        if not done:
            raise RuntimeError(f"simulation did not reach SUCCESS within {max_rounds} rounds")
        return rout


def main(argv=None):
    import argparse
    ap = argparse.ArgumentParser(description="run a COINSTAC computation on N simulated sites")
    ap.add_argument("--comp-dir", required=True, help="directory with local.py / remote.py")
    ap.add_argument("--data-path", required=True, help="simulator data dir (inputspec.json + input/)")
    ap.add_argument("--work-dir", required=True)
    ap.add_argument("--sites", type=int, default=None)
    ap.add_argument("--set", action="append", default=[], help="override key=json_value for all sites")
    ap.add_argument("-v", "--verbose", action="store_true")
    a = ap.parse_args(argv)
    ov = {}
    for kv in a.set:
        k, v = kv.split("=", 1)
        try:
            ov[k] = json.loads(v)
        except json.JSONDecodeError:
            ov[k] = v
    sim = Simulator.from_data_path(a.comp_dir, a.data_path, a.work_dir, a.sites, ov, a.verbose)
    out = sim.run()
    print(json.dumps({"rounds": sim.rounds, "final": clean_json(out)}))


if __name__ == "__main__":
    main()
