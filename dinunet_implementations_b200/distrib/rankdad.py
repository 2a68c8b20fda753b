"""rankDAD in file mode (engine name ``comps/__init__.py:15``).

For every ``nn.Linear`` the site never ships the ``out×in`` weight gradient.  Hooks capture
the layer input ``A [N,in]`` (with a ones column appended when the layer has a bias, so the
bias gradient rides along as the last column of ``ΔᵀA``) and the output delta ``Δ [N,out]``;
``lowrank.dad_compress`` turns the pair into rank-``dad_reduction_rank`` factors.  A module
that is called several times per step (the LSTM cell's ``i2h``/``h2h`` inside the time loop,
``comps/icalstm/models.py:31``) contributes one ``(A, Δ)`` block per call; they are stacked
along ``N``.  Parameters that do not belong to a Linear (BatchNorm scale/shift) travel dense.

Aggregator: stack all sites' factors → ``[S·r, ·]`` → re-compress to rank ``r`` → broadcast.
Site: ``W.grad = Δ_rᵀ·A_r`` and ``optimizer.step()``.

Tunables: ``dad_reduction_rank`` / ``dad_num_pow_iters`` / ``dad_tol``
(``compspec.json:236-238,268-270``).
"""
from __future__ import annotations

import numpy as np
import torch

from . import COINNLearner, COINNReducer, fileio
from . import lowrank as lr
from ..utils.profiler import timed

DAD_FILE = "dad_factors.npz"
DAD_AVG_FILE = "dad_reduced.npz"


class _LinearTap:
    """Forward hook on one ``nn.Linear`` that also taps the gradient of its output.

    Every call of the module registers its own gradient hook that remembers that call's input, so an
    activation is always paired with the delta of the same call — whatever the order in which autograd
    visits the calls (time loops, ``local_iterations`` > 1 micro-batches, a reused module)."""

    def __init__(self, module: torch.nn.Linear):
        self.module = module
        self.pairs = []
        self.calls = 0
        self.h1 = module.register_forward_hook(self._fwd)

    def _fwd(self, mod, inp, out):
        if mod.training and torch.is_grad_enabled() and out.requires_grad:
            x = inp[0].detach()
            x = x.reshape(-1, x.shape[-1])
            self.calls += 1

            def _bwd(g, x=x):
                g = g.detach()
                self.pairs.append((g.reshape(-1, g.shape[-1]), x))
            out.register_hook(_bwd)

    def pop(self, with_ones=True):
        """``(delta [N,out], act [N,in(+1)])`` of every call since the last pop, stacked along N (fp32);
        ``(None, None)`` when the module was not used."""
        pairs, self.pairs, self.calls = self.pairs, [], 0
        if not pairs:
            return None, None
        D = torch.cat([d for d, _ in pairs], 0).float()
        A = torch.cat([a for _, a in pairs], 0).float()
        if with_ones and self.module.bias is not None:
            A = torch.cat([A, torch.ones(A.shape[0], 1, device=A.device)], 1)
        return D, A

    def pop_entries(self):
        """Device-mode protocol: list of ``{"D": delta, "A": act}`` blocks (no ones column)."""
        pairs, self.pairs, self.calls = self.pairs, [], 0
        return [{"D": d, "A": a} for d, a in pairs]

    def remove(self):
        self.h1.remove()


class CapturedTap:
    """Stands in for ``_LinearTap`` when a fused model hands over the (delta, activation) pair of a
    Linear directly: its backward appends ``{"D","A"[,"DT","AT"]}`` blocks to ``model._dad_pairs[key]``
    (one per backward call, so micro-batches stack), and no module hooks or eager forward are needed."""

    def __init__(self, module: torch.nn.Linear, model: torch.nn.Module, key: str):
        self.module, self.model, self.key = module, model, key

    def pop_entries(self):
        pairs = getattr(self.model, "_dad_pairs", None) or {}
        return pairs.pop(self.key, [])

    def pop(self, with_ones=True):
        ent = self.pop_entries()
        if not ent:
            return None, None
        D = torch.cat([e["D"].float() for e in ent], 0)
        A = torch.cat([e["A"].float() for e in ent], 0)
        if with_ones and self.module.bias is not None:
            A = torch.cat([A, torch.ones(A.shape[0], 1, device=A.device)], 1)
        return D, A

    def remove(self):
        pass


def make_taps(nn_modules):
    """``(taps, covered_param_ids)`` for every trainable ``nn.Linear`` of the given models.  A model whose
    fused path bypasses its Linear modules (custom autograd Functions: hooks would never fire) either hands
    the pairs over itself (``dad_capture_keys()`` → ``CapturedTap``) or is switched to its eager path."""
    taps, covered = [], set()
    for m in nn_modules:
        # the fused path only ever runs on a CUDA device with the native library loaded; on a CPU site the model takes
        # its eager path whatever `fused` says, and only module hooks see the pairs
        on_gpu = any(p.is_cuda for p in m.parameters())
        if on_gpu:
            from ..ops import native
            on_gpu = native.available()
        keys = set(m.dad_capture_keys()) if (on_gpu and getattr(m, "fused", False) and hasattr(m, "dad_capture_keys")) else set()
        if keys:
            m._dad_capture = True
        elif hasattr(m, "fused"):
            m.fused = False
        for name, mod in m.named_modules():
            if isinstance(mod, torch.nn.Linear) and mod.weight.requires_grad:
                taps.append(CapturedTap(mod, m, name) if name in keys else _LinearTap(mod))
                covered.add(id(mod.weight))
                if mod.bias is not None:
                    covered.add(id(mod.bias))
    return taps, covered


class DADLearner(COINNLearner):
    def __init__(self, **kw):
        super().__init__(**kw)
        self.rank = int(self.cache.get("dad_reduction_rank", 10))
        self.iters = int(self.cache.get("dad_num_pow_iters", 5))
        self.tol = float(self.cache.get("dad_tol", 1e-3))
        self.taps, covered = make_taps(self.trainer.nn.values())
        self.dense_params = [p for p in self.params if id(p) not in covered]

    @timed("learner_step")
    def step(self):
        name = self.input.get("dad_reduced_file")
        if not name:
            return {}
        arrs = fileio.load_arrays(self.incoming_path(name))
        for i, tap in enumerate(self.taps):
            w = tap.module.weight
            d = torch.from_numpy(arrs[f"d{i}"]).to(w.device)
            a = torch.from_numpy(arrs[f"a{i}"]).to(w.device)
            G = lr.dad_reconstruct(d, a)
            if tap.module.bias is not None:
                tap.module.bias.grad = G[:, -1].contiguous().to(w.dtype)
                G = G[:, :-1]
            w.grad = G.contiguous().to(w.dtype)
        if self.dense_params:
            flat = torch.from_numpy(arrs["dense"])
            off = 0
            for p in self.dense_params:
                n = p.numel()
                p.grad = flat[off:off + n].view_as(p).to(device=p.device, dtype=p.dtype).clone()
                off += n
        self.optimizer_step()
        return {}

    @timed("learner_to_reduce")
    def to_reduce(self):
        its = self.backward()
        if not its:
            return {}, its
        arrs = {}
        seed = int(self.cache.get("seed", 0))
        for i, tap in enumerate(self.taps):
            D, A = tap.pop()
            w = tap.module.weight
            if D is None:
                # a Linear that took no part in this step's graph: its gradient is genuinely zero.  Anything else
                # (hooks bypassed by a fused path, say) must not be papered over with zero factors.
                if w.grad is not None and bool((w.grad != 0).any()):
                    raise RuntimeError(f"rankDAD: no (delta, activation) pair was captured for Linear #{i} "
                                       f"{tuple(w.shape)} although it has a gradient")
                n_in = w.shape[1] + (1 if tap.module.bias is not None else 0)
                D, A = torch.zeros(1, w.shape[0]), torch.zeros(1, n_in)
            d, a, _ = lr.dad_compress(D, A, self.rank, self.iters, self.tol, seed + i)
            arrs[f"d{i}"] = d.cpu().numpy()
            arrs[f"a{i}"] = a.cpu().numpy()
        if self.dense_params:
            arrs["dense"] = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1).float().cpu()
                                       for p in self.dense_params]).numpy()
        name = fileio.save_arrays(self.transfer_path(DAD_FILE), arrs, self.cache.get("precision_bits", "32"))
        return {"dad_file": name, "reduce": True}, its


def _reduce_layer(ds, as_, rank, iters, tol, seed):
    d, a, _ = lr.dad_reduce([torch.from_numpy(x) for x in ds], [torch.from_numpy(x) for x in as_],
                            rank, iters, tol, seed, average=True)
    return d.numpy(), a.numpy()


class DADReducer(COINNReducer):
    @timed("reducer_reduce")
    def reduce(self):
        sites = sorted(self.input.keys())
        payloads = [fileio.load_arrays(self.site_file(s, self.input[s]["dad_file"])) for s in sites]
        rank = int(self.cache.get("dad_reduction_rank", 10))
        iters = int(self.cache.get("dad_num_pow_iters", 5))
        tol = float(self.cache.get("dad_tol", 1e-3))
        seed = int(self.cache.get("seed", 0))
        n_layers = sum(1 for k in payloads[0] if k.startswith("d") and k[1:].isdigit())
        jobs = [([p[f"d{i}"] for p in payloads], [p[f"a{i}"] for p in payloads], rank, iters, tol, seed + 101 * i)
                for i in range(n_layers)]
        if self.mp_pool is not None and n_layers > 1:
            results = self.mp_pool.starmap(_reduce_layer, jobs)
        else:
            results = [_reduce_layer(*j) for j in jobs]
        out = {}
        for i, (d, a) in enumerate(results):
            out[f"d{i}"], out[f"a{i}"] = d, a
        if "dense" in payloads[0]:
            out["dense"] = np.mean(np.stack([p["dense"] for p in payloads]), 0)
        name = fileio.save_arrays(self.transfer_path(DAD_AVG_FILE), out, self.cache.get("precision_bits", "32"))
        return {"dad_reduced_file": name, "update": True}
