"""Device-mode rankDAD and powerSGD: the low-rank engines with
every payload kept in device memory and exchanged through the symmetric arenas.

Math = ``distrib/lowrank.py`` (the oracle).  What runs where:

* rankDAD — ``csrc/dad.cu``: one launch compresses every tapped layer (per-layer cluster subspace iteration on
  tcgen05 / TMA) and all-gathers the factors into every peer's slot from its tail; a second launch recompresses the
  stacked factors and runs the reconstruction GEMM into the flat gradient buffer;
* powerSGD — ``csrc/psgd.cu``: three fused streaming passes per gradient matrix (``E += grad`` with ``P = M·Q``;
  ``Q = Mᵀ·P̂``; on-the-fly ``M̂ = P̂·Q̄ᵀ`` with ``E = M − M̂``), the two all-reduces through the reduce-only mode of the
  fused dSGD kernel (peer loads over NVLink, deterministic order ⇒ bit-identical factors on all sites), batched
  orthogonalisation ``dnb_orthogonalize_batch``;
* optimiser — the fused Adam/SGD kernel on the site's flat parameter buffer.

Every object here drives all sites that live in this process: one with ``torchrun`` (ipc), all
``S`` simulated sites with ``SiteGroup.local`` (tests on one GPU).
"""
from __future__ import annotations

import ctypes
import os

import torch

from ..distrib import lowrank
from ..ops import native
from .comm import SiteGroup
from .dsgd_fused import FusedDSGD, flat_layout


def _r8(n):
    return (int(n) + 7) // 8 * 8


def orthogonalize_batch(mats, eps: float = 1e-8):
    """In-place column orthonormalisation of fp32 matrices ``[n, r]`` (unit inner stride)."""
    if not mats:
        return
    lib = native.require_gpu()
    k = len(mats)
    P = (ctypes.c_void_p * k)(*[m.data_ptr() for m in mats])
    n = (ctypes.c_int * k)(*[m.shape[0] for m in mats])
    r = (ctypes.c_int * k)(*[m.shape[1] for m in mats])
    ld = (ctypes.c_int * k)(*[m.stride(0) for m in mats])
    for m in mats:
        assert m.dtype == torch.float32 and m.stride(1) == 1
    native.check(lib.dnb_orthogonalize_batch(P, n, r, ld, ctypes.c_int(k), ctypes.c_float(eps), native.stream()),
                 "dnb_orthogonalize_batch")


class _Exchange:
    """A symmetric payload buffer + the reduce-only kernel: ``mean()`` leaves the cross-site mean
    of every site's payload in ``self.out[site]``."""

    def __init__(self, group: SiteGroup, numel: int, name: str):
        self.red = FusedDSGD(group, numel, optimizer="none", mode="one_shot", name=name, device_state=True)
        self.payload = self.red.grads                         # per-site fp32 views in the arena
        self.out = [torch.zeros(self.red.n, device=group.device) for _ in self.red.sites_here()]
        self._ready = False

    def mean(self):
        lib = native.require_gpu()
        r = self.red
        if not self._ready:
            r._descs = [r._desc(s, i) for i, s in enumerate(r.sites_here())]
            for d, o in zip(r._descs, self.out):
                d.out = o.data_ptr()
            self._ready = True
        st = native.stream()
        for d in r._descs:
            d.step, d.epoch, d.lr = 1, 0, 0.0
            native.check(lib.dnb_dsgd_fused_step(ctypes.byref(d), st), "dnb_dsgd_fused_step(reduce)")
        return self.out


class _LocalOptim:
    """The fused Adam/SGD kernel on one site's own flat buffers (no cross-site traffic)."""

    def __init__(self, device, params, lr, optimizer="adam", momentum=0.0):
        self.params = params
        offs, total = flat_layout(params)
        self.group = SiteGroup.local(1, total * 12 + (4 << 20), device)
        self.f = FusedDSGD(self.group, total, lr=lr, optimizer=optimizer, momentum=momentum, mode="one_shot",
                           name="local", device_state=True, shadow_bf16=True)
        self.holder = torch.nn.ParameterList(params)
        self.f.bind(self.holder, site=0, params=params)
        self.offs, self.total = offs, total

    @property
    def flat_grad(self):
        return self.f.grads[0]

    def zero_grad(self):
        self.f.zero_grad(site=0)

    def step(self):
        self.f.step()


class DevicePowerSGD:
    """powerSGD across the sites of ``group``; ``site_params[s]`` = ordered parameter list of site s.

    Per step and site, all on kernels of this library, all buffers allocated here once:
    pass P per matrix (``E += grad``, ``P = M·Q`` in one sweep) + one segment-copy launch for the dense (1-D)
    gradients → all-reduce(P) → batched orthogonalisation → memset + pass Q per matrix → all-reduce(Q) → finish per
    matrix (``grad = P̂·Q̄ᵀ`` formed on the fly, ``E = M − M̂``) + one segment-copy launch → fused Adam/SGD.
    ``csrc/psgd.cu`` explains why these are HBM-streaming kernels and not tensor-core GEMMs (rank <= 8: ~1 flop/byte).
    """

    def __init__(self, group: SiteGroup, site_params, rank=4, lr=1e-3, optimizer="adam", momentum=0.0, seed=0):
        self.group, self.rank = group, int(rank)
        self.sites = list(range(group.nsites)) if group.backend == "local" else [group.rank]
        assert len(site_params) == len(self.sites)
        dev = group.device
        self.opt = [_LocalOptim(dev, ps, lr, optimizer, momentum) for ps in site_params]
        p0 = site_params[0]
        self.mat_ix = [i for i, p in enumerate(p0) if p.dim() >= 2]
        self.vec_ix = [i for i, p in enumerate(p0) if p.dim() < 2]
        self.shape2d = [(p0[i].shape[0], p0[i].numel() // p0[i].shape[0]) for i in self.mat_ix]
        self.r = [max(1, min(self.rank, n, m)) for n, m in self.shape2d]
        if max(self.r) > 8:
            raise ValueError("device powerSGD supports rank <= 8")
        # payload layouts (offsets aligned to 8 floats)
        self.p_off, cur = [], 0
        for (n, m), r in zip(self.shape2d, self.r):
            self.p_off.append(cur)
            cur += _r8(n * r)
        self.v_off = []
        for i in self.vec_ix:
            self.v_off.append(cur)
            cur += _r8(p0[i].numel())
        self.q_off, qcur = [], 0
        for (n, m), r in zip(self.shape2d, self.r):
            self.q_off.append(qcur)
            qcur += _r8(m * r)
        self.xp = _Exchange(group, cur, "psgd_p")
        self.xq = _Exchange(group, qcur, "psgd_q")
        # per-site state: Q (fp32, flat, same layout as the Q exchange) identical on all sites, error feedback E
        self.Qflat = [torch.zeros(self.xq.red.n, device=dev) for _ in self.sites]
        for qf in self.Qflat:
            for k, (sh, r) in enumerate(zip(self.shape2d, self.r)):
                qf[self.q_off[k]:self.q_off[k] + sh[1] * r].copy_(lowrank.psgd_init_q(sh, r, seed + 17 * k, dev).reshape(-1))
        self.E = [[torch.zeros(n, m, device=dev) for (n, m) in self.shape2d] for _ in self.sites]

    @property
    def Q(self):
        """Per site, per matrix: the warm-start ``Q [m, r]`` (views into the flat state)."""
        return [[qf[self.q_off[k]:self.q_off[k] + sh[1] * r].view(sh[1], r) for k, (sh, r) in enumerate(zip(self.shape2d, self.r))]
                for qf in self.Qflat]

    def zero_grad(self, site=None):
        for li, s in enumerate(self.sites):
            if site is None or s == site:
                self.opt[li].zero_grad()

    def _pass(self, which, li, k, grad, P, Q):
        lib = native.require_gpu()
        n, m = self.shape2d[k]
        d = native.PsgdDesc()
        d.grad, d.E, d.Q, d.P = grad.data_ptr(), self.E[li][k].data_ptr(), Q.data_ptr(), P.data_ptr()
        d.n, d.m, d.r, d.pass_ = n, m, self.r[k], which
        native.check(lib.dnb_psgd_pass(ctypes.byref(d), native.stream()), f"dnb_psgd_pass({which})")

    def _grad2d(self, p):
        g = p.grad
        if not g.is_contiguous():
            raise ValueError("powerSGD device kernels need contiguous gradients")
        return g

    def step(self):
        # ---- pass P (+ dense gradients into the same exchange buffer)
        for li, s in enumerate(self.sites):
            ps = self.opt[li].params
            pay = self.xp.payload[s]
            for k, i in enumerate(self.mat_ix):
                n, m = self.shape2d[k]
                r = self.r[k]
                self._pass(0, li, k, self._grad2d(ps[i]), pay[self.p_off[k]:self.p_off[k] + n * r],
                           self.Qflat[li][self.q_off[k]:self.q_off[k] + m * r])
            native.seg_copy([(ps[i].grad.reshape(-1), pay[self.v_off[j]:self.v_off[j] + ps[i].numel()])
                             for j, i in enumerate(self.vec_ix)])
        Pm = self.xp.mean()
        # ---- orthogonalise, pass Q
        for li, s in enumerate(self.sites):
            mats = [Pm[li][self.p_off[k]:self.p_off[k] + n * r].view(n, r)
                    for k, ((n, m), r) in enumerate(zip(self.shape2d, self.r))]
            orthogonalize_batch(mats)
            pay = self.xq.payload[s]
            native.memset_async(pay)
            for k, ((n, m), r) in enumerate(zip(self.shape2d, self.r)):
                self._pass(1, li, k, pay, mats[k], pay[self.q_off[k]:self.q_off[k] + m * r])
        Qm = self.xq.mean()
        # ---- finish: grad = P̂·Q̄ᵀ, E = M − M̂; dense gradients back; warm start; optimiser
        lib = native.require_gpu()
        lib.dnb_copy_async.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
        for li, s in enumerate(self.sites):
            ps = self.opt[li].params
            for k, i in enumerate(self.mat_ix):
                n, m = self.shape2d[k]
                r = self.r[k]
                self._pass(2, li, k, self._grad2d(ps[i]), Pm[li][self.p_off[k]:self.p_off[k] + n * r],
                           Qm[li][self.q_off[k]:self.q_off[k] + m * r])
            native.seg_copy([(Pm[li][self.v_off[j]:self.v_off[j] + ps[i].numel()], ps[i].grad.reshape(-1))
                             for j, i in enumerate(self.vec_ix)])
            # in place: the warm start survives CUDA-graph replay (plain device-to-device copy, no kernel)
            native.check(lib.dnb_copy_async(self.Qflat[li].data_ptr(), Qm[li].data_ptr(), self.Qflat[li].numel() * 4, 3,
                                            native.stream()), "copy Q")
            self.opt[li].step()


from ..distrib.rankdad import CapturedTap, _LinearTap  # noqa: E402,F401  (re-exported: engine.py, tests)


def _hilo_rows(x16: torch.Tensor, ld: int):
    """fp32 ``[16, n]`` → bf16 ``[32, ld]``: rows 0-15 the bf16 roundings, rows 16-31 the bf16 remainders."""
    n = x16.shape[1]
    out = torch.zeros(32, ld, device=x16.device, dtype=torch.bfloat16)
    hi = x16.to(torch.bfloat16)
    out[:16, :n] = hi
    out[16:, :n] = (x16 - hi.float()).to(torch.bfloat16)
    return out


class _DadWorkspace:
    """Persistent buffers of one layer for one of the two rankDAD kernels (``csrc/dad.cu``)."""

    def __init__(self, dev, N, n_in, n_out, rank, seed, stacked=False):
        f32 = dict(device=dev, dtype=torch.float32)
        b16 = dict(device=dev, dtype=torch.bfloat16)
        self.N, self.n_in, self.n_out = int(N), int(n_in), int(n_out)
        self.ZN = torch.zeros(32, _r8(N), **b16)
        self.ZI = torch.zeros(32, _r8(n_in), **b16)
        self.ZO = torch.zeros(32, _r8(n_out), **b16)
        v0 = torch.zeros(native.DAD_R, n_in, **f32)
        v0[:rank] = lowrank._seeded_normal((n_in, rank), seed, dev, torch.float32).t()
        self.V0 = _hilo_rows(v0, _r8(n_in))
        self.W = torch.zeros(n_out, native.DAD_R, **f32)
        self.U = torch.zeros(n_out, native.DAD_R, **f32)
        self.V = torch.zeros(n_in, native.DAD_R, **f32)
        if stacked:       # mode 1: the kernel builds its own operands from the received factors
            self.A = torch.zeros(N, _r8(n_in), **b16)
            self.AT = torch.zeros(n_in, _r8(N), **b16)
            self.D = torch.zeros(N, _r8(n_out), **b16)
            self.DT = torch.zeros(n_out, _r8(N), **b16)
            self.Uc = torch.zeros(n_out, 64, **b16)
            self.Vc = torch.zeros(n_in, 64, **b16)

    def fill(self, L: "native.DadLayerDesc"):
        L.N, L.n_in, L.n_out = self.N, self.n_in, self.n_out
        L.ZN, L.ZI, L.ZO, L.V0 = self.ZN.data_ptr(), self.ZI.data_ptr(), self.ZO.data_ptr(), self.V0.data_ptr()
        L.ldzn, L.ldzi, L.ldzo = self.ZN.stride(0), self.ZI.stride(0), self.ZO.stride(0)
        L.W, L.U, L.V = self.W.data_ptr(), self.U.data_ptr(), self.V.data_ptr()


def _tma_ok(t):
    return (t is not None and t.dtype == torch.bfloat16 and t.dim() == 2 and t.stride(1) == 1
            and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0)


def dad_kernels_supported(dims, rank):
    """The rankDAD kernels keep one row block of the left factor in registers while orthogonalising:
    ``out <= DAD_MAX_OUT`` per layer; at most ``DAD_MAX_LAYERS`` layers per launch group is handled by
    splitting into several launches, so only the width and the rank bound the envelope."""
    return rank <= native.DAD_R and all(o <= native.DAD_MAX_OUT for o, _ in dims)


class DeviceRankDAD:
    """rankDAD across sites on the device kernels of ``csrc/dad.cu``.

    ``taps[s]``: the taps (``_LinearTap`` / ``CapturedTap``) of site s, same order on every site; ``dense[s]``:
    the parameters not reconstructed from factors (norm layers and, in device mode, every bias: the bias
    gradient ``Σ_rows Δ`` is exact and tiny, so it travels dense instead of as an extra activation column).
    Per step and site: [one pack launch for pairs that arrive in fp32] → compress kernel (all layers; its tail
    is the factor all-gather) → recompress + reconstruct kernel (writes ``W.grad`` into the flat gradient
    buffer) → fused Adam/SGD.  ``dad_tol`` is not evaluated: all ``num_pow_iters`` iterations run (no host sync).
    """

    def __init__(self, group: SiteGroup, site_params, taps, dense, rank=10, num_pow_iters=5, tol=1e-3, lr=1e-3,
                 optimizer="adam", momentum=0.0, seed=0, nsplit=3, cluster=None):
        self.group = group
        self.sites = list(range(group.nsites)) if group.backend == "local" else [group.rank]
        self.S = group.nsites
        self.rank, self.iters, self.tol, self.seed = int(rank), int(num_pow_iters), float(tol), int(seed)
        self.nsplit = int(nsplit)
        self.cluster = int(cluster or os.environ.get("DINUNET_DAD_CLUSTER", "8"))
        dev = group.device
        self.opt = [_LocalOptim(dev, ps, lr, optimizer, momentum) for ps in site_params]
        self.taps, self.dense = taps, dense
        t0 = taps[0]
        self.dims = [(t.module.weight.shape[0], t.module.weight.shape[1]) for t in t0]
        if not dad_kernels_supported(self.dims, self.rank):
            raise ValueError(f"rankDAD device kernels need out <= {native.DAD_MAX_OUT} per Linear and rank <= "
                             f"{native.DAD_R}; got dims {self.dims}, rank {self.rank}")
        R = native.DAD_R
        # slot layout (floats): per layer U [out,16] then V [in,16]; then the dense gradients
        self.f_off, cur = [], 0
        for (o, i) in self.dims:
            self.f_off.append(cur)
            cur += (o + i) * R
        self.dense_off = cur
        self.spans = []
        for li in range(len(self.sites)):
            pid = {id(p): k for k, p in enumerate(site_params[li])}
            sp = [(self.opt[li].offs[pid[id(p)]], p.numel()) for p in dense[li]]
            self.spans.append(sp)
        if len(self.spans[0]) > native.DAD_MAX_SPANS:
            raise ValueError(f"rankDAD device kernels take at most {native.DAD_MAX_SPANS} dense parameters")
        cur += sum(n for _, n in self.spans[0])
        self.slot = (cur + 3) // 4 * 4
        self.fact = group.alloc("dad.fact", 2 * self.S * self.slot, torch.float32)
        self.flags = group.alloc("dad.flags", 64, torch.int32)
        self.state = [torch.zeros(4, device=dev, dtype=torch.int32) for _ in self.sites]
        self.ws0 = [dict() for _ in self.sites]          # (layer, N) -> _DadWorkspace, compress
        n1 = self.nsplit * self.S * R
        self.ws1 = [[_DadWorkspace(dev, n1, i, o, self.rank, self.seed + 101 * k, stacked=True)
                     for k, (o, i) in enumerate(self.dims)] for _ in self.sites]
        self._pack_bufs = [dict() for _ in self.sites]
        self._desc1 = [None for _ in self.sites]
        self.groups = [list(range(a, min(a + native.DAD_MAX_LAYERS, len(self.dims))))
                       for a in range(0, len(self.dims), native.DAD_MAX_LAYERS)]

    def zero_grad(self, site=None):
        for li, s in enumerate(self.sites):
            if site is None or s == site:
                self.opt[li].zero_grad()

    # ------------------------------------------------------------------ operands
    def _operands(self, li, k, entries, packs):
        """bf16 (D, DT, A, AT) of layer k; blocks that arrive without both orientations (or in fp32) are queued
        for the pack kernel into persistent buffers."""
        o, i = self.dims[k]
        if len(entries) > 1:        # local_iterations > 1: stack the micro-batches along N
            ent = {"D": torch.cat([e["D"].float() for e in entries], 0), "A": torch.cat([e["A"].float() for e in entries], 0)}
        else:
            ent = entries[0]
        out = {}
        N = ent["D"].shape[0]
        for key, tkey, width in (("D", "DT", o), ("A", "AT", i)):
            x, xt = ent[key], ent.get(tkey)
            if _tma_ok(x) and _tma_ok(xt):
                out[key], out[tkey] = x, xt
                continue
            bufs = self._pack_bufs[li]
            bk = (k, key, N)
            if bk not in bufs:
                bufs[bk] = (torch.zeros(N, _r8(width), device=x.device, dtype=torch.bfloat16),
                            torch.zeros(width, _r8(N), device=x.device, dtype=torch.bfloat16))
            rows, colsT = bufs[bk]
            src = x if (x.dtype == torch.float32 and x.stride(1) == 1) else x.float().contiguous()
            packs.append((src, rows, colsT, N, width))
            out[key], out[tkey] = rows[:, :width], colsT[:, :N]
        return out, N

    def _run_packs(self, packs):
        lib = native.require_gpu()
        for a in range(0, len(packs), native.DAD_MAX_PACK):
            d = native.DadPackDesc()
            chunk = packs[a:a + native.DAD_MAX_PACK]
            for j, (src, rows, colsT, N, width) in enumerate(chunk):
                d.src[j], d.rows[j], d.colsT[j] = src.data_ptr(), rows.data_ptr(), colsT.data_ptr()
                d.lds[j], d.ldr[j], d.ldt[j] = src.stride(0), rows.stride(0), colsT.stride(0)
                d.N[j], d.d[j] = N, width
            d.count = len(chunk)
            self._keep = getattr(self, "_keep", [])
            self._keep.append([p[0] for p in chunk])      # sources stay alive until the launch has been queued
            native.check(lib.dnb_dad_pack(ctypes.byref(d), native.stream()), "dnb_dad_pack")
        self._keep = []

    def _common(self, d, li, s, mode):
        g = self.group
        fp, gp = g.ptrs("dad.fact"), g.ptrs("dad.flags")
        for r in range(g.nsites):
            d.fact[r], d.flags[r] = fp[r], gp[r]
        d.slot_floats, d.dense_off = self.slot, self.dense_off
        for j, (off, n) in enumerate(self.spans[li]):
            d.span_off[j], d.span_len[j] = off, n
        d.n_spans = len(self.spans[li])
        d.flat_grad = self.opt[li].flat_grad.data_ptr()
        d.dev_state = self.state[li].data_ptr()
        d.grad_scale = 1.0 / self.S
        d.mode, d.iters, d.cluster, d.nsplit = mode, self.iters, self.cluster, self.nsplit
        d.nranks, d.rank, d.use_flags, d.rank_cols = g.nsites, s, int(g.use_flags), self.rank

    def _launch_groups(self, build, li, s, mode):
        """One launch for the model's tapped layers (``step`` refuses models with more than DAD_MAX_LAYERS)."""
        lib = native.require_gpu()
        for gi, layers in enumerate(self.groups):
            d = native.DadDesc()
            self._common(d, li, s, mode)
            for j, k in enumerate(layers):
                build(d.layer[j], k)
            d.n_layers = len(layers)
            if gi != 0:
                d.n_spans = 0
            native.check(lib.dnb_dad_step(ctypes.byref(d), native.stream()), f"dnb_dad_step(mode {mode})")

    def step(self):
        if len(self.groups) > 1:
            raise NotImplementedError(f"more than {native.DAD_MAX_LAYERS} tapped Linear layers per model")
        R = native.DAD_R
        held = []
        for li, s in enumerate(self.sites):
            packs, ops = [], []
            for k, tap in enumerate(self.taps[li]):
                entries = tap.pop_entries()
                if not entries:
                    raise RuntimeError(f"rankDAD: Linear #{k} {self.dims[k]} produced no (delta, activation) pair this "
                                       "step (is the model's fused path bypassing its hooks?)")
                ops.append(self._operands(li, k, entries, packs))
            if packs:
                self._run_packs(packs)
            held.append(ops)

            def build0(L, k, li=li, ops=ops):
                o, i = self.dims[k]
                t, N = ops[k]
                ws = self.ws0[li].get((k, N))
                if ws is None:
                    ws = self.ws0[li][(k, N)] = _DadWorkspace(self.group.device, N, i, o, self.rank, self.seed + k)
                ws.fill(L)
                L.A, L.AT, L.D, L.DT = t["A"].data_ptr(), t["AT"].data_ptr(), t["D"].data_ptr(), t["DT"].data_ptr()
                L.lda, L.ldat, L.ldd, L.lddt = t["A"].stride(0), t["AT"].stride(0), t["D"].stride(0), t["DT"].stride(0)
                L.fact_off = self.f_off[k]
            self._launch_groups(build0, li, s, 0)
        for li, s in enumerate(self.sites):
            def build1(L, k, li=li):
                ws = self.ws1[li][k]
                ws.fill(L)
                L.A, L.AT, L.D, L.DT = ws.A.data_ptr(), ws.AT.data_ptr(), ws.D.data_ptr(), ws.DT.data_ptr()
                L.lda, L.ldat, L.ldd, L.lddt = ws.A.stride(0), ws.AT.stride(0), ws.D.stride(0), ws.DT.stride(0)
                L.fact_off = self.f_off[k]
                L.Uc, L.Vc = ws.Uc.data_ptr(), ws.Vc.data_ptr()
                w = self.taps[li][k].module.weight
                L.G, L.ldg = w.grad.data_ptr(), w.shape[1]
            self._launch_groups(build1, li, s, 1)
            self.opt[li].step()
        self._held = held          # operand tensors of this step stay referenced until the next one
