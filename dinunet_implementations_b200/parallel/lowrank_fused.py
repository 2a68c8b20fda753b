"""Device-mode rankDAD and powerSGD: the low-rank engines with
every payload kept in device memory and exchanged through the symmetric arenas.

Math = ``distrib/lowrank.py`` (the oracle).  What runs where:

* big products (``P = M·Q``, ``Q = Mᵀ·P̂``, ``M̂ = P̂·Qᵀ``, rankDAD reconstruction ``ΔᵀA``) — the
  tcgen05/TMA GEMM (``ops.gemm``), bf16 operands, fp32 accumulation in TMEM;
* orthogonalisation — ``dnb_orthogonalize_batch`` (one CTA per factor, all layers in one launch);
* cross-site exchange — the reduce-only mode of the fused dSGD kernel pulling every peer's payload
  over NVLink (deterministic summation order ⇒ bit-identical factors on all sites).  rankDAD's
  all-gather of factors is expressed as the same sum: each site fills only its own slot of an
  ``[S, payload]`` buffer, so the sum *is* the concatenation;
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
from ..ops.elementwise import cast_bf16, transpose_bf16
from ..ops.gemm import gemm_bf16
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


def _bf16_pad_cols(x: torch.Tensor, cols: int | None = None):
    """fp32 ``[a, b]`` → bf16 with the leading dimension padded to a multiple of 8 (TMA row pitch)."""
    a, b = x.shape
    out = torch.zeros(a, _r8(cols or b), device=x.device, dtype=torch.bfloat16)
    out[:, :b].copy_(x)
    return out[:, :b]


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

    def step(self):
        self.f.step()


class DevicePowerSGD:
    """powerSGD across the sites of ``group``; ``site_params[s]`` = ordered parameter list of site s."""

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
        # per-site state: Q (fp32 [m,r]) identical on all sites, error feedback E (fp32 [n,m])
        self.Q = [[lowrank.psgd_init_q(sh, r, seed + 17 * k, dev) for k, (sh, r) in enumerate(zip(self.shape2d, self.r))]
                  for _ in self.sites]
        self.E = [[torch.zeros(n, m, device=dev) for (n, m) in self.shape2d] for _ in self.sites]

    def zero_grad(self, site=None):
        for li, s in enumerate(self.sites):
            if site is None or s == site:
                self.opt[li].flat_grad.zero_()

    def step(self):
        S = len(self.sites)
        Mb, MbT = [[] for _ in range(S)], [[] for _ in range(S)]
        # ---- phase P: M = grad + E ; P = M·Q ; dense 1-D grads ride along
        for li, s in enumerate(self.sites):
            ps = self.opt[li].params
            pay = self.xp.payload[s]
            for k, i in enumerate(self.mat_ix):
                n, m = self.shape2d[k]
                r = self.r[k]
                E = self.E[li][k]
                E.add_(ps[i].grad.view(n, m))                                  # E now holds M
                mb = _bf16_pad_cols(E)
                Mb[li].append(mb)
                MbT[li].append(transpose_bf16(mb))                             # [m, n] for Q = Mᵀ·P̂
                qt = _bf16_pad_cols(self.Q[li][k].t().contiguous())            # [r, m]
                gemm_bf16(mb, qt, out=pay[self.p_off[k]:self.p_off[k] + n * r].view(n, r))
            for j, i in enumerate(self.vec_ix):
                pay[self.v_off[j]:self.v_off[j] + ps[i].numel()].copy_(ps[i].grad.reshape(-1))
        Pm = self.xp.mean()
        # ---- orthogonalise, Q = Mᵀ·P̂
        for li, s in enumerate(self.sites):
            mats = [Pm[li][self.p_off[k]:self.p_off[k] + n * r].view(n, r)
                    for k, ((n, m), r) in enumerate(zip(self.shape2d, self.r))]
            orthogonalize_batch(mats)
            pay = self.xq.payload[s]
            for k, ((n, m), r) in enumerate(zip(self.shape2d, self.r)):
                pt = _bf16_pad_cols(mats[k].t().contiguous())                  # [r, n]
                gemm_bf16(MbT[li][k], pt, out=pay[self.q_off[k]:self.q_off[k] + m * r].view(m, r))
        Qm = self.xq.mean()
        # ---- finish: M̂ = P̂·Q̄ᵀ -> grad ; E = M − M̂ ; dense grads ; optimiser
        for li, s in enumerate(self.sites):
            ps = self.opt[li].params
            for k, i in enumerate(self.mat_ix):
                n, m = self.shape2d[k]
                r = self.r[k]
                P_hat = Pm[li][self.p_off[k]:self.p_off[k] + n * r].view(n, r)
                Qk = Qm[li][self.q_off[k]:self.q_off[k] + m * r].view(m, r)
                g = ps[i].grad.view(n, m)
                gemm_bf16(_bf16_pad_cols(P_hat), _bf16_pad_cols(Qk), out=g)    # K = r (zero-filled to a k-tile)
                self.E[li][k].sub_(g)
                self.Q[li][k].copy_(Qk)          # in place: the warm start survives CUDA-graph replay
            for j, i in enumerate(self.vec_ix):
                ps[i].grad.copy_(Pm[li][self.v_off[j]:self.v_off[j] + ps[i].numel()].view_as(ps[i]))
            self.opt[li].step()


class CapturedTap:
    """Stands in for ``distrib.rankdad._LinearTap`` when the fused kernels hand over the
    (delta, activation) pair of a Linear directly (``model._dad_pairs[key]``), so no module hooks
    and no eager forward are needed."""

    def __init__(self, module: torch.nn.Linear, model: torch.nn.Module, key: str):
        self.module, self.model, self.key = module, model, key

    def pop(self):
        pairs = getattr(self.model, "_dad_pairs", None) or {}
        if self.key not in pairs:
            return None, None
        D, A = pairs.pop(self.key)
        D, A = D.float(), A.float()
        if self.module.bias is not None:
            A = torch.cat([A, torch.ones(A.shape[0], 1, device=A.device)], 1)
        return D, A


class DeviceRankDAD:
    """rankDAD across sites.  ``taps[s]`` = list of ``distrib.rankdad._LinearTap`` of site s (same
    order on every site); ``dense[s]`` = parameters not covered by a tapped Linear."""

    def __init__(self, group: SiteGroup, site_params, taps, dense, rank=10, num_pow_iters=5, tol=1e-3, lr=1e-3,
                 optimizer="adam", momentum=0.0, seed=0):
        self.group = group
        self.sites = list(range(group.nsites)) if group.backend == "local" else [group.rank]
        self.S = group.nsites
        self.rank, self.iters, self.tol, self.seed = int(rank), int(num_pow_iters), float(tol), int(seed)
        dev = group.device
        self.opt = [_LocalOptim(dev, ps, lr, optimizer, momentum) for ps in site_params]
        self.taps, self.dense = taps, dense
        t0 = taps[0]
        self.dims = [(t.module.weight.shape[0], t.module.weight.shape[1] + (1 if t.module.bias is not None else 0))
                     for t in t0]
        r = self.rank
        # one slot per site: [d_0 (r x out), a_0 (r x in), d_1, ...]; dense grads are averaged separately
        self.f_off, cur = [], 0
        for (o, i) in self.dims:
            self.f_off.append((cur, cur + _r8(r * o)))
            cur += _r8(r * o) + _r8(r * i)
        self.slot = cur
        self.xf = _Exchange(group, self.slot * self.S, "dad_factors")
        n_dense = sum(p.numel() for p in dense[0])
        self.xd = _Exchange(group, max(n_dense, 8), "dad_dense") if n_dense else None

    def zero_grad(self, site=None):
        for li, s in enumerate(self.sites):
            if site is None or s == site:
                self.opt[li].flat_grad.zero_()

    # The layers are independent of one another and each one's subspace iteration is a strictly serial chain
    # of tiny kernels (a few SMs each), so the per-layer work is spread over side streams: concurrent in
    # eager mode, parallel branches when the step is captured in a CUDA graph.
    _N_STREAMS = int(os.environ.get("DINUNET_DAD_STREAMS", "8"))

    def _streams(self):
        if getattr(self, "_side", None) is None:
            self._side = [torch.cuda.Stream(device=self.group.device) for _ in range(self._N_STREAMS)]
        return self._side

    def _fork(self):
        main = torch.cuda.current_stream(self.group.device)
        side = self._streams()
        for st in side:
            st.wait_stream(main)
        return main, side

    @staticmethod
    def _join(main, side):
        for st in side:
            main.wait_stream(st)

    def step(self):
        r = self.rank
        for li, s in enumerate(self.sites):
            pay = self.xf.payload[s]
            pay.zero_()                                           # other sites' slots must stay zero
            base = s * self.slot
            pairs = [tap.pop() for tap in self.taps[li]]          # on the stream that produced them
            main, side = self._fork()
            for k, (D, A) in enumerate(pairs):
                o, i = self.dims[k]
                with torch.cuda.stream(side[k % len(side)]):
                    if D is None:
                        D, A = torch.zeros(1, o, device=pay.device), torch.zeros(1, i, device=pay.device)
                    d, a, _ = lowrank.dad_compress(D, A, r, self.iters, self.tol, self.seed + k)
                    od, oa = self.f_off[k]
                    pay[base + od: base + od + r * o].copy_(d.reshape(-1))
                    pay[base + oa: base + oa + r * i].copy_(a.reshape(-1))
            self._join(main, side)
            if self.xd is not None:
                flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1)
                                  for p in self.dense[li]])
                self.xd.payload[s][:flat.numel()].copy_(flat)
        Fm = self.xf.mean()                                       # mean of disjoint slots = concat / S
        Dm = self.xd.mean() if self.xd is not None else None
        for li, s in enumerate(self.sites):
            cat = Fm[li] * float(self.S)                          # undo the mean: plain concatenation
            main, side = self._fork()
            for k, tap in enumerate(self.taps[li]):
                o, i = self.dims[k]
                od, oa = self.f_off[k]
                with torch.cuda.stream(side[k % len(side)]):
                    ds = [cat[q * self.slot + od: q * self.slot + od + r * o].view(r, o) for q in range(self.S)]
                    as_ = [cat[q * self.slot + oa: q * self.slot + oa + r * i].view(r, i) for q in range(self.S)]
                    d, a, _ = lowrank.dad_reduce(ds, as_, r, self.iters, self.tol, self.seed + 101 * k, average=True)
                    # reconstruction GEMM  G[out, in] = d^T a  (K = r)
                    w = tap.module.weight
                    G = torch.empty(o, i, device=w.device)
                    gemm_bf16(_bf16_pad_cols(d.t().contiguous()), _bf16_pad_cols(a.t().contiguous()), out=G)
                    if tap.module.bias is not None:
                        tap.module.bias.grad.copy_(G[:, -1])
                        w.grad.copy_(G[:, :-1])
                    else:
                        w.grad.copy_(G)
            self._join(main, side)
            if Dm is not None:
                off = 0
                for p in self.dense[li]:
                    p.grad.copy_(Dm[li][off:off + p.numel()].view_as(p))
                    off += p.numel()
            self.opt[li].step()
