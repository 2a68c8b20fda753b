"""Device-mode dSGD: flat symmetric buffers + the fused reduce/optimiser kernel (K1).

``FusedDSGD`` owns, inside a ``SiteGroup`` arena, one flat fp32 parameter buffer, one flat
gradient buffer in the wire dtype (``precision_bits``: fp32 / bf16 / fp16) and optionally a
bf16 shadow of the parameters for the tensor-core kernels.  ``bind`` re-points a module's
parameters (and, for an fp32 wire, their ``.grad``) at slices of those buffers, so backward
writes gradients straight into peer-visible memory and the updated weights appear in place —
there is no gather, no file, no host copy between backward and the next forward.

One ``step()`` = one kernel launch per site (``csrc/dsgd.cu``): cross-site barrier → mean of the
sites' gradients → unscale → Adam/SGD → push updated parameters → barrier.
"""
from __future__ import annotations

import ctypes
import os

import torch

from ..ops import native
from .comm import SiteGroup

_WIRE = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}


def wire_dtype_from_bits(precision_bits, prefer_bf16=True):
    if str(precision_bits) == "16":
        return torch.bfloat16 if prefer_bf16 else torch.float16
    return torch.float32


ALIGN = 64   # every parameter starts on a 64-element boundary: 16-byte aligned in bf16 for TMA / vector loads


def flat_layout(params):
    """``(offsets, total)`` for a list of parameters laid out with ``ALIGN``-element alignment."""
    offs, cur = [], 0
    for p in params:
        offs.append(cur)
        cur += (p.numel() + ALIGN - 1) // ALIGN * ALIGN
    return offs, cur


def ordered_params(module: torch.nn.Module):
    """Trainable parameters in the module's preferred flat order (``flat_param_order()`` if it
    defines one — e.g. both directions' i2h weights adjacent so one GEMM covers them)."""
    if hasattr(module, "flat_param_order"):
        ps = [p for p in module.flat_param_order() if p.requires_grad]
        seen = {id(p) for p in ps}
        ps += [p for p in module.parameters() if p.requires_grad and id(p) not in seen]
        return ps
    return [p for p in module.parameters() if p.requires_grad]


class FusedDSGD:
    def __init__(self, group: SiteGroup, numel: int, wire_dtype=torch.float32, lr=1e-3, betas=(0.9, 0.999), eps=1e-8,
                 weight_decay=0.0, optimizer="adam", momentum=0.0, shadow_bf16=False, mode="auto", loss_scale=1.0,
                 name="dsgd", device_state=False):
        self.group = group
        self.numel = int(numel)
        self.n = (self.numel + 7) // 8 * 8
        self.wire_dtype = wire_dtype
        self.lr, self.betas, self.eps, self.weight_decay = float(lr), betas, float(eps), float(weight_decay)
        self.opt = {"adam": 0, "sgd": 1, "none": 2}[optimizer]
        self.momentum = float(momentum)
        self.loss_scale = float(loss_scale)
        # one-shot pull below ~1 MB of gradients (pure latency regime)
        if mode == "auto":
            mode = "one_shot" if self.n * 4 <= (1 << 20) else "two_shot"
            # NVLS (opt-in, DINUNET_NVLS=1): reduce inside the NVSwitch and publish with one multicast store.  Per GPU the
            # reduction then receives 1/N of the buffer and sends all of it, the broadcast the other way round, so the busier
            # direction carries about what the P2P two-shot carries ((N-1)/N): measured at 2 GPUs it is not faster than the
            # P2P kernel (profiles/r2_agg_microbench_2gpu.jsonl), so it is not the default.
            if (mode == "two_shot" and group.mc_base and group.nsites > 1 and optimizer != "none"
                    and os.environ.get("DINUNET_NVLS", "0") == "1"):
                mode = "nvls"
        if mode == "nvls" and not (group.mc_base and group.nsites > 1):
            raise ValueError("mode='nvls' needs a multi-GPU group whose arena has a multicast mapping")
        self.mode = {"two_shot": 0, "one_shot": 1, "nvls": 2}[mode]
        self.params = group.alloc(f"{name}.params", self.n, torch.float32)
        self.grads = group.alloc(f"{name}.grads", self.n, wire_dtype)
        self.shadow = group.alloc(f"{name}.shadow", self.n, torch.bfloat16) if shadow_bf16 else None
        # every instance owns its signal pad: two reducers that shared one would see each other's barrier
        # generations and run ahead of their peers (their epochs advance independently)
        self._flag_name = f"{name}.flags"
        group.alloc(self._flag_name, native.SIGNAL_WORDS, torch.int32)
        self._names = (f"{name}.params", f"{name}.grads", f"{name}.shadow" if shadow_bf16 else None)
        dev = group.device
        n_states = group.nsites if group.backend == "local" else 1
        self.m = [torch.zeros(self.n, device=dev) for _ in range(n_states)]
        self.v = [torch.zeros(self.n, device=dev) for _ in range(n_states)]
        self.out = None
        self.t = 0
        # device-resident {steps done, scratch, lr bits, 0} per locally driven site: lets the launch be
        # captured once in a CUDA graph and replayed (step count / barrier epoch advance on the device)
        self.dev_state = None
        if device_state:
            self.dev_state = [torch.zeros(4, device=dev, dtype=torch.int32) for _ in range(n_states)]
            self.set_lr(self.lr)
        self._descs = None
        self._grad_staging = None   # per-site fp32 grads when the wire is 16-bit
        self._stage_clean = False   # True once an in-kernel pack has cleared the staging gradients
        # Race-detection aid: with DINUNET_DEBUG_POISON=1 the peer-visible gradient buffer is
        # filled with NaN right after every aggregation launch.  The kernel's closing barrier guarantees
        # that all peers have finished reading it by then, and zero_grad() clears it before the next
        # backward — so training is unaffected unless a peer reads a buffer outside its window, in which
        # case the NaN propagates into every site's parameters and is impossible to miss.
        self.debug_poison = os.environ.get("DINUNET_DEBUG_POISON", "0") not in ("", "0")

    # ------------------------------------------------------------------ module binding
    def sites_here(self):
        return list(range(self.group.nsites)) if self.group.backend == "local" else [self.group.rank]

    def set_lr(self, lr):
        self.lr = float(lr)
        if self.dev_state is not None:
            bits = torch.tensor([self.lr], dtype=torch.float32).view(torch.int32)
            for st in self.dev_state:
                st[2:3].copy_(bits)

    def bind(self, module: torch.nn.Module, site=None, params=None):
        """Move ``module``'s parameters into this site's flat buffer (values preserved).

        Each parameter starts on an ``ALIGN`` boundary; the gaps stay zero in every buffer.  If
        the bf16 shadow exists, ``module._bf16_shadow`` maps ``id(param)`` to its bf16 view so the
        tensor-core kernels read weights without a per-step cast."""
        site = self.group.rank if site is None else site
        flat_p, flat_g = self.params[site], self.grads[site]
        ps = params if params is not None else ordered_params(module)
        offs, total = flat_layout(ps)
        if total != self.numel:
            raise ValueError(f"module needs {total} flat elements, FusedDSGD was sized for {self.numel}")
        staging = None
        if self.wire_dtype != torch.float32:
            staging = torch.zeros(self.n, device=self.group.device)
            self._grad_staging = self._grad_staging or {}
            self._grad_staging[site] = staging
        shadow_map = {}
        self._span = getattr(self, "_span", {})
        self._span.update({id(p): (off, p.numel()) for p, off in zip(ps, offs)})
        with torch.no_grad():
            for p, off in zip(ps, offs):
                n = p.numel()
                flat_p[off:off + n].copy_(p.detach().reshape(-1))
                p.data = flat_p[off:off + n].view_as(p)
                gsrc = flat_g if staging is None else staging
                p.grad = gsrc[off:off + n].view_as(p)
                if self.shadow is not None:
                    shadow_map[id(p)] = self.shadow[site][off:off + n].view_as(p)
        if self.shadow is not None:
            self.shadow[site].copy_(flat_p)
            module._bf16_shadow = shadow_map
            for m in module.modules():
                m._bf16_shadow = shadow_map
        return ps

    def sync_shadow(self, site=None):
        """Re-derive the bf16 shadow from the fp32 parameters (after a checkpoint load or any other
        in-place write to the parameters that did not go through the kernel)."""
        if self.shadow is not None:
            for s in ([site] if site is not None else self.sites_here()):
                self.shadow[s].copy_(self.params[s])

    @property
    def inkernel_pack(self):
        """16-bit wire: the aggregation kernel itself casts the fp32 staging gradients into the wire buffer and
        clears the staging buffer for the next step (no cast launch, no zero-fill launch).  Not for several
        simulated sites in one process: their kernels run one after the other, so every site's wire buffer has
        to be complete before the first kernel starts."""
        return bool(self._grad_staging) and (self.group.backend in ("ipc", "symm") or self.group.nsites == 1)

    def pack_grads(self, site=None):
        """16-bit wire only: cast the fp32 staging gradients into the peer-visible wire buffer."""
        if self._grad_staging and not self.inkernel_pack:
            for s in ([site] if site is not None else self.sites_here()):
                self.grads[s].copy_(self._grad_staging[s])

    def set_overwritten(self, params):
        """``params``: parameters whose gradients the model's kernels overwrite every step — ``zero_grad`` then clears
        only the rest (merged into as few memsets as possible)."""
        skip = sorted(self._span[id(p)] for p in params if id(p) in getattr(self, "_span", {}))
        spans, cur = [], 0
        for off, n in skip:
            if off > cur:
                spans.append((cur, off - cur))
            cur = max(cur, off + (n + ALIGN - 1) // ALIGN * ALIGN)
        if cur < self.n:
            spans.append((cur, self.n - cur))
        self._zero_spans = spans

    def zero_grad(self, site=None):
        if self.inkernel_pack and self._stage_clean:
            return                                  # the last aggregation left the staging gradients at zero
        for s in ([site] if site is not None else self.sites_here()):
            buf = self._grad_staging[s] if self._grad_staging else self.grads[s]
            spans = getattr(self, "_zero_spans", None)
            if spans is None:
                native.memset_async(buf)
            else:
                for off, n in spans:
                    native.memset_async(buf[off:off + n])

    # ------------------------------------------------------------------ the step
    def _desc(self, site, state_ix):
        d = native.DsgdDesc()
        g = self.group
        gp, pp = g.ptrs(self._names[1]), g.ptrs(self._names[0])
        sp = g.ptrs(self._names[2]) if self._names[2] else [0] * g.nsites
        fp = g.ptrs(self._flag_name)
        for r in range(g.nsites):
            d.grads[r], d.params[r], d.shadow_bf16[r], d.flags[r] = gp[r], pp[r], sp[r] or None, fp[r]
        d.m, d.v = self.m[state_ix].data_ptr(), self.v[state_ix].data_ptr()
        d.out = self.out.data_ptr() if self.out is not None else None
        d.n, d.nranks, d.rank = self.n, g.nsites, site
        d.wire, d.mode, d.opt = _WIRE[self.wire_dtype], self.mode, self.opt
        if self.mode == 2:
            d.grads_mc, d.params_mc = g.mc_ptr(self._names[1]), g.mc_ptr(self._names[0])
            d.shadow_mc = g.mc_ptr(self._names[2]) if self._names[2] else None
        d.use_flags = int(g.use_flags)
        d.beta1, d.beta2, d.eps = self.betas[0], self.betas[1], self.eps
        d.weight_decay, d.momentum = self.weight_decay, self.momentum
        d.grad_scale = 1.0 / (g.nsites * self.loss_scale)
        d.dev_state = self.dev_state[state_ix].data_ptr() if self.dev_state is not None else None
        if self.inkernel_pack and self.opt != 2:
            d.stage, d.zero_stage = self._grad_staging[site].data_ptr(), 1
        return d

    def step(self, lr=None):
        """Aggregate + update.  ipc: one launch on this rank.  local: one launch per simulated site."""
        lib = native.require_gpu()
        if self._descs is None:
            self._descs = [self._desc(s, i) for i, s in enumerate(self.sites_here())]
        self.pack_grads()
        self.t += 1
        self._stage_clean = self.inkernel_pack and self.opt != 2
        st = native.stream()
        if self.dev_state is not None:
            # replayable launch: the kernel reads/advances the step itself; epoch base stays 0
            if lr is not None and float(lr) != self.lr:
                self.set_lr(lr)
            for d in self._descs:
                d.step, d.epoch, d.lr = 1, 0, self.lr
                native.check(lib.dnb_dsgd_fused_step(ctypes.byref(d), st), "dnb_dsgd_fused_step")
            self._poison()
            return
        epoch = self.group.next_epoch()
        for d in self._descs:
            d.step, d.epoch, d.lr = self.t, epoch, float(self.lr if lr is None else lr)
            native.check(lib.dnb_dsgd_fused_step(ctypes.byref(d), st), "dnb_dsgd_fused_step")
        self._poison()

    def _poison(self):
        if self.debug_poison and self.opt != 2:
            for s in self.sites_here():
                self.grads[s].fill_(float("nan"))

    def allreduce_mean(self, out: torch.Tensor):
        """Reduce-only use of the same kernel: ``out = mean over sites of grads`` (local site)."""
        assert self.opt == 2, "construct with optimizer='none'"
        self.out = out
        self._descs = None
        self.step()
        return out
