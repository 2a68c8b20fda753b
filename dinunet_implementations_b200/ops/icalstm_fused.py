"""Fused sm_100a forward/backward of the whole ICA-LSTM (one ``autograd.Function``).

Forward (8 launches, 6 on the critical path): time-major cast → encoder GEMM(+bias+ReLU, +transposed
copy) → input-projection GEMM for all timesteps and both directions → persistent LSTM recurrence (the
W_hh transposes for the BPTT kernel run beside it on a side stream) → classifier head (2 launches;
draws its own dropout mask).  Backward (13 launches, 5 on the critical path): head (2) → LSTM BPTT (the
transposes that do not depend on it run beside it) → K-major copy of dP → tcgen05 weight- and
data-gradient GEMMs on three parallel streams (ReLU mask, bias column-sums and transposed store fused in
the epilogue).  Reference math: ``comps/icalstm/models.py:69-110``.

Shapes the kernels do not take natively are padded, not sent to the eager path: a per-direction hidden
size that is not a multiple of 32 (``_Padded``: exact zero-padding, e.g. the reference inputspec's
``hidden_size 348``) and a feature count ``comps*window`` that is not a multiple of 8 (``_pitched_copy``).

Gradients are *accumulated* into ``param.grad`` (allocated on demand) — the same contract as
autograd — and the Function returns no parameter gradients itself; the device engine points
``param.grad`` at slices of the peer-visible flat gradient buffer, so the backward kernels write
straight into the memory the fused dSGD kernel aggregates.

bf16 copies of the weights come from the engine's parameter shadow when bound
(``model._bf16_shadow``), otherwise they are cast per call.
"""
from __future__ import annotations

import ctypes

import torch

from . import native
from .native import vp
from .gemm import gemm_bf16
from .lstm import lstm_forward, lstm_backward, lstm_supported
from .elementwise import cast_time_major, transpose_bf16, cast_bf16
from .head import head_forward, head_backward


def supported(model, x):
    if not native.available():
        return False
    lstm = model.lstm
    H, D = lstm.hidden_size, lstm.num_direction
    cls = model.classifier
    # H that is not a multiple of 32 (the reference inputspec's hidden_size 348 -> 174 per direction)
    # runs zero-padded to the next multiple (``_Padded``), so only the padded size has to fit
    return (x.dim() == 4 and x.shape[0] <= 64 and lstm_supported(_ceil32(H), D) and model.input_size % 8 == 0
            and cls[1].in_features == H * D and cls[1].out_features % 16 == 0
            and cls[4].in_features % 2 == 0 and (not model.training or x.shape[0] > 1))


def _pitched_copy(model, w):
    st = model.__dict__.setdefault("_pitched", {})
    key = (w.device.index, tuple(w.shape))
    if key not in st:
        st[key] = torch.zeros(w.shape[0], (w.shape[1] + 7) // 8 * 8, device=w.device, dtype=torch.bfloat16)
    v = st[key][:, :w.shape[1]]
    v.copy_(w)
    return v


def _ceil32(h):
    return (int(h) + 31) // 32 * 32


class _Padded:
    """Zero-padded operands for a per-direction hidden size ``H`` that is not a multiple of 32.

    The recurrence kernels tile the hidden units in blocks of 32 (one CTA of the cluster each), so
    ``H = 174`` runs as ``Hp = 192``.  Padding is exact, not approximate: a padded unit has zero
    weights and biases, so its pre-activations are 0, ``g = tanh(0) = 0`` keeps its cell at 0 and
    ``h = o * tanh(0) = 0``; nothing reads it (zero columns in ``W_hh`` and in the first classifier
    layer), so its gradient is 0 too.  Buffers are persistent (allocated once per model and device,
    the pad region is written only at allocation), which keeps the step CUDA-graph capturable:
    per step there are 9 strided copies in (bf16 shadows -> padded operands), one memset of the
    padded gradient staging area and 9 strided adds out (staging -> ``param.grad``).
    """

    def __init__(self, model, dev):
        cells = list(model.lstm.lstms)
        self.D, self.H, self.E = len(cells), model.lstm.hidden_size, model.input_size
        self.Hp = _ceil32(self.H)
        self.N1 = model.classifier[1].out_features
        D, Hp, E, N1 = self.D, self.Hp, self.E, self.N1
        z16 = lambda *sh: torch.zeros(*sh, device=dev, dtype=torch.bfloat16)
        self.w_ih, self.whh, self.W1 = z16(D * 4 * Hp, E), z16(D, 4 * Hp, Hp), z16(N1, D * Hp)
        self.b_ih = torch.zeros(D * 4 * Hp, device=dev)
        self.b_hh = torch.zeros(D, 4 * Hp, device=dev)
        sizes = dict(g_w_ih=D * 4 * Hp * E, g_whh=D * 4 * Hp * Hp, g_W1=N1 * D * Hp, g_b_ih=D * 4 * Hp, g_b_hh=D * 4 * Hp)
        self.stage = torch.zeros(sum(sizes.values()), device=dev)
        off = 0
        for k, n in sizes.items():
            setattr(self, k, self.stage[off:off + n])
            off += n
        self.g_w_ih = self.g_w_ih.view(D * 4 * Hp, E)
        self.g_whh = self.g_whh.view(D, 4 * Hp, Hp)
        self.g_W1 = self.g_W1.view(N1, D * Hp)
        self.g_b_ih = self.g_b_ih.view(D, 4 * Hp)
        self.g_b_hh = self.g_b_hh.view(D, 4 * Hp)

    @staticmethod
    def of(model, dev):
        st = model.__dict__.setdefault("_pad_state", {})
        key = (dev.type, dev.index)
        if key not in st:
            st[key] = _Padded(model, dev)
        return st[key]

    def fill(self, model):
        D, H, Hp, E, N1 = self.D, self.H, self.Hp, self.E, self.N1
        for d, c in enumerate(model.lstm.lstms):
            self.w_ih.view(D, 4, Hp, E)[d, :, :H].copy_(_w16(model, c.i2h.weight).view(4, H, E))
            self.whh.view(D, 4, Hp, Hp)[d, :, :H, :H].copy_(_w16(model, c.h2h.weight).view(4, H, H))
            self.b_ih.view(D, 4, Hp)[d, :, :H].copy_(c.i2h.bias.detach().view(4, H))
            self.b_hh.view(D, 4, Hp)[d, :, :H].copy_(c.h2h.bias.detach().view(4, H))
        self.W1.view(N1, D, Hp)[:, :, :H].copy_(_w16(model, model.classifier[1].weight).view(N1, D, H))

    def unpad(self, t):
        """``[rows, D*Hp] -> [rows, D*H]`` (copy)."""
        return t.reshape(t.shape[0], self.D, self.Hp)[:, :, :self.H].reshape(t.shape[0], self.D * self.H)

    def gates(self, t, d):
        """Direction ``d`` of a ``[rows, D*4*Hp]`` gate tensor without the pad units -> ``[rows, 4H]``."""
        return t.reshape(t.shape[0], self.D, 4, self.Hp)[:, d, :, :self.H].reshape(t.shape[0], 4 * self.H)

    def scatter_grads(self, model):
        """Add the staged (padded) gradients into the real ``param.grad`` tensors."""
        D, H, Hp, E, N1 = self.D, self.H, self.Hp, self.E, self.N1
        for d, c in enumerate(model.lstm.lstms):
            _grad_buf(c.i2h.weight).view(4, H, E).add_(self.g_w_ih.view(D, 4, Hp, E)[d, :, :H])
            _grad_buf(c.h2h.weight).view(4, H, H).add_(self.g_whh.view(D, 4, Hp, Hp)[d, :, :H, :H])
            _grad_buf(c.i2h.bias).view(4, H).add_(self.g_b_ih.view(D, 4, Hp)[d, :, :H])
            _grad_buf(c.h2h.bias).view(4, H).add_(self.g_b_hh.view(D, 4, Hp)[d, :, :H])
        _grad_buf(model.classifier[1].weight).view(N1, D, H).add_(self.g_W1.view(N1, D, Hp)[:, :, :H])


def _w16(model, p):
    sh = getattr(model, "_bf16_shadow", None)
    if sh is not None and id(p) in sh:
        return sh[id(p)]
    return cast_bf16(p.detach())


def _grad_buf(p):
    if p.grad is None:
        p.grad = torch.zeros_like(p)
    return p.grad


_SIDE = {}


def _side_streams(dev):
    key = (dev.type, dev.index)
    if key not in _SIDE:
        _SIDE[key] = (torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev))
    return _SIDE[key]


def _adjacent(a, b):
    """True when ``b`` starts exactly where ``a`` ends inside the SAME storage (flat buffers)."""
    return (a.untyped_storage().data_ptr() == b.untyped_storage().data_ptr()
            and a.is_contiguous() and b.is_contiguous()
            and a.data_ptr() + a.numel() * a.element_size() == b.data_ptr())


def _span(a, shape, stride):
    """View of ``shape`` starting at ``a`` and running on into the tensors that follow it."""
    return torch.as_strided(a, shape, stride, storage_offset=a.storage_offset())


class _ICALstmFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, model, *params):
        lib = native.require_gpu()
        B, S = x.shape[0], x.shape[1]
        enc_l = model.encoder[0]
        cells = list(model.lstm.lstms)
        D, H = len(cells), model.lstm.hidden_size
        E = model.input_size
        cls = model.classifier
        training = model.training

        # 1. cast + time-major (the transposed copy the encoder weight gradient needs is made in the backward
        # pass, on a side stream under the BPTT kernel, instead of here on the critical path)
        xb, _ = cast_time_major(x.float(), want_transpose=False)
        # 2. encoder: relu(x W^T + b), plus the transposed copy the i2h weight-gradient GEMM needs
        BSp = (S * B + 7) // 8 * 8
        encT = torch.empty(E, BSp, device=x.device, dtype=torch.bfloat16)[:, :S * B]
        w_enc = _w16(model, enc_l.weight)
        if w_enc.shape[1] % 8:
            # comps*window not a multiple of 8 (e.g. the class defaults 53 x 20): TMA needs a 16-byte row
            # pitch, so the operand is a persistent copy with the pitch rounded up (the activations already are)
            w_enc = _pitched_copy(model, w_enc)
        enc = gemm_bf16(xb, w_enc, out_dtype=torch.bfloat16, bias_n=enc_l.bias.detach(), relu=True, out_t=encT)
        # 3. input projection for every timestep, both directions
        pad = None
        if H % 32:
            pad = _Padded.of(model, x.device)
            pad.fill(model)
            H = pad.Hp
        w_ih = [_w16(model, c.i2h.weight) for c in cells] if pad is None else [pad.w_ih]
        PX = torch.empty(S * B, D * 4 * H, device=x.device, dtype=torch.float32)
        if pad is not None:
            gemm_bf16(enc, pad.w_ih, out=PX, bias_n=pad.b_ih)
        elif D == 2 and _adjacent(w_ih[0], w_ih[1]) and _adjacent(cells[0].i2h.bias.detach(), cells[1].i2h.bias.detach()):
            w_cat = _span(w_ih[0], (D * 4 * H, E), (E, 1))
            b_cat = _span(cells[0].i2h.bias.detach(), (D * 4 * H,), (1,))
            gemm_bf16(enc, w_cat, out=PX, bias_n=b_cat)
        else:
            for d in range(D):
                gemm_bf16(enc, w_ih[d], out=PX[:, d * 4 * H:(d + 1) * 4 * H], bias_n=cells[d].i2h.bias.detach())
        # 4. recurrence
        if pad is None:
            whh = [_w16(model, c.h2h.weight) for c in cells]
            bhh = [c.h2h.bias.detach() for c in cells]
        else:
            whh, bhh = [pad.whh[d] for d in range(D)], [pad.b_hh[d] for d in range(D)]
        whhT = None
        if torch.is_grad_enabled() or any(ctx.needs_input_grad):
            # K-major copy of W_hh for the BPTT kernel's tensor-memory operand, made on a side stream
            # while the recurrence runs (a parallel branch when the step is a CUDA graph)
            main = torch.cuda.current_stream(x.device)
            s1, _ = _side_streams(x.device)
            s1.wait_stream(main)
            with torch.cuda.stream(s1):
                whhT = [transpose_bf16(w) for w in whh]
                ev_whhT = torch.cuda.Event()
                ev_whhT.record(s1)
        Hmean, hT, cT, C, Hprev = lstm_forward(PX, whh, bhh, B, S, H, D)
        if whhT is not None:
            main.wait_event(ev_whhT)
        if pad is not None:
            hT, cT = pad.unpad(hT), pad.unpad(cT)
        # 5. head
        N0 = D * H
        p_drop = cls[0].p
        gen_drop = None
        if training and p_drop > 0:
            # the head kernel draws the dropout mask itself from (seed, device-resident step counter): no RNG
            # launches, and the captured step replays with a fresh mask every time
            st = model.__dict__.setdefault("_drop_state", {})
            key = (x.device.type, x.device.index)
            if key not in st:
                st[key] = (torch.zeros((), device=x.device, dtype=torch.int64), torch.initial_seed())
            gen_drop = (float(p_drop), st[key][1], st[key][0])
        W1 = _w16(model, cls[1].weight) if pad is None else pad.W1
        W2, W3 = _w16(model, cls[4].weight), _w16(model, cls[6].weight)
        bn = cls[2]
        logits, head_saved = head_forward(Hmean, None, W1, cls[1].bias.detach(), bn.weight.detach(), bn.bias.detach(),
                                          bn.running_mean, bn.running_var, W2, cls[4].bias.detach(), W3,
                                          cls[6].bias.detach(), training, float(bn.momentum or 0.1), float(bn.eps),
                                          gen_drop=gen_drop, num_batches_tracked=bn.num_batches_tracked if training else None)

        ctx.set_materialize_grads(False)       # no zero tensors for the (unused) gradients of hT / cT
        ctx.model = model
        ctx.dims = (B, S, E, H, D)
        ctx.pad = pad
        ctx.saved = (xb, enc, encT, PX, C, Hprev, whh, w_ih, head_saved)
        ctx.whhT = whhT
        ctx.mark_non_differentiable(hT, cT)
        return logits, hT, cT

    @staticmethod
    def backward(ctx, dlogits, _dh, _dc):
        lib = native.require_gpu()
        if dlogits is None:      # only the (non-differentiable) final states were used downstream
            return (None, None) + (None,) * (len(ctx.needs_input_grad) - 2)
        model = ctx.model
        B, S, E, H, D = ctx.dims
        (xb, enc, encT, P, C, Hprev, whh, w_ih, head_saved) = ctx.saved
        ctx.saved = None
        enc_l = model.encoder[0]
        cells = list(model.lstm.lstms)
        cls = model.classifier
        bn = cls[2]
        dev = dlogits.device

        capture = getattr(model, "_dad_capture", False)
        pad = ctx.pad                               # H here is the padded size when pad is set
        if pad is not None:
            pad.stage.zero_()
        # 6. head backward (accumulates into the parameter gradients)
        grads = dict(dW1=_grad_buf(cls[1].weight) if pad is None else pad.g_W1,
                     db1=_grad_buf(cls[1].bias), dbn_w=_grad_buf(bn.weight),
                     dbn_b=_grad_buf(bn.bias), dW2=_grad_buf(cls[4].weight), db2=_grad_buf(cls[4].bias),
                     dW3=_grad_buf(cls[6].weight), db3=_grad_buf(cls[6].bias))
        dHmean = head_backward(dlogits, head_saved, grads, accumulate=True, want_dz2=capture)
        if capture:
            # rankDAD: ship factors of (delta, activation) instead of the weight gradients.  The head pairs are
            # fp32 tensors the head kernels already produced (dz1 / dz2 are exported by the backward launches);
            # the aggregation step packs them to bf16 in both orientations with one launch (dnb_dad_pack).
            hs = head_saved
            pairs = model.__dict__.setdefault("_dad_pairs", {})
            pairs.setdefault("classifier.6", []).append({"D": hs["dlogits32"], "A": hs["a2"]})
            pairs.setdefault("classifier.4", []).append({"D": hs["dz2"], "A": hs["a1"]})
            pairs.setdefault("classifier.1", []).append({"D": hs["dz1"], "A": hs["a0"] if pad is None else pad.unpad(hs["a0"])})

        # Independent pieces of the backward run on side streams (parallel branches when the step is
        # captured in a CUDA graph): the transposes that do not depend on dP hide under the BPTT kernel
        # (which occupies only 2 x H/32 SMs), and the three weight-gradient chains run side by side.
        main = torch.cuda.current_stream(dev)
        s1, s2 = _side_streams(dev)
        if pad is not None:
            w_cat = pad.w_ih
        elif D == 2 and _adjacent(w_ih[0], w_ih[1]):
            w_cat = _span(w_ih[0], (D * 4 * H, E), (E, 1))
        else:
            w_cat = torch.cat(w_ih, 0)
        s1.wait_stream(main)
        with torch.cuda.stream(s1):
            HprevT = transpose_bf16(Hprev)          # [D*H,  S*B]
            w_ihT = transpose_bf16(w_cat)           # [E, D*4H]
            xT = transpose_bf16(xb)                 # [comps*window, S*B]
            ev_side = torch.cuda.Event()
            ev_side.record(s1)

        # 7. BPTT
        if pad is None:
            db_ih = [_grad_buf(c.i2h.bias) for c in cells]
            db_hh = [_grad_buf(c.h2h.bias) for c in cells]
        else:
            db_ih, db_hh = [pad.g_b_ih[d] for d in range(D)], [pad.g_b_hh[d] for d in range(D)]
        dP = lstm_backward(P, C, whh, dHmean, B, S, H, D, db_ih, db_hh, accumulate=True, whhT=ctx.whhT)
        ev_dp = torch.cuda.Event()
        ev_dp.record(main)

        if capture:
            # LSTM / encoder pairs: every (delta, activation) matrix in both orientations, bf16, exactly the
            # operands the weight-gradient GEMMs of the dSGD path consume — the rankDAD kernels take their place.
            BSp = (S * B + 7) // 8 * 8
            s2.wait_event(ev_dp)
            s2.wait_event(ev_side)
            with torch.cuda.stream(s2):
                d_enc = torch.empty(S * B, E, device=dev, dtype=torch.bfloat16)
                d_encT = torch.empty(E, BSp, device=dev, dtype=torch.bfloat16)[:, :S * B]
                gemm_bf16(dP, w_ihT, out=d_enc, out_t=d_encT, mask=enc, colsum=_grad_buf(enc_l.bias))
            dPT = transpose_bf16(dP)                # [D*4H, S*B]
            main.wait_stream(s1)
            main.wait_stream(s2)
            for dd in range(D):
                sl = slice(dd * 4 * H, (dd + 1) * 4 * H)
                if pad is None:
                    delta, deltaT = dP[:, sl], dPT[sl]
                    hp, hpT = Hprev[:, dd * H:(dd + 1) * H], HprevT[dd * H:(dd + 1) * H]
                    e_ih = {"D": delta, "DT": deltaT, "A": enc, "AT": encT}
                    e_hh = {"D": delta, "DT": deltaT, "A": hp, "AT": hpT}
                else:       # zero-padded hidden units: hand over compact copies (packed by the aggregation step)
                    delta = pad.gates(dP, dd)
                    e_ih = {"D": delta, "A": enc}
                    e_hh = {"D": delta, "A": Hprev[:, dd * H:dd * H + pad.H]}
                pairs.setdefault(f"lstm.lstms.{dd}.i2h", []).append(e_ih)
                pairs.setdefault(f"lstm.lstms.{dd}.h2h", []).append(e_hh)
            pairs.setdefault("encoder.0", []).append({"D": d_enc, "DT": d_encT, "A": xb, "AT": xT})
            if pad is not None:
                pad.scatter_grads(model)
            return (None, None) + (None,) * (len(ctx.needs_input_grad) - 2)

        # branch A (s2): through the input projection and the encoder ReLU (bias gradient as column sums,
        # result stored transposed = K-major operand of the next GEMM), then the encoder weights
        BSp = (S * B + 7) // 8 * 8
        s2.wait_event(ev_dp)
        s2.wait_event(ev_side)
        with torch.cuda.stream(s2):
            d_encT = torch.empty(E, BSp, device=dev, dtype=torch.bfloat16)[:, :S * B]
            gemm_bf16(dP, w_ihT, want_out=False, out_t=d_encT, mask=enc, colsum=_grad_buf(enc_l.bias))
            gemm_bf16(d_encT, xT, out=_grad_buf(enc_l.weight), accumulate=True, split_k=0)
        # main: K-major copy of dP, then the input-projection weights
        dPT = transpose_bf16(dP)                    # [D*4H, S*B]
        ev_dpt = torch.cuda.Event()
        ev_dpt.record(main)
        # branch B (s1): recurrent weights
        s1.wait_event(ev_dpt)
        with torch.cuda.stream(s1):
            for dd, c in enumerate(cells):
                gemm_bf16(dPT[dd * 4 * H:(dd + 1) * 4 * H], HprevT[dd * H:(dd + 1) * H],
                          out=_grad_buf(c.h2h.weight) if pad is None else pad.g_whh[dd], accumulate=True, split_k=0)
        g_ih = [_grad_buf(c.i2h.weight) for c in cells]
        if pad is not None:
            gemm_bf16(dPT, encT, out=pad.g_w_ih, accumulate=True, split_k=0)
        elif D == 2 and _adjacent(g_ih[0], g_ih[1]):
            gemm_bf16(dPT, encT, out=_span(g_ih[0], (D * 4 * H, E), (E, 1)), accumulate=True, split_k=0)
        else:
            for dd in range(D):
                gemm_bf16(dPT[dd * 4 * H:(dd + 1) * 4 * H], encT, out=g_ih[dd], accumulate=True, split_k=0)
        main.wait_stream(s1)
        main.wait_stream(s2)
        if pad is not None:
            pad.scatter_grads(model)
        return (None, None) + (None,) * (len(ctx.needs_input_grad) - 2)


def icalstm_forward(model, x):
    params = [p for p in model.parameters() if p.requires_grad]
    logits, hT, cT = _ICALstmFn.apply(x, model, *params)
    return logits, (hT, cT)
