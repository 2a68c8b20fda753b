"""ICA time-course bidirectional LSTM classifier.

Reproduces the reference model (``comps/icalstm/models.py:5-110``) including its quirks
(SURVEY §9):

* gate layout along the 4H axis is ``[i | f | o | g]``; ``i, f, o`` pass through **sigmoid
  twice**, ``g`` through tanh once (``models.py:32-38``);
* ``hidden_size`` is split across the two directions (``:55``); ``num_layers`` is accepted and
  ignored (``:52``);
* the reverse direction runs on the time-flipped input and its outputs are concatenated
  **without flipping back** (``:62-65``);
* per-window encoder ``Linear(comps*window → input_size) + ReLU`` (``:87``), mean over time
  (``:109``), classifier ``Dropout(.25) → Linear(H,256) → BatchNorm1d(256) → ReLU →
  Linear(256,64) → ReLU → Linear(64,num_cls)`` (``:95-103``).

Parameter/buffer names equal the reference's (``encoder.0.*``, ``lstm.lstms.<d>.i2h|h2h.*``,
``classifier.<1|2|4|6>.*``), so ``state_dict``s interchange.

The eager path differs from the reference only in *how* it computes: the encoder is one
batched GEMM instead of a Python loop over the batch (``:107``) and the input projection
``i2h`` is hoisted out of the time loop.  With ``fused=True`` on CUDA the forward/backward run
in the sm_100a kernels of ``ops.lstm`` / ``ops.gemm`` / ``ops.head``.
"""
from __future__ import annotations

import torch
import torch.nn as nn


class LSTMCell(nn.Module):
    def __init__(self, input_size, hidden_size, bias=True):
        super().__init__()
        self.input_size, self.hidden_size = int(input_size), int(hidden_size)
        self.i2h = nn.Linear(self.input_size, 4 * self.hidden_size, bias=bias)
        self.h2h = nn.Linear(self.hidden_size, 4 * self.hidden_size, bias=bias)

    def forward(self, x, h=None):
        B, T, _ = x.shape
        H = self.hidden_size
        if h is None:
            h_t = x.new_zeros(B, H)
            c_t = x.new_zeros(B, H)
        else:
            h_t, c_t = h
        xp = self.i2h(x)                                   # [B,T,4H], hoisted out of the loop
        outs = []
        for t in range(T):
            pre = xp[:, t] + self.h2h(h_t)
            ifo = torch.sigmoid(torch.sigmoid(pre[:, :3 * H]))   # double sigmoid (reference quirk)
            g = torch.tanh(pre[:, 3 * H:])
            i, f, o = ifo[:, :H], ifo[:, H:2 * H], ifo[:, 2 * H:]
            c_t = f * c_t + i * g
            h_t = o * torch.tanh(c_t)
            outs.append(h_t)
        return torch.stack(outs, 1), (h_t, c_t)


class LSTM(nn.Module):
    def __init__(self, input_size, hidden_size, bidirectional=True, num_layers=1, bias=True):
        super().__init__()
        self.input_size = int(input_size)
        self.num_layers = num_layers            # stored, unused — as in the reference
        self.bidirectional = bool(bidirectional)
        self.num_direction = 2 if self.bidirectional else 1
        self.hidden_size = int(hidden_size) // self.num_direction
        self.lstms = nn.ModuleList([LSTMCell(self.input_size, self.hidden_size, bias) for _ in range(self.num_direction)])

    def forward(self, x, h=None):
        seq, (h_t, c_t) = self.lstms[0](x, h)
        if self.bidirectional:
            rseq, (rh, rc) = self.lstms[1](torch.flip(x, (1,)), h)
            seq = torch.cat([seq, rseq], 2)     # reverse outputs stay in processing order
            h_t = torch.cat([h_t, rh], 1)
            c_t = torch.cat([c_t, rc], 1)
        return seq, (h_t, c_t)


class ICALstm(nn.Module):
    def __init__(self, input_size=256, hidden_size=256, bidirectional=True, num_cls=2, num_comps=53, window_size=20,
                 num_layers=1, fused=True):
        super().__init__()
        self.input_size, self.hidden_size = int(input_size), int(hidden_size)
        self.num_comp, self.window_size = int(num_comps), int(window_size)
        self.num_cls = int(num_cls)
        self.fused = fused
        self.encoder = nn.Sequential(nn.Linear(self.num_comp * self.window_size, self.input_size), nn.ReLU())
        self.lstm = LSTM(self.input_size, self.hidden_size, bidirectional, num_layers)
        self.classifier = nn.Sequential(
            nn.Dropout(0.25), nn.Linear(self.hidden_size, 256), nn.BatchNorm1d(256), nn.ReLU(),
            nn.Linear(256, 64), nn.ReLU(), nn.Linear(64, self.num_cls))

    def flat_param_order(self):
        """Order used when the parameters are packed into one flat buffer: the two directions'
        i2h weights (and biases) adjacent, so a single GEMM projects both directions."""
        cells = list(self.lstm.lstms)
        order = [self.encoder[0].weight] + [c.i2h.weight for c in cells] + [c.h2h.weight for c in cells]
        order += [self.classifier[1].weight, self.classifier[4].weight, self.classifier[6].weight]
        order += [c.i2h.bias for c in cells] + [c.h2h.bias for c in cells]
        return order

    def dad_capture_keys(self):
        """Linear modules whose (delta, activation) pairs the fused backward can hand to rankDAD."""
        return [n for n, m in self.named_modules() if isinstance(m, nn.Linear)]

    def forward_eager(self, x):
        B, S = x.shape[0], x.shape[1]
        e = self.encoder(x.reshape(B * S, -1)).view(B, S, -1)
        o, h = self.lstm(e)
        return self.classifier(o.mean(1)), h

    def forward(self, x):
        if self.fused and x.is_cuda:
            from ..ops import icalstm_fused as _f
            if _f.supported(self, x):
                return _f.icalstm_forward(self, x)
            if not getattr(self, "_warned_eager", False):
                self._warned_eager = True
                from ..utils import logger
                logger.warn(f"ICALstm: input {tuple(x.shape)} / hidden {self.lstm.hidden_size} per direction is outside the "
                            "fused kernels' envelope (batch <= 64, hidden <= 256 per direction, input_size % 8 == 0); "
                            "running the eager PyTorch time loop", True)
        return self.forward_eager(x)
