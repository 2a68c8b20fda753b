"""Host-side pieces of the device mode that do not need a GPU: flat parameter layout, the zero-padding index
logic of the fused ICA-LSTM path, the wire-format mapping, the HostEngine update rule."""
import torch

from dinunet_implementations_b200.models import ICALstm
from dinunet_implementations_b200.parallel.dsgd_fused import ALIGN, flat_layout, ordered_params, wire_dtype_from_bits


def test_flat_layout_is_aligned_and_ordered_for_one_projection_gemm():
    m = ICALstm(input_size=32, hidden_size=64, num_comps=8, window_size=5, num_cls=2)
    ps = ordered_params(m)
    assert len(ps) == len([p for p in m.parameters() if p.requires_grad]) and len({id(p) for p in ps}) == len(ps)
    offs, total = flat_layout(ps)
    assert all(o % ALIGN == 0 for o in offs) and total % ALIGN == 0
    for o, p, nxt in zip(offs, ps, offs[1:] + [total]):
        assert nxt - o >= p.numel() and nxt - o < p.numel() + ALIGN
    # both directions' i2h weights (and biases) are adjacent: with 4H*E a multiple of ALIGN they are contiguous,
    # which is what lets ONE GEMM project both directions
    cells = list(m.lstm.lstms)
    i0, i1 = [next(k for k, p in enumerate(ps) if p is c.i2h.weight) for c in cells]
    assert i1 == i0 + 1 and offs[i1] - offs[i0] == cells[0].i2h.weight.numel()
    b0, b1 = [next(k for k, p in enumerate(ps) if p is c.i2h.bias) for c in cells]
    assert b1 == b0 + 1


def test_wire_dtype_from_bits():
    assert wire_dtype_from_bits("16") is torch.bfloat16 and wire_dtype_from_bits(16, prefer_bf16=False) is torch.float16
    assert wire_dtype_from_bits("32") is torch.float32 and wire_dtype_from_bits(None) is torch.float32


def test_padding_index_logic_of_the_fused_path():
    """``_Padded`` (hidden 348 -> 174 per direction -> 192): what the kernels see is the real tensors embedded in
    zeros, and gradients come back from exactly the embedded positions (checked here with CPU tensors)."""
    from dinunet_implementations_b200.ops.icalstm_fused import _Padded, _ceil32
    m = ICALstm(input_size=16, hidden_size=2 * 20, num_comps=4, window_size=4, num_cls=2)   # H = 20 -> Hp = 32
    pad = _Padded(m, torch.device("cpu"))
    D, H, Hp, E, N1 = pad.D, pad.H, pad.Hp, pad.E, pad.N1
    assert (H, Hp, _ceil32(174)) == (20, 32, 192)
    # unpad / gates pick the real units out of padded activations
    t = torch.arange(3 * D * Hp, dtype=torch.float32).view(3, D * Hp)
    u = pad.unpad(t)
    assert u.shape == (3, D * H) and torch.equal(u[:, :H], t[:, :H]) and torch.equal(u[:, H:], t[:, Hp:Hp + H])
    g = torch.arange(2 * D * 4 * Hp, dtype=torch.float32).view(2, D * 4 * Hp)
    g1 = pad.gates(g, 1)
    assert g1.shape == (2, 4 * H)
    for q in range(4):
        assert torch.equal(g1[:, q * H:(q + 1) * H], g[:, (4 + q) * Hp:(4 + q) * Hp + H])
    # scatter_grads adds the staged values at the embedded positions only
    for p in m.parameters():
        p.grad = None
    pad.stage.copy_(torch.arange(pad.stage.numel(), dtype=torch.float32) % 97 + 1)
    pad.scatter_grads(m)
    c1 = list(m.lstm.lstms)[1]
    assert torch.equal(c1.i2h.weight.grad.view(4, H, E), pad.g_w_ih.view(D, 4, Hp, E)[1, :, :H])
    assert torch.equal(c1.h2h.weight.grad.view(4, H, H), pad.g_whh.view(D, 4, Hp, Hp)[1, :, :H, :H])
    assert torch.equal(c1.h2h.bias.grad.view(4, H), pad.g_b_hh.view(D, 4, Hp)[1, :, :H])
    assert torch.equal(m.classifier[1].weight.grad.view(N1, D, H), pad.g_W1.view(N1, D, Hp)[:, :, :H])
    pad.scatter_grads(m)                                              # accumulates
    assert torch.equal(c1.i2h.bias.grad.view(4, H), 2 * pad.g_b_ih.view(D, 4, Hp)[1, :, :H])


def test_host_engine_dsgd_matches_plain_adam():
    """One process, dSGD: the HostEngine step is zero_grad -> backward -> Adam, nothing else."""
    from dinunet_implementations_b200.parallel.host_engine import HostEngine
    from dinunet_implementations_b200.tasks.fs import FreeSurferTrainer
    cache = {"task_id": "FS-Classification", "input_size": 66, "hidden_sizes": [16, 8], "num_class": 2, "seed": 3,
             "learning_rate": 1e-2, "monitor_metric": "auc", "gpus": [], "agg_engine": "dSGD"}
    tr = FreeSurferTrainer(cache=dict(cache), input={}, state={})
    tr.init_nn(init_optim=False)
    ref = FreeSurferTrainer(cache=dict(cache), input={}, state={})
    ref.init_nn(init_optim=False)
    eng = HostEngine(tr, track_metrics=True)
    opt = torch.optim.Adam(ref.nn["fs_net"].parameters(), lr=1e-2)
    g = torch.Generator().manual_seed(0)
    for _ in range(3):
        batch = {"inputs": torch.rand(8, 66, generator=g), "labels": torch.randint(0, 2, (8,), generator=g), "ix": torch.arange(8)}
        eng.train_step(batch)
        opt.zero_grad()
        ref.train_mode()
        ref.iteration(batch)["loss"].backward()
        opt.step()
    for a, b in zip(tr.nn["fs_net"].parameters(), ref.nn["fs_net"].parameters()):
        assert torch.allclose(a, b, atol=1e-7)
    avg, met = eng.epoch_metrics()
    assert avg.get()[0] > 0
