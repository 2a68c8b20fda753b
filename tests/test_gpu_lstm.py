"""Persistent tcgen05 LSTM recurrence (forward + BPTT) vs the eager fp32 module that is itself
checked against the reference's models.py on CPU (tests/test_models_vs_reference.py)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(B, S, H, D, E=64, seed=0):
    from dinunet_implementations_b200.models import LSTM
    torch.manual_seed(seed)
    lstm = LSTM(E, D * H, bidirectional=(D == 2)).cuda()
    with torch.no_grad():
        for cell in lstm.lstms:     # the kernel keeps W_hh in bf16: give the oracle the same values
            cell.h2h.weight.copy_(cell.h2h.weight.to(torch.bfloat16).float())
    enc = torch.randn(B, S, E, device="cuda")
    return lstm, enc


def _px(lstm, enc, B, S, H, D):
    # time-major rows: row = t*B + b ; columns: [dir0 4H | dir1 4H]
    cols = [cell.i2h(enc) for cell in lstm.lstms]                     # each [B,S,4H]
    px = torch.cat(cols, 2).permute(1, 0, 2).reshape(S * B, D * 4 * H).contiguous()
    return px


@pytest.mark.parametrize("nbe", ["auto", "4", "8", "16"])      # batch sub-tile per cluster (csrc/lstm.cu pick_nbe)
@pytest.mark.parametrize("B,S,H,D", [(16, 98, 192, 2), (5, 7, 32, 1), (16, 12, 64, 2), (40, 9, 96, 2), (3, 20, 256, 2)])
def test_lstm_forward_and_backward(B, S, H, D, nbe, monkeypatch):
    from dinunet_implementations_b200.ops.lstm import lstm_forward, lstm_backward
    if nbe == "auto":
        monkeypatch.delenv("DNB_LSTM_NBE", raising=False)
    else:
        monkeypatch.setenv("DNB_LSTM_NBE", nbe)
    lstm, enc = _setup(B, S, H, D)
    enc.requires_grad_(True)
    seq, (hT_ref, cT_ref) = lstm(enc)
    hmean_ref = seq.mean(1)
    G = torch.randn(B, D * H, device="cuda")
    (hmean_ref * G).sum().backward()

    with torch.no_grad():
        px = _px(lstm, enc, B, S, H, D)
        whh = [c.h2h.weight.detach().to(torch.bfloat16).contiguous() for c in lstm.lstms]
        bhh = [c.h2h.bias.detach().contiguous() for c in lstm.lstms]
        Hmean, hT, cT, C, Hprev = lstm_forward(px, whh, bhh, B, S, H, D)
        torch.cuda.synchronize()
        assert torch.allclose(Hmean, hmean_ref, atol=2e-2), (Hmean - hmean_ref).abs().max()
        assert torch.allclose(hT, hT_ref, atol=3e-2), (hT - hT_ref).abs().max()
        assert torch.allclose(cT, cT_ref, atol=5e-2), (cT - cT_ref).abs().max()

        db_ih = [torch.zeros(4 * H, device="cuda") for _ in range(D)]
        db_hh = [torch.zeros(4 * H, device="cuda") for _ in range(D)]
        dP = lstm_backward(px, C, whh, G.contiguous(), B, S, H, D, db_ih, db_hh)
        # same launch with the recurrent weights handed over pre-transposed (no in-kernel transposition)
        db2 = [torch.zeros(4 * H, device="cuda") for _ in range(D)]
        dP2 = lstm_backward(px, C, whh, G.contiguous(), B, S, H, D, db2, None, whhT=[w.t().contiguous() for w in whh])
        torch.cuda.synchronize()
        assert torch.equal(dP2, dP) and all(torch.allclose(x, y, atol=1e-6, rtol=1e-5) for x, y in zip(db2, db_ih))
        dPf = dP.float()
        enc_rows = enc.detach().permute(1, 0, 2).reshape(S * B, -1)
        for d, cell in enumerate(lstm.lstms):
            dPd = dPf[:, d * 4 * H:(d + 1) * 4 * H]
            scale = max(cell.i2h.weight.grad.abs().max().item(), 1e-3)
            dW_ih = dPd.t() @ enc_rows
            assert torch.allclose(dW_ih, cell.i2h.weight.grad, atol=0.03 * scale), \
                ((dW_ih - cell.i2h.weight.grad).abs().max(), scale)
            dW_hh = dPd.t() @ Hprev[:, d * H:(d + 1) * H].float()
            s2 = max(cell.h2h.weight.grad.abs().max().item(), 1e-3)
            assert torch.allclose(dW_hh, cell.h2h.weight.grad, atol=0.03 * s2), \
                ((dW_hh - cell.h2h.weight.grad).abs().max(), s2)
            s3 = max(cell.i2h.bias.grad.abs().max().item(), 1e-3)
            assert torch.allclose(db_ih[d], cell.i2h.bias.grad, atol=0.03 * s3)
            assert torch.allclose(db_ih[d], db_hh[d], atol=1e-4 * s3 + 1e-6)
        # gradient w.r.t. the projected input = dP itself -> d(enc) through i2h
        d_enc = torch.zeros_like(enc_rows)
        for d, cell in enumerate(lstm.lstms):
            d_enc += dPf[:, d * 4 * H:(d + 1) * 4 * H] @ cell.i2h.weight.detach()
        d_enc = d_enc.view(S, B, -1).permute(1, 0, 2)
        s4 = max(enc.grad.abs().max().item(), 1e-3)
        assert torch.allclose(d_enc, enc.grad, atol=0.03 * s4), ((d_enc - enc.grad).abs().max(), s4)
