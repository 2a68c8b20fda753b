"""Our model classes vs the reference's own ``models.py`` files (loaded by path): same
state_dict keys, same forward values, same gradients, quirks included."""
import torch

from dinunet_implementations_b200.models import MSANNet, ICALstm


def _grads(net, out_fn):
    net.zero_grad()
    out_fn().backward()
    return {k: p.grad.clone() for k, p in net.named_parameters()}


def test_msannet_matches_reference(ref_fs_models):
    torch.manual_seed(0)
    ref = ref_fs_models.MSANNet(66, [256, 128, 64, 32], 2)
    ours = MSANNet(66, [256, 128, 64, 32], 2)
    assert list(ref.state_dict().keys()) == list(ours.state_dict().keys())
    ours.load_state_dict(ref.state_dict())
    x = torch.rand(16, 66)
    for mode in ("train", "eval"):               # batch statistics in eval too
        getattr(ref, mode)(); getattr(ours, mode)()
        assert torch.allclose(ref(x), ours(x), atol=1e-6)
    gr = _grads(ref, lambda: ref(x).logsumexp(1).sum())
    go = _grads(ours, lambda: ours(x).logsumexp(1).sum())
    for k in gr:
        assert torch.allclose(gr[k], go[k], atol=1e-6), k


def test_icalstm_matches_reference_including_quirks(ref_ica_models):
    torch.manual_seed(1)
    kw = dict(input_size=64, hidden_size=48, num_comps=10, window_size=5, num_cls=2)
    ref = ref_ica_models.ICALstm(**kw)
    ours = ICALstm(**kw)
    assert list(ref.state_dict().keys()) == list(ours.state_dict().keys())   # incl. BN running stats buffers
    ours.load_state_dict(ref.state_dict())
    x = torch.randn(6, 9, 10, 5)
    ref.eval(); ours.eval()
    (lr_, (hr, cr)), (lo, (ho, co)) = ref(x), ours(x)
    assert torch.allclose(lr_, lo, atol=1e-5)
    assert torch.allclose(hr, ho, atol=1e-6) and torch.allclose(cr, co, atol=1e-6)
    # sequence outputs: reverse half must be in processing (time-reversed) order
    e = ours.encoder(x.reshape(6 * 9, -1)).view(6, 9, -1)
    so, _ = ours.lstm(e)
    sr, _ = ref.lstm(e)
    assert torch.allclose(so, sr, atol=1e-6)
    gr = _grads(ref, lambda: ref(x)[0].logsumexp(1).sum())
    go = _grads(ours, lambda: ours(x)[0].logsumexp(1).sum())
    for k in gr:
        assert torch.allclose(gr[k], go[k], atol=2e-5), (k, (gr[k] - go[k]).abs().max())


def test_icalstm_full_size_shapes():
    net = ICALstm(input_size=256, hidden_size=384, num_comps=100, window_size=10, num_cls=2)
    n = sum(p.numel() for p in net.parameters())
    assert n == 1063106
    assert net.lstm.lstms[0].h2h.weight.shape == (768, 192)
