"""Classifier-head kernels (forward + backward) vs torch fp32 on identical inputs."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mods(N0, N1, N2, N3):
    torch.manual_seed(0)
    head = torch.nn.Sequential(torch.nn.Linear(N0, N1), torch.nn.BatchNorm1d(N1), torch.nn.ReLU(),
                               torch.nn.Linear(N1, N2), torch.nn.ReLU(), torch.nn.Linear(N2, N3)).cuda()
    with torch.no_grad():
        for m in head:
            if isinstance(m, torch.nn.Linear):
                m.weight.copy_(m.weight.to(torch.bfloat16).float())
        head[1].weight.uniform_(0.5, 1.5)
        head[1].bias.uniform_(-0.2, 0.2)
        head[1].running_mean.uniform_(-0.1, 0.1)
        head[1].running_var.uniform_(0.8, 1.2)
    return head


@pytest.mark.parametrize("B", [16, 7, 33, 64])
@pytest.mark.parametrize("training", [True, False])
@pytest.mark.parametrize("use_drop", [False, True])
def test_head_forward_backward(B, training, use_drop):
    from dinunet_implementations_b200.ops.head import head_forward, head_backward
    N0, N1, N2, N3 = 384, 256, 64, 2
    head = _mods(N0, N1, N2, N3)
    head.train(training)
    x = torch.randn(B, N0, device="cuda", requires_grad=True)
    drop = (torch.rand(B, N0, device="cuda") > 0.25).float() / 0.75 if use_drop else None
    rm0, rv0 = head[1].running_mean.clone(), head[1].running_var.clone()
    rm, rv = rm0.clone(), rv0.clone()
    logits, saved = head_forward(x.detach(), drop, head[0].weight.detach().to(torch.bfloat16), head[0].bias.detach(),
                                 head[1].weight.detach(), head[1].bias.detach(), rm, rv,
                                 head[3].weight.detach().to(torch.bfloat16), head[3].bias.detach(),
                                 head[5].weight.detach().to(torch.bfloat16), head[5].bias.detach(), training)
    ref = head(x * drop if use_drop else x)
    torch.cuda.synchronize()
    assert torch.allclose(logits, ref, atol=2e-4, rtol=1e-4), (logits - ref).abs().max()
    if training:
        assert torch.allclose(rm, head[1].running_mean, atol=1e-5) and torch.allclose(rv, head[1].running_var, atol=1e-5)
    G = torch.randn(B, N3, device="cuda")
    (ref * G).sum().backward()
    grads = {k: torch.zeros_like(p) for k, p in dict(dW1=head[0].weight, db1=head[0].bias, dbn_w=head[1].weight,
                                                     dbn_b=head[1].bias, dW2=head[3].weight, db2=head[3].bias,
                                                     dW3=head[5].weight, db3=head[5].bias).items()}
    dx = head_backward(G, saved, grads, accumulate=True)
    torch.cuda.synchronize()
    pairs = dict(dW1=head[0].weight, db1=head[0].bias, dbn_w=head[1].weight, dbn_b=head[1].bias, dW2=head[3].weight,
                 db2=head[3].bias, dW3=head[5].weight, db3=head[5].bias)
    for k, p in pairs.items():
        s = max(p.grad.abs().max().item(), 1e-3)
        assert torch.allclose(grads[k], p.grad, atol=2e-3 * s + 1e-5), (k, (grads[k] - p.grad).abs().max().item(), s)
    s = max(x.grad.abs().max().item(), 1e-3)
    assert torch.allclose(dx, x.grad, atol=2e-3 * s), (dx - x.grad).abs().max()
    # accumulate contract: a second call adds
    head_backward(G, saved, grads, accumulate=True)
    torch.cuda.synchronize()
    assert torch.allclose(grads["dW3"], 2 * head[5].weight.grad, atol=4e-3 * max(head[5].weight.grad.abs().max().item(), 1e-3))


def test_head_in_kernel_dropout_and_batch_counter():
    """gen_drop: the kernel draws the mask (no RNG launches), hands it to the backward, advances the counter."""
    from dinunet_implementations_b200.ops.head import head_forward, head_backward
    B, N0, N1, N2, N3 = 32, 384, 256, 64, 2
    head = _mods(N0, N1, N2, N3)
    head.train(True)
    x = torch.randn(B, N0, device="cuda", requires_grad=True)
    state = torch.zeros((), device="cuda", dtype=torch.int64)
    nbt = torch.zeros((), device="cuda", dtype=torch.int64)
    args = (head[0].weight.detach().to(torch.bfloat16), head[0].bias.detach(), head[1].weight.detach(), head[1].bias.detach(),
            head[1].running_mean.clone(), head[1].running_var.clone(), head[3].weight.detach().to(torch.bfloat16),
            head[3].bias.detach(), head[5].weight.detach().to(torch.bfloat16), head[5].bias.detach(), True)
    masks = []
    for step in range(3):
        logits, saved = head_forward(x.detach(), None, *args, gen_drop=(0.25, 1234, state), num_batches_tracked=nbt)
        torch.cuda.synchronize()
        m = saved["drop"]
        assert all(v == 0.0 or abs(v - 1.0 / 0.75) < 1e-6 for v in torch.unique(m).tolist())
        keep = (m > 0).float().mean().item()
        assert abs(keep - 0.75) < 0.03, keep
        assert torch.equal(saved["a0"], x.detach() * m)
        assert int(state) == step + 1 and int(nbt) == step + 1
        masks.append(m.clone())
    assert not torch.equal(masks[0], masks[1]) and not torch.equal(masks[1], masks[2])
    # same (seed, step) -> same mask; the emitted mask reproduces the torch result
    state2 = torch.zeros((), device="cuda", dtype=torch.int64)
    _, saved2 = head_forward(x.detach(), None, *args, gen_drop=(0.25, 1234, state2))
    assert torch.equal(saved2["drop"], masks[0])
    ref = head(x * m)
    assert torch.allclose(logits, ref, atol=2e-4, rtol=1e-4)
    G = torch.randn(B, N3, device="cuda")
    (ref * G).sum().backward()
    grads = {k: torch.zeros_like(p) for k, p in dict(dW1=head[0].weight, db1=head[0].bias, dbn_w=head[1].weight,
                                                     dbn_b=head[1].bias, dW2=head[3].weight, db2=head[3].bias,
                                                     dW3=head[5].weight, db3=head[5].bias).items()}
    dx = head_backward(G, saved, grads, accumulate=True)
    s = max(x.grad.abs().max().item(), 1e-3)
    assert torch.allclose(dx, x.grad, atol=2e-3 * s)
