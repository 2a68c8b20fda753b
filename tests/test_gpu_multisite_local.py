"""Multi-site logic on ONE GPU: S simulated sites share a ``SiteGroup.local`` arena and the fused dSGD kernel
(``csrc/dsgd.cu``) aggregates them exactly as it does over NVLink (same kernel, flag barrier off).  Covers what the
single-GPU test box can say about federated correctness: all sites stay bit-identical, the loss curve follows the
reference math of file mode (mean of the sites' gradients + torch.optim.Adam) step for step, and a
separable problem is actually learnt by the fused multi-site path."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _bind_sites(models, wire=torch.float32, lr=1e-2, shadow=False):
    from dinunet_implementations_b200.parallel.comm import SiteGroup
    from dinunet_implementations_b200.parallel.dsgd_fused import FusedDSGD, flat_layout, ordered_params
    S = len(models)
    params = [ordered_params(m) for m in models]
    total = flat_layout(params[0])[1]
    g = SiteGroup.local(S, total * 16 + (8 << 20))
    f = FusedDSGD(g, total, wire_dtype=wire, lr=lr, shadow_bf16=shadow)
    for s, m in enumerate(models):
        f.bind(m, site=s, params=params[s])
    return g, f, params


def test_device_dsgd_follows_the_file_mode_math_step_for_step():
    """fp32 wire, eager models on the GPU: 2 sites x 25 steps against 'mean of the sites' gradients + torch Adam'."""
    from dinunet_implementations_b200.models import MSANNet
    torch.manual_seed(0)
    S, B, steps = 2, 16, 25
    base = MSANNet(66, [64, 32], 2, fused=False).cuda()
    dev_models = [copy.deepcopy(base) for _ in range(S)]
    ref_models = [copy.deepcopy(base) for _ in range(S)]
    ref_opts = [torch.optim.Adam(m.parameters(), lr=1e-2) for m in ref_models]
    g, f, params = _bind_sites(dev_models, lr=1e-2)
    gen = torch.Generator(device="cuda").manual_seed(1)
    for step in range(steps):
        xs = [torch.rand(B, 66, device="cuda", generator=gen) for _ in range(S)]
        ys = [(x[:, :8].sum(1) > 4).long() for x in xs]
        f.zero_grad()
        dev_loss, ref_loss = [], []
        for s in range(S):
            l = torch.nn.functional.cross_entropy(dev_models[s](xs[s]), ys[s])
            l.backward()
            dev_loss.append(float(l))
        f.step()
        for s in range(S):
            ref_opts[s].zero_grad()
            l = torch.nn.functional.cross_entropy(ref_models[s](xs[s]), ys[s])
            l.backward()
            ref_loss.append(float(l))
        for ps in zip(*[list(m.parameters()) for m in ref_models]):      # the aggregator: unweighted mean over sites
            mean = torch.stack([p.grad for p in ps]).mean(0)
            for p in ps:
                p.grad = mean.clone()
        for o in ref_opts:
            o.step()
        for a, b in zip(dev_loss, ref_loss):
            assert abs(a - b) < 1e-3, (step, dev_loss, ref_loss)
    torch.cuda.synchronize()
    for p0, p1 in zip(params[0], params[1]):
        assert torch.equal(p0, p1)
    g.close()


@pytest.mark.parametrize("wire", [torch.float32, torch.bfloat16])
def test_four_simulated_sites_stay_identical_and_learn(wire):
    """S = 4 ICA-LSTM sites on the fused model kernels, different data per site, separable labels: the weights are
    bit-identical on all sites after every step and the pooled validation AUC ends above 0.8."""
    from dinunet_implementations_b200.models import ICALstm
    from dinunet_implementations_b200.metrics import AUCROCMetrics
    torch.manual_seed(2)
    S, B, T, C, W = 4, 8, 8, 8, 5
    base = ICALstm(input_size=32, hidden_size=64, num_comps=C, window_size=W, num_cls=2, fused=True).cuda()
    base.classifier[0].p = 0.0
    models = [copy.deepcopy(base) for _ in range(S)]
    g, f, params = _bind_sites(models, wire=wire, lr=3e-3, shadow=True)
    for m in models:
        m.train()
    gen = torch.Generator(device="cuda").manual_seed(3)

    def batch():
        y = torch.randint(0, 2, (B,), device="cuda", generator=gen)
        x = torch.randn(B, T, C, W, device="cuda", generator=gen)
        x[:, :, 0, :] += (y.float() * 2 - 1).view(B, 1, 1) * 0.8          # class shifts the first component
        return x, y
    for step in range(150):
        f.zero_grad()
        for s in range(S):
            x, y = batch()
            torch.nn.functional.cross_entropy(models[s](x)[0], y).backward()
        f.step()
        if step % 50 == 49:
            torch.cuda.synchronize()
            for s in range(1, S):
                for p0, ps in zip(params[0], params[s]):
                    assert torch.equal(p0, ps), (step, s)
    auc = AUCROCMetrics()
    models[0].eval()
    with torch.no_grad():
        for _ in range(20):
            x, y = batch()
            auc.add(torch.softmax(models[0](x)[0], 1)[:, 1], y)
    score = auc.auc if not callable(auc.auc) else auc.auc()
    assert score > 0.8, score
    g.close()
