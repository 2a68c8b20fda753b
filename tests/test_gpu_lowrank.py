"""Device-mode powerSGD / rankDAD vs the fp32 oracle math (distrib/lowrank.py); S sites simulated
on one GPU with SiteGroup.local."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_orthogonalize_kernel_matches_oracle():
    from dinunet_implementations_b200.distrib import lowrank
    from dinunet_implementations_b200.parallel.lowrank_fused import orthogonalize_batch
    torch.manual_seed(0)
    mats = [torch.randn(n, r, device="cuda") for n, r in ((4096, 4), (300, 10), (64, 16), (1000, 1))]
    mats[1][:, 3] = mats[1][:, 0] * 2.0                       # rank-deficient column must vanish
    refs = [lowrank.orthogonalize(m.double()).float() for m in mats]
    orthogonalize_batch(mats)
    torch.cuda.synchronize()
    for m, r in zip(mats, refs):
        assert torch.allclose(m, r, atol=2e-4), (m - r).abs().max()
    assert mats[1][:, 3].abs().max() < 1e-4


def _nets(S, seed=0):
    torch.manual_seed(seed)
    base = torch.nn.Sequential(torch.nn.Linear(40, 64), torch.nn.ReLU(), torch.nn.BatchNorm1d(64), torch.nn.Linear(64, 3)).cuda()
    nets = []
    for _ in range(S):
        n = torch.nn.Sequential(torch.nn.Linear(40, 64), torch.nn.ReLU(), torch.nn.BatchNorm1d(64), torch.nn.Linear(64, 3)).cuda()
        n.load_state_dict(base.state_dict())
        nets.append(n)
    return nets


def test_device_powersgd_matches_oracle_step():
    from dinunet_implementations_b200.distrib import lowrank
    from dinunet_implementations_b200.parallel.comm import SiteGroup
    from dinunet_implementations_b200.parallel.lowrank_fused import DevicePowerSGD
    S, r = 3, 4
    nets = _nets(S)
    params = [[p for p in n.parameters()] for n in nets]
    g = SiteGroup.local(S, 8 << 20)
    eng = DevicePowerSGD(g, params, rank=r, lr=0.1, optimizer="sgd", seed=5)
    w0 = [p.detach().clone() for p in params[0]]
    Q0 = [q.clone() for q in eng.Q[0]]
    xs = [torch.randn(16, 40, device="cuda") for _ in range(S)]
    ys = [torch.randint(0, 3, (16,), device="cuda") for _ in range(S)]
    eng.zero_grad()
    for s in range(S):
        torch.nn.functional.cross_entropy(nets[s](xs[s]), ys[s]).backward()
    grads = [[p.grad.detach().clone() for p in params[s]] for s in range(S)]
    eng.step()
    torch.cuda.synchronize()
    k = 0
    for i, p in enumerate(params[0]):
        if p.dim() >= 2:
            Ms = [grads[s][i].view(p.shape[0], -1).double() for s in range(S)]      # fp64 oracle
            M_hat, Es, _ = lowrank.psgd_oracle_step(Ms, [Q0[k].double()] * S, r)
            M_hat, Es = M_hat.float(), [e.float() for e in Es]
            expect = w0[i] - 0.1 * M_hat.view_as(p)
            scale = (0.1 * M_hat).abs().max().item()
            assert torch.allclose(p.detach(), expect, atol=0.05 * scale + 1e-6), (i, (p - expect).abs().max(), scale)
            assert torch.allclose(eng.E[0][k], Es[0], atol=0.05 * Ms[0].abs().max().item())
            k += 1
        else:
            expect = w0[i] - 0.1 * torch.stack([grads[s][i] for s in range(S)]).mean(0)
            assert torch.allclose(p.detach(), expect, atol=1e-5)
    for s in range(1, S):                                    # every site ends bit-identical
        for a, b in zip(params[0], params[s]):
            assert torch.equal(a, b)
    g.close()


def test_device_rankdad_equals_dsgd_when_rank_covers_batch():
    from dinunet_implementations_b200.distrib.rankdad import _LinearTap
    from dinunet_implementations_b200.parallel.comm import SiteGroup
    from dinunet_implementations_b200.parallel.lowrank_fused import DeviceRankDAD
    S, B = 2, 6
    nets = _nets(S, seed=2)
    params = [[p for p in n.parameters()] for n in nets]
    taps = [[_LinearTap(m) for m in n.modules() if isinstance(m, torch.nn.Linear)] for n in nets]
    dense = [[p for m in n.modules() if isinstance(m, torch.nn.BatchNorm1d) for p in m.parameters()] for n in nets]
    g = SiteGroup.local(S, 8 << 20)
    eng = DeviceRankDAD(g, params, taps, dense, rank=16, num_pow_iters=40, tol=0.0, lr=0.1, optimizer="sgd")
    w0 = [p.detach().clone() for p in params[0]]
    xs = [torch.randn(B, 40, device="cuda") for _ in range(S)]
    ys = [torch.randint(0, 3, (B,), device="cuda") for _ in range(S)]
    eng.zero_grad()
    for s in range(S):
        nets[s].train()
        torch.nn.functional.cross_entropy(nets[s](xs[s]), ys[s]).backward()
    grads = [[p.grad.detach().clone() for p in params[s]] for s in range(S)]
    eng.step()
    torch.cuda.synchronize()
    for i, p in enumerate(params[0]):
        mean_g = torch.stack([grads[s][i] for s in range(S)]).mean(0)
        expect = w0[i] - 0.1 * mean_g
        scale = max((0.1 * mean_g).abs().max().item(), 1e-6)
        assert torch.allclose(p.detach(), expect, atol=0.05 * scale + 1e-6), (i, (p - expect).abs().max(), scale)
    for a, b in zip(params[0], params[1]):
        assert torch.equal(a, b)
    g.close()


@pytest.mark.parametrize("engine", ["powerSGD", "rankDAD"])
def test_device_engine_with_lowrank_aggregation_trains(engine):
    from dinunet_implementations_b200.parallel.engine import DeviceEngine, make_ica_trainer
    tr = make_ica_trainer(0, batch_size=8, hidden_size=64, input_size=32, num_components=8, window_size=5, temporal_size=40)
    tr.cache["agg_engine"] = engine
    tr.nn["net"].classifier[0].p = 0.0
    eng = DeviceEngine(tr)
    g = torch.Generator().manual_seed(0)
    batch = {"inputs": torch.randn(8, 8, 8, 5, generator=g), "labels": torch.randint(0, 2, (8,), generator=g)}
    losses = [float(eng.train_step(batch)["loss"].detach()) for _ in range(12)]
    assert all(l == l for l in losses) and losses[-1] < losses[0], losses


def test_rankdad_step_replays_as_a_cuda_graph():
    """The device rankDAD step (fused backward hands over (delta, activation) pairs, subspace iteration,
    factor exchange, reconstruction, optimiser) has no host syncs: it is captured once and replayed, and
    trains like the eager engine."""
    from dinunet_implementations_b200.parallel.engine import DeviceEngine, make_ica_trainer

    def eng(use_graph):
        tr = make_ica_trainer(0, batch_size=8, hidden_size=64, input_size=32, num_components=8, window_size=5,
                              temporal_size=40, seed=5)
        tr.cache["agg_engine"] = "rankDAD"
        tr.nn["net"].classifier[0].p = 0.0
        return DeviceEngine(tr, use_graph=use_graph, graph_warmup=2)
    g = torch.Generator().manual_seed(1)
    bs = [{"inputs": torch.randn(8, 8, 8, 5, generator=g), "labels": torch.randint(0, 2, (8,), generator=g)} for _ in range(10)]
    e1, e2 = eng(False), eng(True)
    l1 = [float(e1.train_step(b)["loss"].detach()) for b in bs]
    l2 = [float(e2.train_step(b)["loss"].detach()) for b in bs]
    assert e2._graph is not None, "rankDAD step must be graph-capturable"
    assert max(abs(a - b) for a, b in zip(l1, l2)) < 5e-3, (l1, l2)
    p1 = torch.cat([p.detach().flatten() for p in e1.params])
    p2 = torch.cat([p.detach().flatten() for p in e2.params])
    assert (p1 - p2).abs().mean() < 1e-3
    e1.close(); e2.close()


def test_fused_capture_pairs_reproduce_the_weight_gradients():
    """The (delta, activation) pairs handed over by the fused ICA-LSTM backward reconstruct the same
    gradients the fused weight-gradient GEMMs produce (delta^T . activation, bias = column sums)."""
    import copy
    from dinunet_implementations_b200.models import ICALstm
    torch.manual_seed(0)
    kw = dict(input_size=64, hidden_size=128, num_comps=12, window_size=6, num_cls=2)
    a = ICALstm(fused=True, **kw).cuda()
    b = ICALstm(fused=True, **kw).cuda()
    b.load_state_dict(copy.deepcopy(a.state_dict()))
    for m in (a, b):
        m.train(); m.classifier[0].p = 0.0
    b._dad_capture = True
    x = torch.randn(8, 9, 12, 6, device="cuda")
    y = torch.randint(0, 2, (8,), device="cuda")
    torch.nn.functional.cross_entropy(a(x)[0], y).backward()
    torch.nn.functional.cross_entropy(b(x)[0], y).backward()
    torch.cuda.synchronize()
    mods_a = dict(a.named_modules())
    assert set(b._dad_pairs) == set(b.dad_capture_keys())
    for key, (D, A) in b._dad_pairs.items():
        G = D.float().t() @ A.float()
        ref = mods_a[key].weight.grad
        s = max(ref.abs().max().item(), 1e-6)
        assert torch.allclose(G, ref, atol=0.05 * s), (key, (G - ref).abs().max().item(), s)
        bref = mods_a[key].bias.grad
        # a bias in front of BatchNorm has an exactly-zero gradient: what is left is fp32 cancellation
        # noise, so the scale is the magnitude of the summed terms, not of the (vanishing) result
        sb = max(bref.abs().max().item(), 1e-3 * D.float().abs().sum(0).max().item(), 1e-6)
        assert torch.allclose(D.float().sum(0), bref, atol=0.05 * sb), (key, "bias")
