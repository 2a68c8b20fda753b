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
    dense = [[p for p in n.parameters() if p.dim() < 2] for n in nets]       # norm layers and biases travel dense
    g = SiteGroup.local(S, 8 << 20)
    eng = DeviceRankDAD(g, params, taps, dense, rank=16, num_pow_iters=8, tol=0.0, lr=0.1, optimizer="sgd")
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
        assert torch.allclose(p.detach(), expect, atol=0.02 * scale + 1e-6), (i, (p - expect).abs().max(), scale)
    for a, b in zip(params[0], params[1]):
        assert torch.equal(a, b)
    g.close()


def _bf16(x):
    return x.to(torch.bfloat16).float()


@pytest.mark.parametrize("N,n_in,n_out,S", [(1568, 256, 768, 2), (1568, 1000, 256, 3), (16, 384, 256, 4), (300, 66, 40, 8)])
def test_dad_kernels_match_fp64_oracle(N, n_in, n_out, S):
    """K2a + K2b (csrc/dad.cu) against the fp64 restatement of the same algorithm (distrib/lowrank.py): per-site
    rank-10 subspace iteration, stacking, re-compression, reconstruction.  The oracle gets the same bf16-rounded
    (delta, activation) inputs and the same seeded start matrices; what remains is the kernels' hi/lo bf16 iterates."""
    from dinunet_implementations_b200.distrib import lowrank
    from dinunet_implementations_b200.distrib.rankdad import _LinearTap
    from dinunet_implementations_b200.parallel.comm import SiteGroup
    from dinunet_implementations_b200.parallel.lowrank_fused import DeviceRankDAD
    torch.manual_seed(N + n_in)
    r, iters, seed = 10, 5, 3
    lins = [torch.nn.Linear(n_in, n_out, bias=False).cuda() for _ in range(S)]
    for l in lins[1:]:
        l.load_state_dict(lins[0].state_dict())
    params = [[l.weight] for l in lins]
    taps = [[_LinearTap(l)] for l in lins]
    g = SiteGroup.local(S, 64 << 20)
    eng = DeviceRankDAD(g, params, taps, [[] for _ in lins], rank=r, num_pow_iters=iters, tol=0.0, lr=1.0, optimizer="sgd", seed=seed)
    w0 = lins[0].weight.detach().clone()
    # activations with a decaying spectrum (a few dominant directions, like real gradients) so that rank 10 means something
    basis = torch.randn(24, n_in, device="cuda")
    As = [_bf16((torch.randn(N, 24, device="cuda") * torch.logspace(0, -2, 24, device="cuda")) @ basis) for _ in range(S)]
    Ds = [_bf16(torch.randn(N, n_out, device="cuda") * 0.1) for _ in range(S)]
    eng.zero_grad()
    for s in range(S):
        lins[s].train()
        (lins[s](As[s]) * Ds[s]).sum().backward()          # delta = Ds[s] exactly
    eng.step()
    torch.cuda.synchronize()
    G_dev = (w0 - lins[0].weight.detach()).double()
    ds, as_ = [], []
    for s in range(S):
        d, a, _ = lowrank.dad_compress(Ds[s].double(), As[s].double(), r, iters, 0.0, seed + 0)
        ds.append(d); as_.append(a)
    d, a, _ = lowrank.dad_reduce(ds, as_, r, iters, 0.0, seed + 0, average=True)
    G_ref = lowrank.dad_reconstruct(d, a)
    G_true = sum(D.double().t() @ A.double() for D, A in zip(Ds, As)) / S
    err = ((G_dev - G_ref).norm() / G_ref.norm()).item()
    base = ((G_ref - G_true).norm() / G_true.norm()).item()       # what rank 10 itself loses
    assert err < 0.02, (err, base)
    assert ((G_dev - G_true).norm() / G_true.norm()).item() < base + 0.02
    for s in range(1, S):
        assert torch.equal(lins[0].weight, lins[s].weight)        # every site reconstructs the same bits
    g.close()


def test_dad_kernels_exact_for_low_rank_gradients():
    """rank(G) <= r on every site and in the sum: the compressed exchange loses nothing."""
    from dinunet_implementations_b200.distrib.rankdad import _LinearTap
    from dinunet_implementations_b200.parallel.comm import SiteGroup
    from dinunet_implementations_b200.parallel.lowrank_fused import DeviceRankDAD
    torch.manual_seed(7)
    S, N, n_in, n_out, r = 2, 512, 192, 320, 10
    lins = [torch.nn.Linear(n_in, n_out, bias=False).cuda() for _ in range(S)]
    for l in lins[1:]:
        l.load_state_dict(lins[0].state_dict())
    taps = [[_LinearTap(l)] for l in lins]
    g = SiteGroup.local(S, 32 << 20)
    eng = DeviceRankDAD(g, [[l.weight] for l in lins], taps, [[] for _ in lins], rank=r, num_pow_iters=5, lr=1.0, optimizer="sgd")
    w0 = lins[0].weight.detach().clone()
    # every site's activations live in the same 5-dimensional subspace: rank(sum_s D_s^T A_s) <= 5
    basis = _bf16(torch.randn(5, n_in, device="cuda"))
    As = [_bf16(torch.randint(-2, 3, (N, 5), device="cuda").float()) @ basis for _ in range(S)]
    As = [_bf16(a) for a in As]
    Ds = [_bf16(torch.randn(N, n_out, device="cuda")) for _ in range(S)]
    eng.zero_grad()
    for s in range(S):
        lins[s].train()
        (lins[s](As[s]) * Ds[s]).sum().backward()
    eng.step()
    torch.cuda.synchronize()
    G_dev = (w0 - lins[0].weight.detach()).double()
    G_true = sum(D.double().t() @ A.double() for D, A in zip(Ds, As)) / S
    rel = ((G_dev - G_true).norm() / G_true.norm()).item()
    assert rel < 5e-3, rel
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
    for key, entries in b._dad_pairs.items():
        assert len(entries) == 1
        D, A = entries[0]["D"], entries[0]["A"]
        if "DT" in entries[0]:      # both orientations are handed over for the TMA-fed rankDAD kernels
            assert torch.equal(entries[0]["DT"].float(), D.float().t()) and torch.equal(entries[0]["AT"].float(), A.float().t())
        G = D.float().t() @ A.float()
        ref = mods_a[key].weight.grad
        s = max(ref.abs().max().item(), 1e-6)
        assert torch.allclose(G, ref, atol=0.05 * s), (key, (G - ref).abs().max().item(), s)
        bref = mods_a[key].bias.grad
        # a bias in front of BatchNorm has an exactly-zero gradient: what is left is fp32 cancellation
        # noise, so the scale is the magnitude of the summed terms, not of the (vanishing) result
        sb = max(bref.abs().max().item(), 1e-3 * D.float().abs().sum(0).max().item(), 1e-6)
        assert torch.allclose(D.float().sum(0), bref, atol=0.05 * sb), (key, "bias")


def test_file_mode_rankdad_learner_on_a_fused_model(tmp_path):
    """File-mode rankDAD on a CUDA site whose model runs the fused kernels (its nn.Linear modules are never called, so
    module hooks cannot fire): the learner must take the (delta, activation) pairs from the fused backward instead of
    shipping zero factors.  rank >= rows of the head pairs, so those layers reconstruct exactly."""
    from dinunet_implementations_b200.distrib import lowrank
    from dinunet_implementations_b200.distrib.rankdad import make_taps, CapturedTap
    from dinunet_implementations_b200.models import ICALstm
    torch.manual_seed(0)
    kw = dict(input_size=32, hidden_size=64, num_comps=8, window_size=5, num_cls=2)
    net = ICALstm(fused=True, **kw).cuda()
    twin = ICALstm(fused=True, **kw).cuda()               # same weights, ordinary (non-capturing) fused backward
    twin.load_state_dict(net.state_dict())
    for m in (net, twin):
        m.train(); m.classifier[0].p = 0.0
    taps, covered = make_taps([net])
    assert all(isinstance(t, CapturedTap) for t in taps) and net.fused
    x = torch.randn(6, 7, 8, 5, device="cuda")
    y = torch.randint(0, 2, (6,), device="cuda")
    torch.nn.functional.cross_entropy(net(x)[0], y).backward()
    torch.nn.functional.cross_entropy(twin(x)[0], y).backward()
    torch.cuda.synchronize()
    ref = dict(twin.named_modules())
    names = [n for n, m in net.named_modules() if isinstance(m, torch.nn.Linear)]
    for name, tap in zip(names, taps):
        D, A = tap.pop()                               # fp32, ones column appended for the bias
        assert D is not None, "a fused model must hand its pairs over"
        wg = ref[name].weight.grad
        G = D.t() @ A
        s = max(wg.abs().max().item(), 1e-6)
        assert torch.allclose(G[:, :-1], wg, atol=0.05 * s), (name, (G[:, :-1] - wg).abs().max().item(), s)
        d, a, _ = lowrank.dad_compress(D, A, 10, 5, 1e-3, 0)
        if D.shape[0] <= 10:
            assert torch.allclose(lowrank.dad_reconstruct(d, a), G, atol=1e-4 * max(G.abs().max().item(), 1e-6) + 1e-7)
