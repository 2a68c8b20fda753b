"""K1 fused dSGD kernel vs a plain PyTorch fp32 restatement (mean of grads -> torch.optim step).
S sites are simulated on one GPU with SiteGroup.local (same kernel, no cross-GPU flags)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _reference(params0, grads, opt_kind, lr, steps_grads):
    p = params0.clone().requires_grad_(True)
    opt = (torch.optim.Adam([p], lr=lr) if opt_kind == "adam" else torch.optim.SGD([p], lr=lr, momentum=0.9))
    for gs in steps_grads:
        p.grad = torch.stack(gs).float().mean(0)
        opt.step()
    return p.detach()


@pytest.mark.parametrize("mode", ["two_shot", "one_shot"])
@pytest.mark.parametrize("wire", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("opt_kind", ["adam", "sgd"])
def test_fused_dsgd_matches_torch(mode, wire, opt_kind):
    from dinunet_implementations_b200.parallel.comm import SiteGroup
    from dinunet_implementations_b200.parallel.dsgd_fused import FusedDSGD
    S, n = 4, 100003
    torch.manual_seed(0)
    g = SiteGroup.local(S, arena_bytes=8 << 20)
    f = FusedDSGD(g, n, wire_dtype=wire, lr=1e-2, optimizer=opt_kind, momentum=0.9, shadow_bf16=True, mode=mode)
    p0 = torch.randn(n, device="cuda")
    for s in range(S):
        f.params[s][:n].copy_(p0)
    steps = []
    for t in range(3):
        gs = [torch.randn(n, device="cuda") * 0.1 for _ in range(S)]
        gs_wire = [x.to(wire) for x in gs]
        for s in range(S):
            f.grads[s][:n].copy_(gs_wire[s])
        steps.append([x.float() for x in gs_wire])
        f.step()
    ref = _reference(p0, None, opt_kind, 1e-2, steps)
    torch.cuda.synchronize()
    for s in range(S):
        assert torch.allclose(f.params[s][:n], ref, atol=2e-5, rtol=1e-5), (s, (f.params[s][:n] - ref).abs().max())
        assert torch.equal(f.params[s], f.params[0])                        # every site bit-identical
        assert torch.equal(f.shadow[s], f.params[s].to(torch.bfloat16))     # bf16 shadow tracks the master copy
    g.close()


def test_reduce_only_mode():
    from dinunet_implementations_b200.parallel.comm import SiteGroup
    from dinunet_implementations_b200.parallel.dsgd_fused import FusedDSGD
    S, n = 3, 4096
    g = SiteGroup.local(S, arena_bytes=1 << 20)
    f = FusedDSGD(g, n, optimizer="none", mode="one_shot")
    xs = [torch.randn(n, device="cuda") for _ in range(S)]
    for s in range(S):
        f.grads[s].copy_(xs[s])
    out = torch.empty(n, device="cuda")
    f.allreduce_mean(out)
    assert torch.allclose(out, torch.stack(xs).mean(0), atol=1e-6)
    g.close()


def test_bind_module_trains_like_torch_adam():
    from dinunet_implementations_b200.parallel.comm import SiteGroup
    from dinunet_implementations_b200.parallel.dsgd_fused import FusedDSGD
    torch.manual_seed(1)
    S = 2
    nets = [torch.nn.Sequential(torch.nn.Linear(20, 32), torch.nn.ReLU(), torch.nn.Linear(32, 3)).cuda() for _ in range(S)]
    for nnet in nets[1:]:
        nnet.load_state_dict(nets[0].state_dict())
    ref = torch.nn.Sequential(torch.nn.Linear(20, 32), torch.nn.ReLU(), torch.nn.Linear(32, 3)).cuda()
    ref.load_state_dict(nets[0].state_dict())
    ropt = torch.optim.Adam(ref.parameters(), lr=1e-2)
    from dinunet_implementations_b200.parallel.dsgd_fused import flat_layout, ordered_params
    numel = flat_layout(ordered_params(ref))[1]
    g = SiteGroup.local(S, arena_bytes=1 << 20)
    f = FusedDSGD(g, numel, lr=1e-2)
    for s in range(S):
        f.bind(nets[s], site=s)
    for t in range(4):
        xs = [torch.randn(16, 20, device="cuda") for _ in range(S)]
        ys = [torch.randint(0, 3, (16,), device="cuda") for _ in range(S)]
        f.zero_grad()
        for s in range(S):
            torch.nn.functional.cross_entropy(nets[s](xs[s]), ys[s]).backward()
        f.step()
        ropt.zero_grad()
        # mean over sites of per-site mean losses == the dSGD update
        (sum(torch.nn.functional.cross_entropy(ref(xs[s]), ys[s]) for s in range(S)) / S).backward()
        ropt.step()
    for a, b in zip(nets[0].parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=1e-5), (a - b).abs().max()
    for a, b in zip(nets[0].parameters(), nets[1].parameters()):
        assert torch.equal(a, b)
    g.close()
