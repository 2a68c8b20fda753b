"""Engine math vs fp64 restatements."""
import torch

from dinunet_implementations_b200.distrib import lowrank as lr


def test_orthogonalize_orthonormal_and_rank_deficient():
    torch.manual_seed(0)
    P = torch.randn(200, 6, dtype=torch.float64)
    Q = lr.orthogonalize(P)
    assert torch.allclose(Q.t() @ Q, torch.eye(6, dtype=torch.float64), atol=1e-10)
    # span preserved
    assert torch.allclose(Q @ (Q.t() @ P), P, atol=1e-9)
    P[:, 3] = P[:, 0] * 2  # rank deficient: that column must vanish, not explode
    Q = lr.orthogonalize(P)
    assert Q[:, 3].abs().max() < 1e-6 and torch.isfinite(Q).all()


def test_dad_exact_when_batch_le_rank():
    torch.manual_seed(1)
    D, A = torch.randn(8, 30, dtype=torch.float64), torch.randn(8, 20, dtype=torch.float64)
    d, a, used = lr.dad_compress(D, A, rank=10)
    assert used == 0
    assert torch.allclose(lr.dad_reconstruct(d, a), D.t() @ A, atol=1e-12)


def test_dad_exact_when_true_rank_le_r():
    torch.manual_seed(2)
    # 64 samples but the gradient has rank 4
    basis_d, basis_a = torch.randn(4, 30, dtype=torch.float64), torch.randn(4, 20, dtype=torch.float64)
    c = torch.randn(64, 4, dtype=torch.float64)
    D, A = c @ basis_d, torch.linalg.qr(torch.randn(64, 64, dtype=torch.float64))[0][:, :4] @ basis_a
    G = D.t() @ A
    d, a, _ = lr.dad_compress(D, A, rank=10, num_pow_iters=3, tol=0.0)
    assert torch.allclose(lr.dad_reconstruct(d, a), G, atol=1e-8 * G.abs().max())


def test_dad_is_best_rank_r_after_enough_iters():
    torch.manual_seed(3)
    D, A = torch.randn(64, 40, dtype=torch.float64), torch.randn(64, 50, dtype=torch.float64)
    G = D.t() @ A
    d, a, _ = lr.dad_compress(D, A, rank=10, num_pow_iters=60, tol=0.0)
    err = (lr.dad_reconstruct(d, a) - G).norm()
    s = torch.linalg.svdvals(G)
    best = s[10:].pow(2).sum().sqrt()
    assert err <= best * 1.02


def test_dad_reduce_equals_sum_of_site_grads_when_low_rank():
    torch.manual_seed(4)
    sites = []
    for s in range(3):
        D, A = torch.randn(3, 16, dtype=torch.float64), torch.randn(3, 12, dtype=torch.float64)
        sites.append(lr.dad_compress(D, A, rank=10)[:2] + (D.t() @ A,))
    d, a, _ = lr.dad_reduce([x[0] for x in sites], [x[1] for x in sites], rank=10, num_pow_iters=4, tol=0.0)
    mean_G = sum(x[2] for x in sites) / 3
    assert torch.allclose(lr.dad_reconstruct(d, a), mean_G, atol=1e-8)


def test_powersgd_residual_plus_approx_is_input_and_exact_for_low_rank():
    torch.manual_seed(5)
    S, n, m, r = 4, 48, 32, 4
    Ms = [torch.randn(n, m, dtype=torch.float64) for _ in range(S)]
    Q0 = lr.psgd_init_q((n, m), r, seed=7, dtype=torch.float64)
    M_hat, Es, Q = lr.psgd_oracle_step(Ms, [Q0] * S, r)
    for M, E in zip(Ms, Es):
        assert torch.allclose(M_hat + E, M, atol=1e-12)
    # low-rank mean gradient: one step with warm Q recovers it exactly after two steps
    U, V = torch.randn(n, 2, dtype=torch.float64), torch.randn(m, 2, dtype=torch.float64)
    Ms = [U @ V.t() for _ in range(S)]
    M_hat, _, Q = lr.psgd_oracle_step(Ms, [Q0] * S, r)
    assert torch.allclose(M_hat, U @ V.t(), atol=1e-8)


def test_linear_taps_pair_each_call_with_its_own_delta():
    """Two micro-batches (local_iterations = 2) and a module called in a time loop: delta^T . activation over the
    stacked blocks must equal autograd's accumulated weight gradient (the bias rides along as the ones column)."""
    import torch
    from dinunet_implementations_b200.distrib.rankdad import _LinearTap
    torch.manual_seed(0)
    lin = torch.nn.Linear(5, 3)
    cell = torch.nn.Linear(4, 4)
    taps = [_LinearTap(lin), _LinearTap(cell)]
    lin.train(); cell.train()
    for mb in range(2):
        x = torch.randn(6, 5)
        h = torch.zeros(6, 4)
        for t in range(3):                              # the same module three times per forward
            h = torch.tanh(cell(h) + lin(x)[:, :1])
        (h.sum() / 2).backward()
    for tap, mod in zip(taps, (lin, cell)):
        D, A = tap.pop()
        G = D.t() @ A
        assert torch.allclose(G[:, :-1], mod.weight.grad, atol=1e-5), (G[:, :-1] - mod.weight.grad).abs().max()
        assert torch.allclose(G[:, -1], mod.bias.grad, atol=1e-5)
        assert tap.pop() == (None, None)


def test_make_taps_uses_hooks_on_a_cpu_site_even_for_fused_capable_models():
    """ICALstm defaults to fused=True, but on a CPU site its forward is the eager one: rankDAD must tap the nn.Linear
    modules there (the fused hand-over only exists on CUDA)."""
    import torch
    from dinunet_implementations_b200.distrib.rankdad import make_taps, _LinearTap
    from dinunet_implementations_b200.models import ICALstm
    torch.manual_seed(0)
    net = ICALstm(input_size=8, hidden_size=8, num_comps=3, window_size=4, num_cls=2, fused=True)
    taps, covered = make_taps([net])
    assert taps and all(isinstance(t, _LinearTap) for t in taps)
    net.train()
    x = torch.randn(4, 5, 3, 4)
    torch.nn.functional.cross_entropy(net(x)[0], torch.tensor([0, 1, 0, 1])).backward()
    for t in taps:
        D, A = t.pop()
        G = D.t() @ A
        assert torch.allclose(G[:, :-1], t.module.weight.grad, atol=1e-5)


def test_hi_lo_stacking_carries_fp32_factors_through_bf16_operands():
    """The recompression kernel (csrc/dad.cu, mode 1) stacks every received factor three times — D side as bf16
    hi / hi / lo parts, A side as hi / lo / hi — so that a product of bf16 operands equals the fp32 product of the
    factors up to the lo*lo term.  Same arithmetic on the CPU: relative error ~1e-5 instead of ~4e-3 for plain bf16."""
    import torch
    torch.manual_seed(0)
    S, r, n_out, n_in = 4, 16, 96, 80
    D = torch.randn(S * r, n_out)
    A = torch.randn(S * r, n_in)
    G = D.double().t() @ A.double()

    def split(x):
        hi = x.to(torch.bfloat16).float()
        lo = (x - hi).to(torch.bfloat16).float()
        return hi, lo
    Dh, Dl = split(D)
    Ah, Al = split(A)
    plain = Dh.double().t() @ Ah.double()
    stacked = torch.cat([Dh, Dh, Dl]).double().t() @ torch.cat([Ah, Al, Ah]).double()
    e_plain = ((plain - G).norm() / G.norm()).item()
    e_stack = ((stacked - G).norm() / G.norm()).item()
    assert e_plain > 1e-3 and e_stack < 5e-5, (e_plain, e_stack)
