"""FreeSurfer MLP: tcgen05 GEMM with fused BatchNorm+ReLU epilogue, and the fused backward, vs eager."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_gemm_batchnorm_epilogue_matches_torch():
    from dinunet_implementations_b200.ops.gemm import gemm_bf16
    torch.manual_seed(0)
    for (N_feat, B, K) in [(256, 16, 72), (128, 37, 256), (4096, 16, 4096), (64, 200, 128)]:
        W = (torch.randn(N_feat, K, device="cuda") * 0.2).to(torch.bfloat16)
        x = torch.randn(B, K, device="cuda").to(torch.bfloat16)
        gamma = torch.rand(N_feat, device="cuda") + 0.5
        beta = torch.randn(N_feat, device="cuda") * 0.1
        Bp = (B + 7) // 8 * 8
        xhat = torch.empty(N_feat, Bp, device="cuda")[:, :B]
        invstd = torch.empty(N_feat, device="cuda")
        out_feat = torch.empty(N_feat, Bp, device="cuda", dtype=torch.bfloat16)[:, :B]
        out_batch = torch.empty(B, N_feat, device="cuda", dtype=torch.bfloat16)
        gemm_bf16(W, x, out=out_feat, out_t=out_batch, relu=True, bn=dict(gamma=gamma, beta=beta, eps=1e-5, xhat=xhat, invstd=invstd))
        z = x.float() @ W.float().t()                                   # [B, N]
        ref = torch.relu(torch.nn.functional.batch_norm(z, None, None, gamma, beta, True, 0.0, 1e-5))
        torch.cuda.synchronize()
        assert torch.allclose(out_batch.float(), ref, atol=3e-2, rtol=2e-2), (N_feat, B, K, (out_batch.float() - ref).abs().max())
        assert torch.equal(out_feat.t().contiguous(), out_batch)
        zn = (z - z.mean(0)) / torch.sqrt(z.var(0, unbiased=False) + 1e-5)
        assert torch.allclose(xhat.t(), zn, atol=2e-3, rtol=1e-3)


@pytest.mark.parametrize("hidden,B", [([256, 128, 64, 32], 16), ([256, 128, 64, 32], 50), ([512, 512], 16)])
def test_msannet_fused_matches_eager(hidden, B):
    from dinunet_implementations_b200.models import MSANNet
    torch.manual_seed(1)
    ref = MSANNet(66, hidden, 2, fused=False).cuda()
    with torch.no_grad():
        for p in ref.parameters():
            if p.dim() >= 2:
                p.copy_(p.to(torch.bfloat16).float())
    fused = MSANNet(66, hidden, 2, fused=True).cuda()
    fused.load_state_dict(copy.deepcopy(ref.state_dict()))
    x = torch.rand(B, 66, device="cuda").to(torch.bfloat16).float()
    y = torch.randint(0, 2, (B,), device="cuda")
    o_r, o_f = ref(x), fused(x)
    torch.cuda.synchronize()
    assert torch.allclose(o_f, o_r, atol=5e-2, rtol=5e-2), (o_f - o_r).abs().max()
    torch.nn.functional.cross_entropy(o_r, y).backward()
    torch.nn.functional.cross_entropy(o_f, y).backward()
    torch.cuda.synchronize()
    top = max(p.grad.abs().max().item() for p in ref.parameters())
    for (n, pr), (_, pf) in zip(ref.named_parameters(), fused.named_parameters()):
        assert pf.grad is not None, n
        if pr.grad.abs().max().item() < 1e-4 * top:
            continue
        cos = torch.nn.functional.cosine_similarity(pf.grad.flatten(), pr.grad.flatten(), dim=0).item()
        rel = ((pf.grad - pr.grad).norm() / pr.grad.norm()).item()
        assert cos > 0.98 and rel < 0.2, (n, cos, rel)


@pytest.mark.parametrize("hidden,B,in_size,classes", [([256, 128, 64, 32], 16, 66, 2), ([256, 128, 64, 32], 48, 66, 2),
                                                      ([256, 128, 64, 32], 32, 66, 2), ([40, 24], 2, 17, 5),
                                                      ([37, 21, 9], 5, 13, 3), ([128] * 8, 33, 128, 16)])
def test_whole_network_kernels_match_the_reference_model_file(hidden, B, in_size, classes, ref_fs_models):
    """csrc/mlp_small.cu (one launch forward, one backward) against the REFERENCE's own comps/fs/models.py on CUDA in
    fp32; the weights are bf16-representable so the only difference is the summation order: tight bounds, every
    parameter gradient, the input gradient, accumulation over two backward passes, eval mode (batch statistics too)."""
    from dinunet_implementations_b200.models import MSANNet
    from dinunet_implementations_b200.ops import mlp_small, native
    torch.manual_seed(3)
    ref = ref_fs_models.MSANNet(in_size, hidden, classes).cuda()
    with torch.no_grad():
        for p in ref.parameters():
            if p.dim() >= 2:
                p.copy_(p.to(torch.bfloat16).float())
        for m in ref.modules():
            if isinstance(m, torch.nn.BatchNorm1d):
                m.weight.uniform_(0.5, 1.5)
                m.bias.normal_(0, 0.2)
    ours = MSANNet(in_size, hidden, classes, fused=True).cuda()
    ours.load_state_dict(copy.deepcopy(ref.state_dict()))
    x = torch.rand(B, in_size, device="cuda")
    y = torch.randint(0, classes, (B,), device="cuda")
    assert mlp_small.supported(ours, x)
    # as under the engine: the kernels read a persistent bf16 copy of the weights (no per-call cast launches)
    ours._bf16_shadow = {id(p): p.detach().to(torch.bfloat16) for p in ours.parameters() if p.dim() >= 2}
    n0 = native.launch_count()
    xr, xo = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    for rep in range(2):                                  # two backward passes: gradients accumulate like autograd's
        lr_ = torch.nn.functional.cross_entropy(ref(xr), y)
        lo_ = torch.nn.functional.cross_entropy(ours(xo), y)
        lr_.backward(); lo_.backward()
    torch.cuda.synchronize()
    assert native.launch_count() - n0 == 4, "one launch forward + one backward per pass"
    assert torch.allclose(ours(x), ref(x), atol=2e-4, rtol=2e-4)
    for (n, pr), (_, po) in zip(ref.named_parameters(), ours.named_parameters()):
        s = max(pr.grad.abs().max().item(), 1e-6)
        assert torch.allclose(po.grad, pr.grad, atol=2e-3 * s + 1e-7), (n, (po.grad - pr.grad).abs().max().item(), s)
    s = max(xr.grad.abs().max().item(), 1e-6)
    assert torch.allclose(xo.grad, xr.grad, atol=2e-3 * s + 1e-7)
    ref.eval(); ours.eval()                               # BatchNorm has no running statistics: eval uses the batch too
    with torch.no_grad():
        assert torch.allclose(ours(x), ref(x), atol=2e-4, rtol=2e-4)


def test_fs_engine_step_is_a_handful_of_launches():
    """FS default net under the device engine: gather-free step = forward, loss, backward, aggregation (+ metrics)."""
    from dinunet_implementations_b200.parallel.engine import DeviceEngine
    from dinunet_implementations_b200.tasks.fs import FreeSurferTrainer
    from dinunet_implementations_b200.ops import native
    cache = {"task_id": "FS-Classification", "input_size": 66, "hidden_sizes": [256, 128, 64, 32], "num_class": 2,
             "monitor_metric": "auc", "batch_size": 16, "learning_rate": 1e-2, "gpus": [0], "seed": 1, "agg_engine": "dSGD",
             "fused": True}
    tr = FreeSurferTrainer(cache=cache, input={}, state={})
    tr.init_nn(init_optim=False)
    eng = DeviceEngine(tr, use_graph=True, graph_warmup=2)
    g = torch.Generator().manual_seed(0)
    batch = {"inputs": torch.rand(16, 66, generator=g), "labels": torch.randint(0, 2, (16,), generator=g), "ix": torch.arange(16)}
    losses = [float(eng.train_step(batch)["loss"].detach()) for _ in range(30)]
    assert eng._graph is not None and eng.kernels_per_step <= 4, eng.kernels_per_step
    assert losses[-1] < losses[0] * 0.7, losses
    eng.close()
