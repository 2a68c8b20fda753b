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
