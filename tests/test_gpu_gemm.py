"""tcgen05 + TMA GEMM vs a PyTorch fp32 reference of the same op (inputs rounded to bf16)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [
    (128, 128, 64), (128, 16, 64), (256, 64, 128), (1568, 256, 1000),   # encoder: K=1000 ragged
    (1568, 1536, 256), (100, 40, 72), (16, 66, 8), (768, 256, 1568), (4096, 16, 4096), (300, 1000, 1568),
]


def _mk(M, N, K, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    Kp = (K + 7) // 8 * 8
    a = torch.zeros(M, Kp, device="cuda", dtype=torch.bfloat16)
    b = torch.zeros(N, Kp, device="cuda", dtype=torch.bfloat16)
    a[:, :K] = (torch.randn(M, K, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
    b[:, :K] = (torch.randn(N, K, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
    return a[:, :K], b[:, :K]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("block_n", [0, 16, 32, 64, 128, 256])
def test_gemm_plain(M, N, K, block_n):
    from dinunet_implementations_b200.ops.gemm import gemm_bf16
    a, b = _mk(M, N, K)
    out = gemm_bf16(a, b, block_n=block_n)
    ref = a.float() @ b.float().t()
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    assert err <= 1e-2 * max(1.0, ref.abs().max().item()) * 0.05 + 1e-3, f"max err {err}"


@pytest.mark.parametrize("M,N,K", [(1568, 256, 1000), (200, 72, 136)])
def test_gemm_epilogues(M, N, K):
    from dinunet_implementations_b200.ops.gemm import gemm_bf16
    a, b = _mk(M, N, K, seed=1)
    bias_n = torch.randn(N, device="cuda")
    bias_m = torch.randn(M, device="cuda")
    ref = a.float() @ b.float().t()
    # bias + relu, bf16 out + transposed copy
    out_t = torch.empty(N, (M + 7) // 8 * 8, device="cuda", dtype=torch.bfloat16)[:, :M]
    out = gemm_bf16(a, b, out_dtype=torch.bfloat16, bias_n=bias_n, relu=True, out_t=out_t)
    r = torch.relu(ref + bias_n)
    assert torch.allclose(out.float(), r, atol=0.06, rtol=0.02)
    assert torch.equal(out_t.t().contiguous(), out)
    # row bias, alpha
    out = gemm_bf16(a, b, bias_m=bias_m, alpha=0.5)
    assert torch.allclose(out, 0.5 * ref + bias_m[:, None], atol=2e-2, rtol=1e-3)
    # relu-mask multiply + column sums (encoder backward epilogue)
    mask = (torch.randn(M, N, device="cuda") > 0).to(torch.bfloat16)
    colsum = torch.zeros(N, device="cuda")
    out = gemm_bf16(a, b, mask=mask, colsum=colsum)
    r = ref * mask.float()
    assert torch.allclose(out, r, atol=2e-2, rtol=1e-3)
    assert torch.allclose(colsum, r.sum(0), atol=0.5, rtol=1e-3)
    # split-K and accumulate
    out = gemm_bf16(a, b, split_k=4)
    assert torch.allclose(out, ref, atol=2e-2, rtol=1e-3)
    out2 = gemm_bf16(a, b, out=out.clone(), accumulate=True)
    assert torch.allclose(out2, 2 * ref, atol=4e-2, rtol=1e-3)
