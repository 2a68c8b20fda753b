"""``dnb_head_backward_noalloc`` — the head backward for callers that may not allocate device memory.

Some host programs cannot hand the library a scratch buffer and are not allowed to let it call
``cudaMalloc`` either: their device memory is partitioned up front by another runtime or pool, or
they run under a fixed memory budget where a hidden allocation per call (possible device
synchronisation, possible late failure) is ruled out.  For them the ``[B, N1]`` scratch that
``dnb_head_backward`` expects from its caller lives in a fixed ``__device__`` array inside the
library (limits: ``B <= 64``, ``N1 <= 512``).

The test drives both entry points through the C ABI (ctypes) on identical inputs and requires
identical results, for two calls in a row with different shapes — the second call finds the first
call's leftovers in the shared array and must not be affected by them — and checks that shapes
beyond the built-in limits are refused with an error instead of overrunning the array.

The file name sorts last on purpose: a run that stops at the first failure has then already
executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(B, N0, N1, N2, N3, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    t = dict(
        dlogits=r(B, N3), a0=r(B, N0), xhat=r(B, N1), invstd=torch.rand(N1, device="cuda", generator=g) + 0.5,
        a1=torch.relu(r(B, N1)), a2=torch.relu(r(B, N2)),
        W1=(r(N1, N0) * 0.05).to(torch.bfloat16), W2=(r(N2, N1) * 0.05).to(torch.bfloat16),
        W3=(r(N3, N2) * 0.1).to(torch.bfloat16), bn_w=torch.rand(N1, device="cuda", generator=g) + 0.5,
    )
    return t


def _run(entry, t, dims, dz1):
    from dinunet_implementations_b200.ops import native
    B, N0, N1, N2, N3 = dims
    out = dict(dW1=torch.zeros(N1, N0), db1=torch.zeros(N1), dbn_w=torch.zeros(N1), dbn_b=torch.zeros(N1),
               dW2=torch.zeros(N2, N1), db2=torch.zeros(N2), dW3=torch.zeros(N3, N2), db3=torch.zeros(N3),
               dx=torch.zeros(B, N0))
    out = {k: v.cuda() for k, v in out.items()}
    d = native.HeadBwdDesc()
    for k in ("dlogits", "a0", "xhat", "invstd", "a1", "a2", "W1", "W2", "W3", "bn_w"):
        setattr(d, k, t[k].data_ptr())
    d.drop = None
    for k, v in out.items():
        setattr(d, k, v.data_ptr())
    d.dz1 = dz1.data_ptr() if dz1 is not None else None
    d.training, d.accumulate = 1, 1
    d.B, d.N0, d.N1, d.N2, d.N3 = dims
    rc = entry(ctypes.byref(d), native.stream())
    torch.cuda.synchronize()
    return rc, out


def _torch_reference(t, dims):
    """The same backward in plain fp32 PyTorch (the weights are bf16 values, everything else fp32)."""
    B = dims[0]
    W1, W2, W3 = t["W1"].float(), t["W2"].float(), t["W3"].float()
    dl = t["dlogits"]
    dz2 = (dl @ W3) * (t["a2"] > 0)
    dy = (dz2 @ W2) * (t["a1"] > 0)
    dxh = dy * t["bn_w"]
    dz1 = t["invstd"] / B * (B * dxh - dxh.sum(0) - t["xhat"] * (dxh * t["xhat"]).sum(0))
    return dict(dW3=dl.t() @ t["a2"], db3=dl.sum(0), dW2=dz2.t() @ t["a1"], db2=dz2.sum(0),
                dbn_w=(dy * t["xhat"]).sum(0), dbn_b=dy.sum(0), dW1=dz1.t() @ t["a0"], db1=dz1.sum(0), dx=dz1 @ W1)


def test_noalloc_head_backward_matches_caller_scratch_entry_point():
    from dinunet_implementations_b200.ops import native
    lib = native.require_gpu()
    lib.dnb_last_error.restype = ctypes.c_char_p
    assert hasattr(lib, "dnb_head_backward_noalloc")
    # two calls in a row; the second is smaller, so the shared array still holds rows/columns
    # written by the first one
    for dims, seed in (((64, 384, 256, 64, 2), 1), ((16, 384, 256, 64, 2), 2), ((5, 64, 512, 32, 3), 3)):
        t = _inputs(*dims, seed)
        dz1 = torch.empty(dims[0], dims[2], device="cuda")
        rc_ref, ref = _run(lib.dnb_head_backward, t, dims, dz1)
        rc_new, new = _run(lib.dnb_head_backward_noalloc, t, dims, None)
        assert rc_ref == 0 and rc_new == 0, lib.dnb_last_error()
        ref32 = _torch_reference(t, dims)
        for k, v in ref32.items():                       # both entry points against plain fp32 PyTorch
            s = max(v.abs().max().item(), 1e-3)
            assert torch.allclose(new[k], v, atol=5e-3 * s + 1e-5), (dims, k, (new[k] - v).abs().max().item(), s)
        for k in ref:
            s = max(ref[k].abs().max().item(), 1e-3)
            assert torch.allclose(new[k], ref[k], atol=2e-3 * s + 1e-5), (dims, k, (new[k] - ref[k]).abs().max().item())


def test_noalloc_call_allocates_no_device_memory():
    """The point of the entry point: free device memory is the same before and after a call (the address lookup of
    the first call has already happened above or happens in the warm-up call here)."""
    from dinunet_implementations_b200.ops import native
    lib = native.require_gpu()
    dims = (16, 384, 256, 64, 2)
    t = _inputs(*dims, 9)
    outs = [_run(lib.dnb_head_backward_noalloc, t, dims, None)[1]]          # warm-up: resolves the symbol address
    from dinunet_implementations_b200.ops.native import HeadBwdDesc
    B, N0, N1, N2, N3 = dims
    out = {k: torch.zeros_like(v) for k, v in outs[0].items()}
    d = HeadBwdDesc()
    for k in ("dlogits", "a0", "xhat", "invstd", "a1", "a2", "W1", "W2", "W3", "bn_w"):
        setattr(d, k, t[k].data_ptr())
    for k, v in out.items():
        setattr(d, k, v.data_ptr())
    d.training, d.accumulate = 1, 1
    d.B, d.N0, d.N1, d.N2, d.N3 = dims
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert lib.dnb_head_backward_noalloc(ctypes.byref(d), native.stream()) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)
    for k in out:
        s = max(outs[0][k].abs().max().item(), 1e-3)
        mult = 1 if k == "dx" else 3          # accumulate=1: the parameter gradients of three calls add up, dx is overwritten
        assert torch.allclose(out[k], mult * outs[0][k], atol=6e-3 * s + 1e-5), k


def test_noalloc_rejects_shapes_beyond_the_builtin_scratch():
    from dinunet_implementations_b200.ops import native
    lib = native.require_gpu()
    lib.dnb_last_error.restype = ctypes.c_char_p
    dims = (8, 64, 1024, 32, 2)          # N1 = 1024 > 512
    t = _inputs(*dims, 4)
    rc, _ = _run(lib.dnb_head_backward_noalloc, t, dims, None)
    assert rc != 0
    assert b"exceeds the built-in scratch" in lib.dnb_last_error()
