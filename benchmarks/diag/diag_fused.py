import sys, copy, torch
sys.path.insert(0, '.')
sys.path.insert(0, "tests"); from test_gpu_icalstm_fused import _pair
cfgs = [(dict(input_size=64, hidden_size=128, num_comps=12, window_size=6, num_cls=3), 7, 11),
        (dict(input_size=32, hidden_size=64, num_comps=9, window_size=5, num_cls=2), 33, 5),
        (dict(input_size=256, hidden_size=384, num_comps=100, window_size=10, num_cls=2), 16, 98)]
for kw, B, S in cfgs:
    for mode in ("eval", "train"):
        ref, fused = _pair(kw)
        for m in (ref, fused):
            m.train(mode == "train"); m.classifier[0].p = 0.0
        x = torch.randn(B, S, kw["num_comps"], kw["window_size"], device="cuda")
        y = torch.randint(0, kw["num_cls"], (B,), device="cuda")
        lo_r, _ = ref(x); lo_f, _ = fused(x)
        torch.nn.functional.cross_entropy(lo_r, y).backward()
        torch.nn.functional.cross_entropy(lo_f, y).backward()
        torch.cuda.synchronize()
        print("==", kw["hidden_size"], B, S, mode, "logit err", (lo_f - lo_r).abs().max().item())
        for (n, pr), (_, pf) in zip(ref.named_parameters(), fused.named_parameters()):
            scale = pr.grad.abs().max().item(); err = (pf.grad - pr.grad).abs().max().item()
            cos = torch.nn.functional.cosine_similarity(pf.grad.flatten(), pr.grad.flatten(), dim=0).item()
            print(f"  {n:32s} scale {scale:.3e} err {err:.3e} rel {err/max(scale,1e-12):.3f} cos {cos:.5f}")
