"""Run under torchrun on >= 2 GPUs: the fused dSGD kernel over real CUDA-IPC peer memory.

Checks (1) raw kernel: every rank ends with the parameters torch.optim.Adam produces from the
mean of all ranks' gradients, bit-identical on all ranks, for both kernel modes and wire dtypes;
(2) DeviceEngine: a few ICA-LSTM training steps with different data per site leave identical
weights on every site, in eager and in CUDA-graph stepping."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ.setdefault("DINUNET_QUIET", "1")

from dinunet_implementations_b200.parallel.engine import init_distributed_from_env, DeviceEngine, make_ica_trainer  # noqa: E402
from dinunet_implementations_b200.parallel.comm import SiteGroup  # noqa: E402
from dinunet_implementations_b200.parallel.dsgd_fused import FusedDSGD  # noqa: E402


def gather(t):
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t.contiguous())
    return out


def main():
    rank, world, local_rank = init_distributed_from_env()
    dev = torch.device(f"cuda:{local_rank}")
    n = 300007
    probe = SiteGroup.from_torch_distributed(1 << 20, dev)
    has_mc = bool(probe.mc_base)
    dist.barrier()
    probe.close(barrier=dist.barrier)
    if rank == 0:
        print(f"arena backend: {probe.backend}, NVLS multicast mapping: {has_mc}", flush=True)
    for mode in ("two_shot", "one_shot") + (("nvls",) if has_mc else ()):
        for wire in (torch.float32, torch.bfloat16):
            g = SiteGroup.from_torch_distributed(16 << 20, dev)
            f = FusedDSGD(g, n, wire_dtype=wire, lr=1e-2, mode=mode, shadow_bf16=True)
            torch.manual_seed(7)
            p0 = torch.randn(n, device=dev)
            f.params[rank][:n].copy_(p0)
            ref = p0.clone().requires_grad_(True)
            opt = torch.optim.Adam([ref], lr=1e-2)
            for step in range(3):
                torch.manual_seed(100 * step + rank)
                gr = (torch.randn(n, device=dev) * 0.1).to(wire)
                f.grads[rank][:n].copy_(gr)
                torch.cuda.synchronize()
                dist.barrier()
                f.step()
                torch.cuda.synchronize()
                ref.grad = torch.stack([x.float() for x in gather(gr)]).mean(0)
                opt.step()
            mine = f.params[rank][:n].clone()
            # the switch rounds a 16-bit reduction once more (fp32 accumulate, 16-bit result): looser bound there
            atol = 2e-3 if (mode == "nvls" and wire != torch.float32) else 2e-5
            assert torch.allclose(mine, ref.detach(), atol=atol, rtol=1e-5), (mode, wire, (mine - ref).abs().max())
            for other in gather(mine):
                assert torch.equal(other, mine), (mode, wire, "ranks diverged")
            assert torch.equal(f.shadow[rank][:n], mine.to(torch.bfloat16))
            dist.barrier()
            g.close(barrier=dist.barrier)
    if rank == 0:
        print("raw fused dSGD over IPC: OK", flush=True)

    for use_graph in (False, True):
        tr = make_ica_trainer(local_rank, batch_size=8, hidden_size=64, input_size=32, num_components=8, window_size=5,
                              temporal_size=40, seed=3, precision_bits="16")
        tr.nn["net"].classifier[0].p = 0.0
        eng = DeviceEngine(tr, use_graph=use_graph, graph_warmup=2)
        gen = torch.Generator().manual_seed(1000 + rank)          # different data on every site
        losses = []
        for _ in range(6):
            batch = {"inputs": torch.randn(8, 8, 8, 5, generator=gen), "labels": torch.randint(0, 2, (8,), generator=gen)}
            losses.append(float(eng.train_step(batch)["loss"].detach()))
        torch.cuda.synchronize()
        flat = torch.cat([p.detach().flatten() for p in eng.params])
        for other in gather(flat):
            assert torch.equal(other, flat), ("engine", use_graph, "sites diverged")
        assert all(l == l for l in losses)
        dist.barrier()
        eng.group.close()
    # a model large enough for the two-shot (shard + push) mode with the 16-bit wire: the aggregation kernel
    # casts its own gradients into the wire buffer, block-aligned with the peers' reads
    tr = make_ica_trainer(local_rank, batch_size=8, hidden_size=256, input_size=128, num_components=32, window_size=10,
                          temporal_size=80, seed=4, precision_bits="16")
    tr.nn["net"].classifier[0].p = 0.0
    eng = DeviceEngine(tr, use_graph=True, graph_warmup=2)
    assert eng.fused.mode in (0, 2) and eng.fused.inkernel_pack, (eng.fused.mode, eng.fused.inkernel_pack)
    gen = torch.Generator().manual_seed(3000 + rank)
    first = None
    for i in range(8):
        batch = {"inputs": torch.randn(8, 8, 32, 10, generator=gen), "labels": torch.randint(0, 2, (8,), generator=gen)}
        l = float(eng.train_step(batch)["loss"].detach())
        first = l if first is None else first
    torch.cuda.synchronize()
    flat = torch.cat([p.detach().flatten() for p in eng.params])
    assert torch.isfinite(flat).all()
    for other in gather(flat):
        assert torch.equal(other, flat), "two-shot + in-kernel pack: sites diverged"
    # the staging gradients are left at zero by the kernel (no zero-fill launch in the step)
    assert float(eng.fused._grad_staging[eng.site].abs().max()) == 0.0
    dist.barrier()
    eng.close()
    if rank == 0:
        print("DeviceEngine multi-site: OK", flush=True)

    # race-detection mode: gradient wire buffers are NaN outside their read window; a peer that reads
    # too early or too late would spread NaN into everyone's parameters
    os.environ["DINUNET_DEBUG_POISON"] = "1"
    tr = make_ica_trainer(local_rank, batch_size=8, hidden_size=64, input_size=32, num_components=8, window_size=5,
                          temporal_size=40, seed=3, precision_bits="16")
    eng = DeviceEngine(tr, use_graph=True, graph_warmup=2)
    assert eng.fused.debug_poison
    gen = torch.Generator().manual_seed(2000 + rank)
    for _ in range(40):
        batch = {"inputs": torch.randn(8, 8, 8, 5, generator=gen), "labels": torch.randint(0, 2, (8,), generator=gen)}
        eng.train_step(batch)
    torch.cuda.synchronize()
    flat = torch.cat([p.detach().flatten() for p in eng.params])
    assert torch.isfinite(flat).all(), "a poisoned gradient buffer was read outside its window"
    assert torch.isnan(eng.fused.grads[rank]).all()
    for other in gather(flat):
        assert torch.equal(other, flat)
    prof = eng.profile_steps([{"inputs": torch.randn(8, 8, 8, 5, generator=gen),
                               "labels": torch.randint(0, 2, (8,), generator=gen)} for _ in range(8)], warmup=3)
    assert set(prof) >= {"forward", "backward", "aggregate+update", "step"} and prof["step"]["calls"] == 5
    dist.barrier()
    eng.close()
    os.environ.pop("DINUNET_DEBUG_POISON")
    if rank == 0:
        print("poisoned-buffer race check + phase profile: OK", {k: round(v["ms_mean"], 3) for k, v in prof.items()}, flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
