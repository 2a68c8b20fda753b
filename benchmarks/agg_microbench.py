#!/usr/bin/env python
"""Aggregation micro-benchmark (BASELINE.json config 5): dSGD gradient aggregation + optimiser,
and the rankDAD factor exchange, 1 MB - 1 GB of fp32 gradients, ours vs the reconstructed NCCL path.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/agg_microbench.py

Arms
  ours        one launch of the fused kernel: barrier, pull peers' gradients over NVLink, mean,
              Adam, push parameters (csrc/dsgd.cu)
  recon-nccl  torch.distributed.all_reduce (NCCL) on the flat gradient + torch.optim.Adam step on
              the flat parameter (what the reference's aggregation becomes when run over NCCL)
  gather      rankDAD factor exchange for a 4096x4096 layer (r*(in+out) floats per site):
              ours (sum-of-disjoint-slots with the reduce-only kernel) vs NCCL all_gather

Times are CUDA-event durations on every rank, max over ranks, median of the timed iterations.
Between iterations a 256 MB buffer is overwritten so nothing is served from L2.
Roofline: the two-shot P2P algorithm moves (N-1)/N * (wire + fp32 parameter) bytes per GPU per direction over
NVLink.  The denominator is MEASURED in this run: every rank pulls 1 GiB from its right neighbour's arena with one
device-to-device cudaMemcpyAsync (all ranks at once, CUDA events, best of 5) -> "peer_copy_gbs"; the recipe's 770 GB/s
is printed beside it.  The NVLS arm (in-switch reduction, multicast stores) needs only 1/N of the bytes per GPU and
can therefore exceed 1.0 of the P2P roofline; its own bound (bytes/N each way) is reported as nvls_roofline_ms.
"""
import json
import os
import statistics
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("DINUNET_QUIET", "1")

from dinunet_implementations_b200.parallel.engine import init_distributed_from_env  # noqa: E402
from dinunet_implementations_b200.parallel.comm import SiteGroup  # noqa: E402
from dinunet_implementations_b200.parallel.dsgd_fused import FusedDSGD  # noqa: E402

NVLINK_GBS = 770.0
HBM_GBS = 6548.5


def timed(fn, iters, warm, flush, world, dev):
    ts = []
    for i in range(warm + iters):
        flush.add_(1.0)                      # evict L2
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if i >= warm:
            ts.append(float(t.item()))
    return statistics.median(ts)


def measure_peer_copy(world, rank, dev):
    """GB/s of a 1 GiB device-to-device copy out of the right neighbour's arena, all ranks concurrently."""
    import ctypes
    from dinunet_implementations_b200.ops import native
    nbytes = 1 << 30
    g = SiteGroup.from_torch_distributed(nbytes + (4 << 20), dev)
    g.alloc("probe", nbytes, torch.uint8)
    dst = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    src = g.ptrs("probe")[(rank + 1) % world]
    lib = native.lib()
    lib.dnb_copy_async.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
    best = 0.0
    for _ in range(6):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        native.check(lib.dnb_copy_async(dst.data_ptr(), src, nbytes, 3, torch.cuda.current_stream().cuda_stream), "copy")
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = max(best, nbytes / (float(t.item()) * 1e-3) / 1e9)
    dist.barrier()
    g.close(barrier=dist.barrier)
    del dst
    torch.cuda.empty_cache()
    return best


def main():
    rank, world, local_rank = init_distributed_from_env()
    dev = torch.device(f"cuda:{local_rank}")
    flush = torch.zeros(64 << 20, device=dev)
    peer_gbs = measure_peer_copy(world, rank, dev) if world > 1 else None
    if rank == 0 and peer_gbs:
        print(json.dumps({"bench": "peer_copy", "n_gpus": world, "peer_copy_gbs": round(peer_gbs, 1),
                          "method": "1 GiB cudaMemcpyAsync D2D from the right neighbour's peer-mapped arena, all ranks "
                                    "concurrently, CUDA events, max over ranks, best of 6", "recipe_gbs": NVLINK_GBS}), flush=True)
    link_gbs = peer_gbs or NVLINK_GBS
    sizes_mb = [1, 4, 16, 64, 256, 1024]
    max_mb = int(os.environ.get("AGG_MAX_MB", "1024"))
    results = []
    for mb in [s for s in sizes_mb if s <= max_mb]:
        n = mb * (1 << 20) // 4
        for wire, wname in ((torch.float32, "fp32"), (torch.bfloat16, "bf16")):
            arena = n * (4 + wire.itemsize) + (4 << 20)
            t_modes = {}
            for mode in ("two_shot", "nvls"):
                g = SiteGroup.from_torch_distributed(arena, dev) if world > 1 else SiteGroup.local(1, arena, dev)
                if mode == "nvls" and not (world > 1 and g.mc_base):
                    g.close(barrier=dist.barrier if world > 1 else None)
                    continue
                f = FusedDSGD(g, n, wire_dtype=wire, lr=1e-3, mode=mode, device_state=True)
                f.grads[rank if world > 1 else 0].normal_()
                t_modes[mode] = timed(f.step, 10, 3, flush, world, dev)
                g.close(barrier=dist.barrier if world > 1 else None)
                del f, g
            t_ours = min(t_modes.values())
            # reconstructed NCCL path
            p = torch.zeros(n, device=dev, requires_grad=True)
            grad = torch.randn(n, device=dev).to(wire)
            opt = torch.optim.Adam([p], lr=1e-3)

            def nccl_step():
                gg = grad.clone()
                if world > 1:
                    dist.all_reduce(gg)
                p.grad = gg.float().div_(world)
                opt.step()
            t_nccl = timed(nccl_step, 10, 3, flush, world, dev)
            del p, grad, opt
            torch.cuda.empty_cache()
            wire_bytes = n * wire.itemsize
            link_bytes = (world - 1) / world * (wire_bytes + n * 4) if world > 1 else 0.0   # pull grads + push fp32 params
            t_link = link_bytes / (link_gbs * 1e9) * 1e3
            t_nvls = ((wire_bytes + n * 4) / world) / (link_gbs * 1e9) * 1e3 if world > 1 else 0.0
            t_hbm = (n * (wire.itemsize * max(world, 1) / max(world, 1) + 4 * 3 + 4 * 3) / world) / (HBM_GBS * 1e9) * 1e3
            roof = max(t_link, t_hbm)
            results.append({"bench": "dsgd_reduce_adam", "n_gpus": world, "grad_mb_fp32": mb, "wire": wname,
                            "ours_ms": round(t_ours, 4), "ours_p2p_ms": round(t_modes["two_shot"], 4),
                            "ours_nvls_ms": round(t_modes["nvls"], 4) if "nvls" in t_modes else None,
                            "nvls_roofline_ms": round(max(t_nvls, t_hbm), 4),
                            "p2p_roofline_frac": round(roof / t_modes["two_shot"], 3),
                            "recon_nccl_ms": round(t_nccl, 4),
                            "speedup": round(t_nccl / t_ours, 2), "roofline_ms": round(roof, 4),
                            "roofline_frac": round(roof / t_ours, 3) if t_ours > 0 else None})
            if rank == 0:
                print(json.dumps(results[-1]), flush=True)
    # rankDAD factor exchange: r = 10, layer 4096 x 4096  -> 10 * 8192 floats per site
    if world > 1:
        L = 10 * 8192
        g = SiteGroup.from_torch_distributed(L * world * 8 + (4 << 20), dev)
        f = FusedDSGD(g, L * world, optimizer="none", mode="one_shot", device_state=True, name="gather")
        out = torch.zeros(f.n, device=dev)
        f.out = out
        f.grads[rank].zero_()
        f.grads[rank][rank * L:(rank + 1) * L].normal_()
        t_ours = timed(f.step, 20, 5, flush, world, dev)
        mine = torch.randn(L, device=dev)
        outs = [torch.empty(L, device=dev) for _ in range(world)]
        t_nccl = timed(lambda: dist.all_gather(outs, mine), 20, 5, flush, world, dev)
        if rank == 0:
            print(json.dumps({"bench": "rankdad_factor_allgather", "n_gpus": world, "floats_per_site": L,
                              "ours_ms": round(t_ours, 4), "nccl_all_gather_ms": round(t_nccl, 4),
                              "speedup": round(t_nccl / t_ours, 2)}), flush=True)
        g.close(barrier=dist.barrier)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
