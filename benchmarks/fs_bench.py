#!/usr/bin/env python
"""FreeSurfer MLP (the reference's DEFAULT task, local.py:12; 66 -> 256 -> 128 -> 64 -> 32 -> 2, batch 16, dSGD) under the
device engine vs the reconstructed baseline (the reference's own comps/fs/models.py eager + NCCL all_reduce + torch Adam).

    python benchmarks/fs_bench.py [--steps 300]            # 1 GPU
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/fs_bench.py

One JSON line per arm: ms/step (CUDA events, max over ranks), samples/s of the whole job, launches per step."""
import argparse
import importlib.util
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("DINUNET_QUIET", "1")

from dinunet_implementations_b200.parallel.engine import DeviceEngine, init_distributed_from_env  # noqa: E402


def timed(step, batches, steps, warmup, world, dev):
    for i in range(warmup):
        step(batches[i % len(batches)])
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(batches[(warmup + i) % len(batches)])
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--batch", type=int, default=16)
    a = ap.parse_args()
    rank, world, local_rank = init_distributed_from_env()
    dev = torch.device(f"cuda:{local_rank}")
    g = torch.Generator().manual_seed(100 + rank)
    batches = [{"inputs": torch.rand(a.batch, 66, generator=g).to(dev), "labels": torch.randint(0, 2, (a.batch,), generator=g).to(dev),
                "ix": torch.arange(a.batch).to(dev)} for _ in range(16)]
    from dinunet_implementations_b200.tasks.fs import FreeSurferTrainer
    cache = {"task_id": "FS-Classification", "input_size": 66, "hidden_sizes": [256, 128, 64, 32], "num_class": 2,
             "monitor_metric": "auc", "batch_size": a.batch, "learning_rate": 1e-3, "gpus": [local_rank], "seed": 1,
             "agg_engine": "dSGD", "fused": True}
    tr = FreeSurferTrainer(cache=cache, input={}, state={})
    tr.init_nn(init_optim=False)
    eng = DeviceEngine(tr, use_graph=True)
    ms = timed(lambda b: eng.train_step(b), batches, a.steps, a.warmup, world, dev)
    if rank == 0:
        print(json.dumps({"config": "FS-MLP 66-256-128-64-32-2 dSGD", "impl": "ours", "n_gpus": world, "batch_per_site": a.batch,
                          "ms_per_step": round(ms / a.steps, 4), "samples_per_sec": round(world * a.batch * a.steps / (ms / 1e3), 1),
                          "our_kernels_per_step": eng.kernels_per_step, "graph": eng._graph is not None}), flush=True)
    # the product path of parallel/federation.py: the split lives on the device, the batch gather is part of the graph,
    # the host only replays (no per-step copies, no per-step Python data handling)
    from dinunet_implementations_b200.data.device import DeviceTable, DeviceBatcher
    tbl = DeviceTable(torch.rand(4096, 66, generator=g).to(dev), torch.randint(0, 2, (4096,), generator=g).to(dev))
    eng.attach_batcher(DeviceBatcher(tbl, a.batch, seed=rank))
    ms = timed(lambda b: eng.train_step_resident(), batches, a.steps, a.warmup, world, dev)
    if rank == 0:
        print(json.dumps({"config": "FS-MLP 66-256-128-64-32-2 dSGD", "impl": "ours (device-resident batches)", "n_gpus": world,
                          "batch_per_site": a.batch, "ms_per_step": round(ms / a.steps, 4),
                          "samples_per_sec": round(world * a.batch * a.steps / (ms / 1e3), 1),
                          "our_kernels_per_step": eng.kernels_per_step, "graph": eng._rgraph is not None}), flush=True)
    eng.close()
    mod = None
    for base in (os.path.join(ROOT, "baseline", "_ref", "reference_src"), "/root/reference"):
        p = os.path.join(base, "comps/fs/models.py")
        if os.path.exists(p):
            spec = importlib.util.spec_from_file_location("_ref_fs_models_bench", p)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            break
    if mod is not None:
        torch.manual_seed(1)
        net = mod.MSANNet(66, [256, 128, 64, 32], 2).to(dev)
        opt = torch.optim.Adam(net.parameters(), lr=1e-3)
        params = list(net.parameters())

        def step(b):
            opt.zero_grad(set_to_none=False)
            out = torch.nn.functional.log_softmax(net(b["inputs"]), 1)
            torch.nn.functional.nll_loss(out, b["labels"]).backward()
            if world > 1:
                flat = torch.cat([q.grad.reshape(-1) for q in params])
                dist.all_reduce(flat)
                flat /= world
                off = 0
                for q in params:
                    q.grad.copy_(flat[off:off + q.numel()].view_as(q))
                    off += q.numel()
            opt.step()
        n = min(a.steps, 100)
        ms = timed(step, batches, n, 5, world, dev)
        if rank == 0:
            print(json.dumps({"config": "FS-MLP 66-256-128-64-32-2 dSGD", "impl": "recon-nccl (reference comps/fs/models.py eager + NCCL + torch Adam)",
                              "n_gpus": world, "batch_per_site": a.batch, "ms_per_step": round(ms / n, 4),
                              "samples_per_sec": round(world * a.batch * n / (ms / 1e3), 1)}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
