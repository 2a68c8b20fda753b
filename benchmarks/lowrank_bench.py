#!/usr/bin/env python
"""BASELINE.json configs 3 and 4 under torchrun (one site per GPU), synthetic data:

  --config ica-rankdad   ICA-LSTM (input 256, hidden 384, S 98), rankDAD rank 10, batch 16 per site
  --config mlp-powersgd  FreeSurfer MLP widened to 66 -> 4096 x 6 -> 2, powerSGD rank 4, batch 16 per site

Arms: `ours` = DeviceEngine (fused sm_100a model kernels, device-mode low-rank engine, peer-memory
exchange); `recon-nccl` = the reference's own eager models.py + the same engine math in torch ops with
NCCL collectives (all_gather of factors / all_reduce of P and Q) + torch.optim.Adam.
Prints one JSON line per arm: samples/s (whole job), rounds/s, ms/step — CUDA events, max over ranks."""
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

from dinunet_implementations_b200.distrib import lowrank  # noqa: E402
from dinunet_implementations_b200.distrib.rankdad import _LinearTap  # noqa: E402
from dinunet_implementations_b200.parallel.engine import DeviceEngine, init_distributed_from_env, make_ica_trainer  # noqa: E402


def ref_module(rel):
    for base in (os.path.join(ROOT, "baseline", "_ref", "reference_src"), "/root/reference"):
        p = os.path.join(base, rel)
        if os.path.exists(p):
            spec = importlib.util.spec_from_file_location("_ref_" + rel.replace("/", "_"), p)
            m = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(m)
            return m
    return None


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


def make_fs_trainer(device_index, hidden, batch, engine, rank):
    from dinunet_implementations_b200.tasks.fs import FreeSurferTrainer
    cache = {"task_id": "FS-Classification", "input_size": 66, "hidden_sizes": hidden, "num_class": 2,
             "monitor_metric": "auc", "batch_size": batch, "learning_rate": 1e-3, "gpus": [device_index], "seed": 1,
             "agg_engine": engine, "powersgd_rank": rank, "fused": True}
    tr = FreeSurferTrainer(cache=cache, input={}, state={})
    tr.init_nn(init_optim=False)
    return tr


def recon_rankdad(net, world, r=10, iters=5, tol=1e-3):
    taps = [_LinearTap(m) for m in net.modules() if isinstance(m, torch.nn.Linear)]
    covered = {id(p) for t in taps for p in (t.module.weight, t.module.bias) if p is not None}
    dense = [p for p in net.parameters() if id(p) not in covered]
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)

    def step(batch):
        net.train()
        opt.zero_grad(set_to_none=False)
        out = net(batch["inputs"])
        out = out[0] if isinstance(out, tuple) else out
        torch.nn.functional.cross_entropy(out, batch["labels"]).backward()
        for k, t in enumerate(taps):
            D, A = t.pop()
            d, a, _ = lowrank.dad_compress(D, A, r, iters, tol, k)
            ds = [torch.empty_like(d) for _ in range(world)]
            as_ = [torch.empty_like(a) for _ in range(world)]
            if world > 1:
                dist.all_gather(ds, d.contiguous())
                dist.all_gather(as_, a.contiguous())
            else:
                ds, as_ = [d], [a]
            dr, ar, _ = lowrank.dad_reduce(ds, as_, r, iters, tol, 101 * k)
            G = dr.t() @ ar
            if t.module.bias is not None:
                t.module.bias.grad = G[:, -1].contiguous()
                G = G[:, :-1]
            t.module.weight.grad = G.contiguous()
        if world > 1:
            for p in dense:
                dist.all_reduce(p.grad)
                p.grad.div_(world)
        opt.step()
    return step


def recon_powersgd(net, world, rank=4):
    ps = list(net.parameters())
    mats = [p for p in ps if p.dim() >= 2]
    vecs = [p for p in ps if p.dim() < 2]
    Q = [lowrank.psgd_init_q((p.shape[0], p.numel() // p.shape[0]), min(rank, p.shape[0]), 17 * i, p.device)
         for i, p in enumerate(mats)]
    E = [torch.zeros(p.shape[0], p.numel() // p.shape[0], device=p.device) for p in mats]
    opt = torch.optim.Adam(ps, lr=1e-3)

    def step(batch):
        net.train()
        opt.zero_grad(set_to_none=False)
        out = net(batch["inputs"])
        torch.nn.functional.cross_entropy(out, batch["labels"]).backward()
        for i, p in enumerate(mats):
            M = p.grad.view(p.shape[0], -1) + E[i]
            P = M @ Q[i]
            if world > 1:
                dist.all_reduce(P); P /= world
            P = lowrank.orthogonalize(P)
            Qn = M.t() @ P
            if world > 1:
                dist.all_reduce(Qn); Qn /= world
            Mh = P @ Qn.t()
            E[i] = M - Mh
            p.grad = Mh.view_as(p).contiguous()
            Q[i] = Qn
        if world > 1:
            for v in vecs:
                dist.all_reduce(v.grad); v.grad.div_(world)
        opt.step()
    return step


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=["ica-rankdad", "mlp-powersgd"])
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--arms", default="ours,recon-nccl")
    a = ap.parse_args()
    rank, world, local_rank = init_distributed_from_env()
    dev = torch.device(f"cuda:{local_rank}")
    g = torch.Generator().manual_seed(100 + rank)
    if a.config == "ica-rankdad":
        batches = [{"inputs": torch.randn(a.batch, 98, 100, 10, generator=g).to(dev),
                    "labels": torch.randint(0, 2, (a.batch,), generator=g).to(dev)} for _ in range(8)]
        label = "ICA-LSTM rankDAD r=10"
    else:
        batches = [{"inputs": torch.rand(a.batch, 66, generator=g).to(dev),
                    "labels": torch.randint(0, 2, (a.batch,), generator=g).to(dev),
                    "ix": torch.arange(a.batch).to(dev)} for _ in range(8)]
        label = "FS-MLP 66-4096x6-2 powerSGD r=4"
    for arm in a.arms.split(","):
        if arm == "ours":
            if a.config == "ica-rankdad":
                tr = make_ica_trainer(local_rank, batch_size=a.batch)
                tr.cache["agg_engine"] = "rankDAD"
            else:
                tr = make_fs_trainer(local_rank, [4096] * 6, a.batch, "powerSGD", 4)
            eng = DeviceEngine(tr, use_graph=True)   # captured when the engine allows it (rankDAD with fused capture); else eager
            step = lambda b: eng.train_step(b)
        else:
            torch.manual_seed(1)
            if a.config == "ica-rankdad":
                mod = ref_module("comps/icalstm/models.py")
                if mod is None:
                    continue
                net = mod.ICALstm(input_size=256, hidden_size=384, num_comps=100, window_size=10, num_cls=2).to(dev)
                step = recon_rankdad(net, world)
            else:
                mod = ref_module("comps/fs/models.py")
                if mod is None:
                    continue
                net = mod.MSANNet(66, [4096] * 6, 2).to(dev)
                step = recon_powersgd(net, world)
        ms = timed(step, batches, a.steps, a.warmup, world, dev)
        if rank == 0:
            print(json.dumps({"config": label, "impl": arm, "n_gpus": world, "batch_per_site": a.batch,
                              "ms_per_step": round(ms / a.steps, 3),
                              "samples_per_sec": round(world * a.batch * a.steps / (ms / 1e3), 1),
                              "rounds_per_sec": round(a.steps / (ms / 1e3), 1)}), flush=True)
        if arm == "ours":
            eng.group.close()
        torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
