#!/usr/bin/env python
"""Micro-benchmark of the rankDAD device kernels (csrc/dad.cu) on the ICA-LSTM layer set of BASELINE config 3:
times the compress kernel (K2a, incl. the factor all-gather tail) and the recompress + reconstruct kernel (K2b)
separately with CUDA events, for several cluster sizes / iteration counts, S simulated sites on one GPU.

    python benchmarks/dad_kernel_bench.py [--sites 1] [--clusters 1,2,4,8] [--iters 5] [--batch 16]
"""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("DINUNET_QUIET", "1")

from dinunet_implementations_b200.ops import native  # noqa: E402
from dinunet_implementations_b200.parallel.comm import SiteGroup  # noqa: E402
from dinunet_implementations_b200.parallel.lowrank_fused import DeviceRankDAD  # noqa: E402


class FakeTap:
    def __init__(self, lin, entry):
        self.module, self.entry = lin, entry

    def pop_entries(self):
        return [self.entry]


def layer_set(batch, S=98):
    N = S * batch
    # (out, in, rows): encoder, i2h x2, h2h x2, classifier 1/4/6
    return [(256, 1000, N), (768, 256, N), (768, 256, N), (768, 192, N), (768, 192, N), (256, 384, batch),
            (64, 256, batch), (2, 64, batch)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sites", type=int, default=1)
    ap.add_argument("--clusters", default="1,2,4,8")
    ap.add_argument("--iters", default="5")
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--nsplit", type=int, default=3)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    shapes = layer_set(a.batch)

    def r8(n):
        return (n + 7) // 8 * 8

    for cl in [int(c) for c in a.clusters.split(",")]:
        for iters in [int(i) for i in a.iters.split(",")]:
            g = SiteGroup.local(a.sites, 256 << 20, dev)
            lins = [[torch.nn.Linear(i, o, bias=False).to(dev) for (o, i, _) in shapes] for _ in range(a.sites)]
            taps = []
            for s in range(a.sites):
                ts = []
                for lin, (o, i, n) in zip(lins[s], shapes):
                    D = torch.zeros(n, r8(o), device=dev, dtype=torch.bfloat16)[:, :o].normal_()
                    A = torch.zeros(n, r8(i), device=dev, dtype=torch.bfloat16)[:, :i].normal_()
                    DT = torch.zeros(o, r8(n), device=dev, dtype=torch.bfloat16)[:, :n].copy_(D.t())
                    AT = torch.zeros(i, r8(n), device=dev, dtype=torch.bfloat16)[:, :n].copy_(A.t())
                    ts.append(FakeTap(lin, {"D": D, "DT": DT, "A": A, "AT": AT}))
                taps.append(ts)
            eng = DeviceRankDAD(g, [[l.weight for l in ls] for ls in lins], taps, [[] for _ in lins], rank=10,
                                num_pow_iters=iters, lr=1e-3, optimizer="sgd", nsplit=a.nsplit, cluster=cl)
            # time the two kernels separately: monkey-patch the launcher
            lib = native.lib()
            times = {0: [], 1: []}
            orig = eng._launch_groups

            def timed(build, li, s, mode, orig=orig):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                orig(build, li, s, mode)
                e1.record()
                times[mode].append((e0, e1))
            eng._launch_groups = timed
            for rep in range(a.reps + 5):
                if rep == 5:
                    torch.cuda.synchronize()
                    times[0].clear(); times[1].clear()
                eng.zero_grad()
                eng.step()
            torch.cuda.synchronize()
            med = {m: sorted(e0.elapsed_time(e1) for e0, e1 in v)[len(v) // 2] * 1e3 for m, v in times.items()}
            # algorithmic numbers per site: flops of the four skinny products (N=32 wide incl. hi/lo) per iteration
            fl = sum(2 * 32 * n * (2 * i + 2 * o) for (o, i, n) in shapes) * iters
            by = sum(2 * (2 * n * i + 2 * n * o) for (o, i, n) in shapes) * iters          # bf16 operand bytes streamed (L2)
            fact = sum(16 * (o + i) * 4 for (o, i, _) in shapes)
            print(json.dumps({"sites": a.sites, "cluster": cl, "iters": iters, "batch": a.batch, "nsplit": a.nsplit,
                              "k2a_us": round(med[0], 1), "k2b_us": round(med[1], 1),
                              "k2a_gflop": round(fl / 1e9, 3), "k2a_l2_stream_mb": round(by / 1e6, 1),
                              "k2a_tflops": round(fl / med[0] / 1e6, 2), "k2a_l2_gbs": round(by / med[0] / 1e3, 1),
                              "factor_bytes_per_site": fact}), flush=True)
            g.close()
            for o_ in eng.opt:
                o_.group.close()


if __name__ == "__main__":
    main()
