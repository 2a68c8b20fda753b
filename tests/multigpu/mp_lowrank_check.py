"""Run under torchrun on >= 2 GPUs: device-mode rankDAD and powerSGD over real peer memory.

rankDAD: the compress kernel's tail stores every site's factors into all peers' slots over NVLink and raises
their flags; the recompress kernel waits for them.  Every site must end each step with bit-identical weights
(all sites recompress the same stack deterministically), in eager and in CUDA-graph stepping, and the loss must
go down on a fixed batch."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ.setdefault("DINUNET_QUIET", "1")

from dinunet_implementations_b200.parallel.engine import init_distributed_from_env, DeviceEngine, make_ica_trainer  # noqa: E402


def gather(t):
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t.contiguous())
    return out


def main():
    rank, world, local_rank = init_distributed_from_env()
    for engine in ("rankDAD", "powerSGD"):
        for use_graph in (False, True):
            tr = make_ica_trainer(local_rank, batch_size=8, hidden_size=64, input_size=32, num_components=8, window_size=5,
                                  temporal_size=40, seed=3)
            tr.cache["agg_engine"] = engine
            tr.nn["net"].classifier[0].p = 0.0
            eng = DeviceEngine(tr, use_graph=use_graph, graph_warmup=2)
            gen = torch.Generator().manual_seed(1000 + rank)          # different data on every site
            batch = {"inputs": torch.randn(8, 8, 8, 5, generator=gen), "labels": torch.randint(0, 2, (8,), generator=gen)}
            losses = []
            for i in range(14):
                losses.append(float(eng.train_step(batch)["loss"].detach()))
                if i in (0, 5, 13):
                    torch.cuda.synchronize()
                    flat = torch.cat([p.detach().flatten() for p in eng.params])
                    assert torch.isfinite(flat).all(), (engine, use_graph, i)
                    for other in gather(flat):
                        assert torch.equal(other, flat), (engine, use_graph, i, "sites diverged")
            if use_graph:
                assert eng._graph is not None, (engine, "step must be graph-capturable")
            assert losses[-1] < losses[0], (engine, use_graph, losses)
            dist.barrier()
            eng.close()
        if rank == 0:
            print(f"DeviceEngine {engine} multi-site: OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
