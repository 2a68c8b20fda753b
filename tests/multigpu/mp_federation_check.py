"""torchrun target: device-mode federation across real GPUs (one site each), unbalanced FS sites + ICA."""
import json
import os
import sys
import tempfile

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ.setdefault("DINUNET_QUIET", "1")

from dinunet_implementations_b200.data import synthetic                       # noqa: E402
from dinunet_implementations_b200.parallel.federation import DeviceFederation  # noqa: E402


def main():
    from dinunet_implementations_b200.parallel.engine import init_distributed_from_env
    import torch.distributed as dist
    rank, world, _ = init_distributed_from_env()
    tmp = os.environ.get("FED_TMP") or os.path.join(tempfile.gettempdir(), f"fed_{os.environ.get('MASTER_PORT', '0')}")
    for task, engine in (("FS-Classification", "dSGD"), ("ICA-Classification", "dSGD"), ("FS-Classification", "rankDAD")):
        root, work = os.path.join(tmp, task, engine, "data"), os.path.join(tmp, task, engine, "work")
        if rank == 0:
            if task.startswith("FS"):
                synthetic.make_fs_dataset(root, site_sizes=[73, 50, 100, 80, 120, 64, 90, 70][:world], seed=4, effect=0.6)
            else:
                synthetic.make_ica_dataset(root, num_sites=world, subjects_per_site=[64, 40, 56, 48, 64, 40, 56, 48][:world],
                                           num_components=16, temporal_size=120, window_size=10, window_stride=10,
                                           input_size=64, hidden_size=64, effect=1.0)
        dist.barrier()                         # dataset written by rank 0 before anyone reads the inputspec
        fed = DeviceFederation(task, root, work, agg_engine=engine,
                               overrides={"epochs": 3, "patience": 10, "num_workers": 0, "batch_size": 8})
        res = fed.run()
        dist.barrier()
        if rank == 0:
            name = "fs_net" if task.startswith("FS") else "net"
            ref = torch.load(os.path.join(work, f"output/local0/simulatorRun/{task}/fold_0/latest.pt"), weights_only=False)
            for r in range(1, world):
                oth = torch.load(os.path.join(work, f"output/local{r}/simulatorRun/{task}/fold_0/latest.pt"), weights_only=False)
                for k, v in ref["models"][name].items():
                    if "running_" in k or "num_batches" in k:
                        continue               # BatchNorm statistics are per-site (local data)
                    assert torch.equal(v, oth["models"][name][k]), (task, engine, r, k)
                assert os.path.exists(os.path.join(work, f"output/local{r}/simulatorRun/{task}_global_results.zip"))
            log = json.load(open(os.path.join(work, f"output/remote/simulatorRun/{task}/fold_0/logs.json")))
            assert log["agg_engine"] == engine and len(log["test_metrics"][0]) == 2
            print(f"federation {task} {engine} on {world} sites: OK test={log['test_metrics']} steps={res[0]['steps']}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
