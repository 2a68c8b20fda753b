"""torchrun target (CPU, gloo): DeviceFederation's host-side logic with one process per site.

Sites of different sizes (the smaller ones wrap around inside an epoch), k-fold, validation / early stop /
best checkpoint / test, metric merging as sufficient statistics, reference-layout outputs."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ.setdefault("DINUNET_QUIET", "1")

from dinunet_implementations_b200.data import synthetic                       # noqa: E402
from dinunet_implementations_b200.parallel.engine import init_distributed_from_env  # noqa: E402
from dinunet_implementations_b200.parallel.federation import DeviceFederation  # noqa: E402


def main():
    import torch.distributed as dist
    rank, world, _ = init_distributed_from_env()
    tmp = os.environ["FED_TMP"]
    root, work = os.path.join(tmp, "data"), os.path.join(tmp, "work")
    if rank == 0:
        synthetic.make_fs_dataset(root, site_sizes=[70, 45, 60][:world], seed=7, effect=0.6)
    dist.barrier()
    fed = DeviceFederation("FS-Classification", root, work,
                           overrides={"epochs": 4, "num_folds": 3, "patience": 2, "num_workers": 0, "batch_size": 8,
                                      "precision_bits": "16"})
    res = fed.run()
    dist.barrier()
    if rank == 0:
        assert len(res) == 3
        for k in range(3):
            ref = torch.load(os.path.join(work, f"output/local0/simulatorRun/FS-Classification/fold_{k}/latest.pt"), weights_only=False)
            for r in range(1, world):
                oth = torch.load(os.path.join(work, f"output/local{r}/simulatorRun/FS-Classification/fold_{k}/latest.pt"), weights_only=False)
                for name, v in ref["models"]["fs_net"].items():
                    assert torch.equal(v, oth["models"]["fs_net"][name]), (k, r, name)
            log = json.load(open(os.path.join(work, f"output/remote/simulatorRun/FS-Classification/fold_{k}/logs.json")))
            assert log["agg_engine"] == "dSGD" and len(log["test_metrics"][0]) == 2 and log["best_val_epoch"] >= 1
            assert len(log["validation_log"]) == len(log["train_log"]) <= 4
            # every site ran the same number of steps: that of the largest site's training split
            steps = [json.load(open(os.path.join(work, f"output/local{r}/simulatorRun/FS-Classification/fold_{k}/logs.json")))["steps"]
                     for r in range(world)]
            assert len(set(steps)) == 1, steps
        for r in range(world):
            assert os.path.exists(os.path.join(work, f"output/local{r}/simulatorRun/FS-Classification_global_results.zip"))
    # second run: pre-training (E10) — the site with the most data trains alone first, its weights start everyone
    work2 = os.path.join(tmp, "work_pre")
    fed2 = DeviceFederation("FS-Classification", root, work2,
                            overrides={"epochs": 2, "num_workers": 0, "batch_size": 8, "pretrain": True,
                                       "pretrain_args": {"epochs": 2, "learning_rate": 0.002}})
    res2 = fed2.run()
    dist.barrier()
    if rank == 0:
        assert os.path.exists(os.path.join(work2, "output/local0/simulatorRun/FS-Classification/fold_0/pretrained_best.pt")), \
            "site 0 has the largest split (70 subjects) and must have pre-trained"
        assert not os.path.exists(os.path.join(work2, "output/local1/simulatorRun/FS-Classification/fold_0/pretrained_best.pt"))
        assert len(res2) == 1 and res2[0]["best_val_epoch"] >= 1
    # the low-rank engines through the same driver: sites must stay identical, training must work
    for eng in ("rankDAD", "powerSGD"):
        work3 = os.path.join(tmp, "work_" + eng)
        fed3 = DeviceFederation("FS-Classification", root, work3, agg_engine=eng,
                                overrides={"epochs": 3, "num_workers": 0, "batch_size": 8, "patience": 10})
        res3 = fed3.run()
        dist.barrier()
        if rank == 0:
            ref = torch.load(os.path.join(work3, "output/local0/simulatorRun/FS-Classification/fold_0/latest.pt"), weights_only=False)
            for r in range(1, world):
                oth = torch.load(os.path.join(work3, f"output/local{r}/simulatorRun/FS-Classification/fold_0/latest.pt"), weights_only=False)
                for name, v in ref["models"]["fs_net"].items():
                    assert torch.allclose(v, oth["models"]["fs_net"][name], atol=1e-6), (eng, r, name)
            log = json.load(open(os.path.join(work3, "output/remote/simulatorRun/FS-Classification/fold_0/logs.json")))
            assert log["agg_engine"] == eng
            assert log["train_log"][-1][0] < log["train_log"][0][0] + 0.05, (eng, log["train_log"])
    if rank == 0:
        print(f"cpu federation on {world} sites: OK auc={[r['global_test_metrics'][0][1] for r in res]} steps={res[0]['steps']}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
