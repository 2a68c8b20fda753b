"""File-mode learners/reducers driven directly (no round machine): S sites, one step."""
import os

import numpy as np
import torch

from dinunet_implementations_b200 import COINNTrainer
from dinunet_implementations_b200.distrib import get_engine, flat_grads, trainable_params


class _T(COINNTrainer):
    def _init_nn_model(self):
        self.nn["net"] = torch.nn.Sequential(torch.nn.Linear(12, 20), torch.nn.ReLU(), torch.nn.BatchNorm1d(20),
                                             torch.nn.Linear(20, 3))

    def iteration(self, batch):
        out = self.nn["net"](batch["x"])
        loss = torch.nn.functional.cross_entropy(out, batch["y"])
        a, m = self.new_averages(), self.new_metrics()
        a.add(loss.item(), len(batch["y"]))
        return {"loss": loss, "averages": a, "metrics": m}


def _site(tmp, i, engine, extra=None):
    st = {"baseDirectory": str(tmp / f"base{i}"), "transferDirectory": str(tmp / f"xfer{i}"), "clientId": f"local{i}"}
    os.makedirs(st["baseDirectory"], exist_ok=True)
    cache = {"seed": 5, "learning_rate": 1e-2, "agg_engine": engine, "local_iterations": 1, "num_class": 3,
             "monitor_metric": "f1", "optimizer": "sgd", **(extra or {})}
    tr = _T(cache=cache, input={}, state=st)
    tr.init_nn()
    g = torch.Generator().manual_seed(100 + i)
    batches = [{"x": torch.randn(8, 12, generator=g), "y": torch.randint(0, 3, (8,), generator=g)} for _ in range(4)]
    it = iter(batches)
    L, _ = get_engine(engine)
    return L(trainer=tr, batch_source=lambda: next(it)), st, batches


def _relay(tmp, sites, engine, outs, cache):
    import shutil
    rst = {"baseDirectory": str(tmp / "rbase"), "transferDirectory": str(tmp / "rxfer")}
    for (_, st, _), _o in zip(sites, outs):
        d = os.path.join(rst["baseDirectory"], st["clientId"])
        os.makedirs(d, exist_ok=True)
        for f in os.listdir(st["transferDirectory"]):
            shutil.move(os.path.join(st["transferDirectory"], f), os.path.join(d, f))
    _, R = get_engine(engine)
    rin = {st["clientId"]: o for (_, st, _), o in zip(sites, outs)}
    rout = R(cache=cache, input=rin, state=rst).reduce()
    for f in os.listdir(rst["transferDirectory"]):
        for (_, st, _) in sites:
            shutil.copy(os.path.join(rst["transferDirectory"], f), os.path.join(st["baseDirectory"], f))
        os.remove(os.path.join(rst["transferDirectory"], f))
    return rout


def _true_mean_grad(sites):
    gs = []
    for (L, _, batches) in sites:
        L.trainer.train_mode()
        L.zero_grad()
        L.trainer.iteration(batches[0])["loss"].backward()
        gs.append(flat_grads(L.params).copy())
        L.zero_grad()
    return np.mean(gs, 0)


def test_dsgd_step_is_mean_gradient_and_sites_stay_identical(tmp_path):
    sites = [_site(tmp_path, i, "dSGD") for i in range(3)]
    w0 = torch.cat([p.detach().reshape(-1) for p in sites[0][0].params]).clone()
    mean_g = _true_mean_grad(sites)
    outs = [L.to_reduce()[0] for (L, _, _) in sites]
    rout = _relay(tmp_path, sites, "dSGD", outs, sites[0][0].cache)
    assert rout["update"]
    for (L, _, _) in sites:
        L.rebind(rout)
        L.step()
    ws = [torch.cat([p.detach().reshape(-1) for p in L.params]) for (L, _, _) in sites]
    assert torch.equal(ws[0], ws[1]) and torch.equal(ws[0], ws[2])       # bit-identical on all sites
    assert torch.allclose(ws[0], w0 - 1e-2 * torch.from_numpy(mean_g), atol=1e-6)


def test_dsgd_fp16_payload(tmp_path):
    sites = [_site(tmp_path, i, "dSGD", {"precision_bits": "16"}) for i in range(2)]
    outs = [L.to_reduce()[0] for (L, _, _) in sites]
    d = os.path.join(sites[0][1]["transferDirectory"], outs[0]["grads_file"])
    assert np.load(d).dtype == np.float16
    rout = _relay(tmp_path, sites, "dSGD", outs, sites[0][0].cache)
    for (L, _, _) in sites:
        L.rebind(rout)
        L.step()


def test_rankdad_exact_for_small_batch(tmp_path):
    # batch 8 <= rank 10: factors are exact, so the step equals the dSGD step
    sites = [_site(tmp_path, i, "rankDAD") for i in range(2)]
    w0 = torch.cat([p.detach().reshape(-1) for p in sites[0][0].params]).clone()
    mean_g = _true_mean_grad(sites)
    outs = [L.to_reduce()[0] for (L, _, _) in sites]
    cache = dict(sites[0][0].cache, dad_num_pow_iters=30, dad_tol=0.0, dad_reduction_rank=16)
    for (L, _, _) in sites:
        L.rank = 16
    # redo with rank 16 >= S*B so the aggregator's re-compression is exact too
    sites = [_site(tmp_path, i, "rankDAD", {"dad_reduction_rank": 16, "dad_num_pow_iters": 30, "dad_tol": 0.0})
             for i in range(2)]
    outs = [L.to_reduce()[0] for (L, _, _) in sites]
    rout = _relay(tmp_path, sites, "rankDAD", outs, cache)
    for (L, _, _) in sites:
        L.rebind(rout)
        L.step()
    ws = [torch.cat([p.detach().reshape(-1) for p in trainable_params(L.trainer.nn)]) for (L, _, _) in sites]
    assert torch.equal(ws[0], ws[1])
    assert torch.allclose(ws[0], w0 - 1e-2 * torch.from_numpy(mean_g), atol=2e-5)


def test_powersgd_two_phase_roundtrip(tmp_path):
    sites = [_site(tmp_path, i, "powerSGD", {"powersgd_rank": 3}) for i in range(2)]
    w0 = [torch.cat([p.detach().reshape(-1) for p in L.params]).clone() for (L, _, _) in sites]
    outs = [L.to_reduce()[0] for (L, _, _) in sites]
    assert all(o["psgd_phase"] == "P" for o in outs)
    rout = _relay(tmp_path, sites, "powerSGD", outs, sites[0][0].cache)
    outs2 = []
    for (L, _, _) in sites:
        L.rebind(rout)
        L.step()
        o, its = L.to_reduce()
        assert o["psgd_phase"] == "Q" and its == []       # no new batches in the second half
        outs2.append(o)
    rout = _relay(tmp_path, sites, "powerSGD", outs2, sites[0][0].cache)
    for (L, _, _) in sites:
        L.rebind(rout)
        L.step()
    ws = [torch.cat([p.detach().reshape(-1) for p in L.params]) for (L, _, _) in sites]
    assert torch.equal(ws[0], ws[1]) and not torch.equal(ws[0], w0[0])
    # error feedback holds what the rank-3 step dropped
    assert all(float(E.abs().sum()) > 0 for E in sites[0][0].E)
