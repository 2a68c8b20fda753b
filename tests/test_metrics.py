import numpy as np
import torch

from dinunet_implementations_b200.metrics import COINNAverages, Prf1a, ConfusionMatrix, AUCROCMetrics, roc_auc


def test_averages_weighted_and_reduce():
    a = COINNAverages()
    a.add(1.0, 10)
    a.add(torch.tensor(3.0), 30)
    assert abs(a.average - 2.5) < 1e-9
    b = COINNAverages()
    b.add(0.0, 40)
    r = COINNAverages().reduce_sites([a.serialize(), b.serialize()])
    assert abs(r.average - 1.25) < 1e-9
    assert r.get() == [1.25]


def test_prf1a_counts():
    m = Prf1a()
    m.add(torch.tensor([1, 1, 0, 0, 1]), torch.tensor([1, 0, 0, 1, 1]))
    assert (m.tp, m.fp, m.tn, m.fn) == (2, 1, 1, 1)
    assert abs(m.precision - 2 / 3) < 1e-9 and abs(m.recall - 2 / 3) < 1e-9
    assert abs(m.accuracy - 0.6) < 1e-9
    r = Prf1a().reduce_sites([m.serialize(), m.serialize()])
    assert r.tp == 4 and abs(r.f1 - m.f1) < 1e-9


def test_confusion_matrix():
    c = ConfusionMatrix(3)
    c.add([0, 1, 2, 2], [0, 1, 2, 1])
    assert c.matrix.sum() == 4 and abs(c.accuracy - 0.75) < 1e-9
    r = ConfusionMatrix(3).reduce_sites([c.serialize()])
    assert np.allclose(r.matrix, c.matrix)


def test_auc_matches_sklearn_with_ties():
    from sklearn.metrics import roc_auc_score
    rng = np.random.default_rng(0)
    y = rng.integers(0, 2, 500)
    s = np.round(rng.random(500) * 0.6 + y * 0.3, 2)  # ties on purpose
    assert abs(roc_auc(s, y) - roc_auc_score(y, s)) < 1e-12
    # hard predictions (the FreeSurfer trainer feeds argmax, comps/fs/__init__.py:57-59)
    p = (s > 0.5).astype(int)
    assert abs(roc_auc(p, y) - roc_auc_score(y, p)) < 1e-12


def test_auc_metric_is_union_not_mean_of_sites():
    rng = np.random.default_rng(1)
    m1, m2 = AUCROCMetrics(), AUCROCMetrics()
    y1, y2 = rng.integers(0, 2, 50), rng.integers(0, 2, 70)
    s1, s2 = rng.random(50) + 0.5 * y1, rng.random(70) * 2 + 0.2 * y2
    m1.add(torch.tensor(s1), torch.tensor(y1))
    m2.add(s2, y2)
    r = AUCROCMetrics().reduce_sites([m1.serialize(), m2.serialize()])
    assert abs(r.auc - roc_auc(np.concatenate([s1, s2]), np.concatenate([y1, y2]))) < 1e-12
    assert r.get() == [round(r.auc, 5)]
    assert 0 <= r.accuracy <= 1 and 0 <= r.f1 <= 1


def test_auc_single_class_is_defined():
    m = AUCROCMetrics()
    m.add([0.2, 0.9], [1, 1])
    assert m.auc == 0.5
