"""Device threshold selection / ROC-AUC against the host implementations in libs/test_model.py
(which tests/test_metrics.py pins against scikit-learn's definitions)."""
import numpy as np
import pytest
import torch

from gnnqc import metrics
from libs.test_model import _confusion_at, _mcc, roc_auc, select_threshold

pytestmark = pytest.mark.gpu


def make(n, seed, kind):
    g = np.random.default_rng(seed)
    y = (g.random(n) < 0.2).astype(np.int64)
    z = g.normal(0, 2.5, n) + 2.0 * y
    if kind == "saturated":
        z *= 12                                  # many exact 1.0f and ~0 scores: long runs of ties
    s = (1.0 / (1.0 + np.exp(-z))).astype(np.float32)
    if kind == "coarse":
        s = np.round(s, 2).astype(np.float32)    # few distinct values, scores exactly on threshold candidates
    return s, y


@pytest.mark.parametrize("n,kind", [(1, "plain"), (37, "plain"), (5000, "plain"), (200000, "plain"), (200000, "saturated"), (50000, "coarse")])
def test_select_threshold_matches_host(n, kind, capsys):
    s, y = make(n, 11 + n, kind)
    want = select_threshold(s, y)
    thr, mcc, conf = metrics.select_threshold(torch.from_numpy(s).cuda(), torch.from_numpy(y).cuda())
    assert np.float32(thr) == np.float32(want)
    tp, fp, tn, fn = _confusion_at(np.array([want], dtype=np.float32), s, y)
    assert conf == (int(tp[0]), int(fp[0]), int(tn[0]), int(fn[0]))
    assert abs(mcc - float(_mcc(tp, fp, tn, fn)[0])) < 1e-12


@pytest.mark.parametrize("n,kind", [(2, "plain"), (5000, "plain"), (300000, "plain"), (300000, "saturated"), (50000, "coarse")])
def test_roc_auc_matches_host(n, kind):
    s, y = make(n, 3 + n, kind)
    y[0], y[-1] = 1, 0
    want = roc_auc(y, s)[3]
    got = metrics.roc_auc(torch.from_numpy(y).cuda(), torch.from_numpy(s).cuda())
    assert abs(got - want) < 1e-12


def test_degenerate_inputs():
    s = torch.tensor([0.2, 0.7, 0.7], device="cuda")
    assert metrics.roc_auc(torch.tensor([1, 1, 1], device="cuda"), s) == 0.0
    with pytest.raises(Exception):
        metrics.select_threshold(torch.tensor([0.2, float("nan")], device="cuda"), torch.tensor([0, 1], device="cuda"))
