"""libs/test_model.py: the vectorised threshold sweep and metrics equal the reference's
scikit-learn formulation (libs/test_model.py:9-17, :38-59 of the reference)."""
import numpy as np
from sklearn.metrics import accuracy_score, auc, matthews_corrcoef, precision_score, recall_score, roc_curve

from libs import test_model


def reference_select_threshold(predictions, truth):
    cands = np.unique(np.round(predictions, 3))
    mcc = np.array([matthews_corrcoef(truth, np.greater(predictions, c)) for c in cands])
    return cands[np.flatnonzero(mcc >= mcc.max() - 1e-12)[0]], mcc.max()


def test_select_threshold_equals_sklearn_sweep():
    rng = np.random.default_rng(0)
    for n, p_pos in ((50, 0.3), (2000, 0.05), (300, 0.5)):
        truth = (rng.random(n) < p_pos).astype(int)
        truth[0], truth[1] = 0, 1
        scores = np.clip(rng.beta(0.5, 3.0, n) + 0.5 * truth * rng.random(n), 0, 1).astype(np.float32)
        want, want_mcc = reference_select_threshold(scores, truth)
        got = test_model.select_threshold(scores, truth)
        assert got == want
        assert abs(matthews_corrcoef(truth, scores > got) - want_mcc) < 1e-12


def test_metrics_equal_sklearn():
    rng = np.random.default_rng(1)
    truth = (rng.random(500) < 0.2).astype(int)
    scores = np.clip(0.6 * truth + 0.5 * rng.random(500), 0, 1)
    pred = scores > 0.55
    mcc, pr, rc, acc, a, fpr, tpr, thr = test_model.calculate_metrics(truth, pred, scores)
    assert abs(mcc - matthews_corrcoef(truth, pred)) < 1e-12
    assert abs(pr - precision_score(truth, pred)) < 1e-12 and abs(rc - recall_score(truth, pred)) < 1e-12
    assert abs(acc - accuracy_score(truth, pred)) < 1e-12
    f2, t2, _ = roc_curve(truth, scores)
    assert abs(a - auc(f2, t2)) < 1e-12
    # degenerate cases: no predicted positives / single class
    assert test_model.calculate_metrics(truth, np.zeros(500, bool), scores)[0] == 0.0
