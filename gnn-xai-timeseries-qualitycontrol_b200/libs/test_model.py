"""Evaluation helpers with the reference's names (libs/test_model.py): select_threshold,
calculate_threshold, calculate_metrics.  Scores come from model.predict (libgnnqc); the
threshold sweep is done in one vectorised pass over the sorted scores instead of one
scikit-learn call per candidate threshold, with identical results (tests/test_metrics.py).
Plotting is out of scope."""
import numpy as np


def _confusion_at(thresholds, scores, truth):
    """TP, FP, TN, FN of `scores > thr` for every thr (vectorised with searchsorted)."""
    order = np.argsort(scores, kind="stable")
    s, y = scores[order], truth[order].astype(np.int64)
    pos_below = np.concatenate([[0], np.cumsum(y)])          # positives among the k smallest scores
    k = np.searchsorted(s, thresholds, side="right")         # scores <= thr are predicted negative
    n, n_pos = len(s), int(y.sum())
    fn = pos_below[k]
    tp = n_pos - fn
    tn = k - fn
    fp = (n - k) - tp
    return tp, fp, tn, fn


def _mcc(tp, fp, tn, fn):
    tp, fp, tn, fn = (np.asarray(v, dtype=np.float64) for v in (tp, fp, tn, fn))
    den = np.sqrt((tp + fp) * (tp + fn) * (tn + fp) * (tn + fn))
    return np.where(den > 0, (tp * tn - fp * fn) / np.where(den > 0, den, 1.0), 0.0)


def select_threshold(predictions, anomaly_flags_true):
    """Candidate thresholds are the distinct scores rounded to 3 decimals; the one with the highest
    Matthews correlation wins, the smallest on ties (reference :9-17)."""
    if getattr(predictions, "is_cuda", False):          # scores still on the GPU: sweep there (gnnqc_select_threshold)
        import torch
        from gnnqc import metrics
        thr, mcc, _ = metrics.select_threshold(predictions, torch.as_tensor(anomaly_flags_true).to(predictions.device))
        print("Max MCC: {:.3f} for threshold: {:.3f}".format(mcc, thr))
        return np.float32(thr)
    scores = np.asarray(predictions).reshape(-1)          # keep the dtype: `score > thr` is evaluated in it
    truth = np.asarray(anomaly_flags_true).reshape(-1)
    candidates = np.unique(np.round(scores, 3))
    mcc = _mcc(*_confusion_at(candidates, scores, truth))
    best = int(np.flatnonzero(mcc >= mcc.max() - 1e-12)[0])
    print("Max MCC: {:.3f} for threshold: {:.3f}".format(mcc[best], candidates[best]))
    return candidates[best]


def calculate_threshold(model_config, preproc_config, val_batches, model, baseline=False):
    """Returns (threshold, anomaly_date_ind) like the reference (:19-36); val_batches is an iterable
    of (inputs, labels) batches."""
    info = model.model_info
    first = int(info[0].numpy())
    anomaly_date_ind = first // int(info[len(info) - 1].numpy()) if preproc_config.ds_type == "soilnet" else first
    if not model_config.calculate_threshold:
        return 0.5, anomaly_date_ind
    batches = list(val_batches)
    scores = model.predict(batches).reshape(-1)
    truth = np.concatenate([np.asarray(b[1]).reshape(-1) for b in batches])
    return select_threshold(scores, truth), anomaly_date_ind


def roc_auc(truth, scores):
    """ROC curve (fpr, tpr, thresholds from high to low, like sklearn.metrics.roc_curve without
    the collinear-point pruning) and its trapezoidal area."""
    truth = np.asarray(truth).reshape(-1).astype(np.int64)
    scores = np.asarray(scores, dtype=np.float64).reshape(-1)
    order = np.argsort(-scores, kind="stable")
    s, y = scores[order], truth[order]
    last_of_run = np.r_[np.nonzero(np.diff(s))[0], len(s) - 1]
    tps = np.cumsum(y)[last_of_run]
    fps = (last_of_run + 1) - tps
    tpr = np.r_[0.0, tps / max(int(y.sum()), 1)]
    fpr = np.r_[0.0, fps / max(int(len(y) - y.sum()), 1)]
    thr = np.r_[np.inf, s[last_of_run]]
    trapezoid = getattr(np, "trapezoid", None) or np.trapz
    return fpr, tpr, thr, float(trapezoid(tpr, fpr))


def calculate_metrics(anomaly_flags_true, anomaly_flags_pred, predictions, model_config=None, threshold=0.5, baseline=False,
                      outpath=None):
    """MCC, precision, recall, accuracy, ROC-AUC, fpr, tpr, thresholds (reference :38-59, minus the plot)."""
    truth = np.asarray(anomaly_flags_true).reshape(-1).astype(bool)
    pred = np.asarray(anomaly_flags_pred).reshape(-1).astype(bool)
    tp, fp = int((pred & truth).sum()), int((pred & ~truth).sum())
    tn, fn = int((~pred & ~truth).sum()), int((~pred & truth).sum())
    mcc = float(_mcc(tp, fp, tn, fn))
    precision = tp / (tp + fp) if tp + fp else 0.0
    recall = tp / (tp + fn) if tp + fn else 0.0
    accuracy = (tp + tn) / max(len(truth), 1)
    fpr, tpr, thr, auc_score = roc_auc(truth, predictions)
    print("MCC: {:.3f}\nPrecision: {:.3f}\nRecall: {:.3f}\nAccuracy: {:.3f}\nAUC: {:.3f} ".format(mcc, precision, recall, accuracy, auc_score))
    return mcc, precision, recall, accuracy, auc_score, fpr, tpr, thr
