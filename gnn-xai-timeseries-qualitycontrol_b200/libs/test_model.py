"""Drop-in for the reference's libs/test_model.py: threshold selection on a validation set
(max MCC over the rounded scores) and the test metrics.  Scoring goes through
model.predict -> libgnnqc; the metric arithmetic is scikit-learn on the host, as in the
reference (:9-17, :38-59).  Plotting is out of scope."""
import numpy as np
from sklearn.metrics import accuracy_score, auc, matthews_corrcoef, precision_score, recall_score, roc_curve


def select_threshold(predictions, anomaly_flags_true):
    threshold_ranges = np.unique(np.round(predictions, 3))
    mcc = [matthews_corrcoef(anomaly_flags_true, np.greater(predictions, thr)) for thr in threshold_ranges]
    best = threshold_ranges[int(np.argmax(mcc))]
    print("Max MCC: {:.3f} for threshold: {:.3f}".format(np.max(mcc), best))
    return best


def calculate_threshold(model_config, preproc_config, val_batches, model, baseline=False):
    """val_batches: iterable of (inputs, labels) batches.  Returns (threshold, anomaly_date_ind)
    like the reference (:19-36)."""
    info = model.model_info
    if preproc_config.ds_type == "soilnet":
        anomaly_date_ind = int(info[0].numpy() / info[len(info) - 1].numpy())
    else:
        anomaly_date_ind = int(info[0].numpy())
    if not model_config.calculate_threshold:
        return 0.5, anomaly_date_ind
    batches = list(val_batches)
    predictions = model.predict(batches).flatten()
    labels = np.concatenate([np.asarray(b[1]).reshape(-1) for b in batches])
    return select_threshold(predictions, labels), anomaly_date_ind


def calculate_metrics(anomaly_flags_true, anomaly_flags_pred, predictions, model_config=None, threshold=0.5, baseline=False,
                      outpath=None):
    mcc = matthews_corrcoef(anomaly_flags_true, anomaly_flags_pred)
    precision = precision_score(anomaly_flags_true, anomaly_flags_pred)
    recall = recall_score(anomaly_flags_true, anomaly_flags_pred)
    accuracy = accuracy_score(anomaly_flags_true, anomaly_flags_pred)
    fpr, tpr, thr = roc_curve(anomaly_flags_true, predictions)
    auc_score = auc(fpr, tpr)
    print("MCC: {:.3f}\nPrecision: {:.3f}\nRecall: {:.3f}\nAccuracy: {:.3f}\nAUC: {:.3f} ".format(
        mcc, precision, recall, accuracy, auc_score))
    return mcc, precision, recall, accuracy, auc_score, fpr, tpr, thr
