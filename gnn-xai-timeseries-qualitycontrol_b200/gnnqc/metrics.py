"""Evaluation on the device: threshold selection (reference libs/test_model.py:9-17) and ROC-AUC
(:49) on CUDA tensors, so the scores never leave the GPU between `predict` and the metrics."""
import ctypes as C

from ._lib import check, lib


def _args(scores, labels):
    import torch
    if not (scores.is_cuda and labels.is_cuda) or scores.device != labels.device:
        raise ValueError("scores and labels must be CUDA tensors on the same device")
    scores = scores.reshape(-1).to(torch.float32).contiguous()
    labels = labels.reshape(-1).to(torch.int64).contiguous()
    if scores.numel() != labels.numel():
        raise ValueError("scores and labels differ in length")
    return scores, labels


def select_threshold(scores, labels):
    """Returns (threshold, mcc, (tp, fp, tn, fn)) - the threshold the reference's select_threshold picks."""
    import torch
    scores, labels = _args(scores, labels)
    thr, mcc, conf = C.c_float(0), C.c_double(0), (C.c_int64 * 4)()
    check(lib.gnnqc_select_threshold(scores.device.index, scores.data_ptr(), labels.data_ptr(), scores.numel(), C.addressof(thr), C.addressof(mcc),
                                     C.addressof(conf), torch.cuda.current_stream(scores.device).cuda_stream))
    return float(thr.value), float(mcc.value), tuple(int(v) for v in conf)


def roc_auc(labels, scores):
    import torch
    scores, labels = _args(scores, labels)
    auc = C.c_double(0)
    check(lib.gnnqc_roc_auc(scores.device.index, scores.data_ptr(), labels.data_ptr(), scores.numel(), C.addressof(auc),
                            torch.cuda.current_stream(scores.device).cuda_stream))
    return float(auc.value)
