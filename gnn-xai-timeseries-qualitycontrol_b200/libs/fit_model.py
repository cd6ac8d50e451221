"""Drop-in for the reference's libs/fit_model.py (calculate_weights :8-25, MCC_custom :28-58,
train_model :61-112): same function names and config keys; the train step itself runs in
libgnnqc.so (gnnqc_train_*), data-parallel over torch.distributed when a process group is up.

Differences that follow from not having Keras: batches are python iterables of
(inputs, labels) with inputs in the offsets layout of include/gnnqc.h (dict) or the plain
[B,T,2] tensor of the baseline; metrics are computed on the host with scikit-learn at epoch end;
PlotLossesKeras is dropped.  Optimisers: adam, sgd, rmsprop with the Keras defaults, as in the reference.
"""
import copy

import numpy as np
import torch
from sklearn.metrics import matthews_corrcoef

from gnnqc import dp


def calculate_weights(model_config, train_dataset_batched):
    wc = model_config.weight_classes
    if not wc.use:
        return None
    if wc.calculate:
        flags = np.concatenate([np.asarray(torch.as_tensor(b[1]).cpu()).reshape(-1) for b in train_dataset_batched])
        total, positives = len(flags), flags.sum()
        return {0: total / (total - positives), 1: 2 * total / positives}
    if wc.class_0 is not None and wc.class_1 is not None:
        return {0: wc.class_0, 1: wc.class_1}
    return {0: 1, 1: 5}


def mcc_metric(y_true, y_pred, threshold=0.5):
    return matthews_corrcoef(np.asarray(y_true).reshape(-1), (np.asarray(y_pred).reshape(-1) > threshold).astype(np.int64))


class MCC_custom:
    """Epoch-end MCC on the training (and validation) batches, like the reference's callback."""

    def __init__(self, train, validation=None):
        self.train, self.validation = train, validation

    def on_epoch_end(self, model, logs):
        def score(batches):
            y = np.concatenate([np.asarray(torch.as_tensor(b[1]).cpu()).reshape(-1) for b in batches])
            return mcc_metric(y, model.predict(batches))
        logs["MCC_score_train"] = np.round(score(self.train), 5)
        if self.validation:
            logs["MCC_score_val"] = np.round(score(self.validation), 5)


def learning_rate_for_epoch(model_config, epoch, lr):
    """LearningRateScheduler(scheduler) of the reference (:96-102): from epoch `after_epochs` on the
    current rate is multiplied by `rate` at every epoch start."""
    sch = model_config.learning_learn_scheduler
    if sch.use and epoch >= sch.after_epochs:
        return lr * sch.rate
    return lr


def _bce_from_scores(p, y, class_weights):
    p = np.clip(np.asarray(p, dtype=np.float64).reshape(-1), 1e-7, 1 - 1e-7)
    y = np.asarray(y, dtype=np.float64).reshape(-1)
    w = np.where(y > 0, class_weights[1], class_weights[0]) if class_weights else 1.0
    return float(np.mean(w * -(y * np.log(p) + (1 - y) * np.log(1 - p))))


def _device_of(native):
    return torch.device("cuda", native.device) if native.device is not None and native.device >= 0 else torch.device("cpu")


def shard_for_rank(native, inputs, labels):
    """This rank's part of one global batch, still wherever the batch lives (host or device): (kind, shard, labels,
    global batch size).  Splitting first lets train_model upload only the shard, one batch ahead of the step."""
    rank, world_size = dp.world()
    labels = torch.as_tensor(labels).to(torch.int64)
    if native.has_gcn and native.soilnet:
        n_seq = int(inputs["node_off"][-1])                      # the loss is the mean over all sensor-sequences of the batch
        sub, lab = dp.shard_soil_batch(inputs, labels, rank, world_size)
        return "soilnet", sub, lab, n_seq
    if native.has_gcn:
        sub, lab = dp.shard_cml_batch(inputs, labels, rank, world_size)
        return "cml", sub, lab, inputs["anom"].shape[0]
    x = torch.as_tensor(inputs)
    xs, lab = dp.shard_seq_batch(x, labels, rank, world_size)
    return "seq", {"x": xs}, lab, x.shape[0]


def upload_shard(native, sub, lab, non_blocking=True):
    dev = _device_of(native)
    sub = {k: (v.to(dev, non_blocking=non_blocking) if isinstance(v, torch.Tensor) else v) for k, v in sub.items()}
    return sub, lab.to(dev, non_blocking=non_blocking)


def step_on_shard(native, kind, sub, lab, global_batch, class_weights, lr):
    """The optimisation step on a shard that is on the device already; returns the global loss as a 1-element device
    tensor (reading it is the caller's synchronisation point).  A rank whose shard is empty (the last, partial batch
    of an epoch can hold fewer windows than there are ranks; Keras' fit trains on it too) skips the device calls and
    contributes zeros to both all-reduces, so the collectives stay matched on every rank."""
    dev = _device_of(native)
    cw = (class_weights[0], class_weights[1]) if class_weights else (1.0, 1.0)
    if kind == "soilnet":
        empty = sub["n_nodes_total"] == 0
        moments = torch.zeros(10, dtype=torch.float64, device=dev) if empty else native.train_moments_soilnet(sub["features"])
        dp.all_reduce_sum(moments)
        if empty:
            rvec = torch.zeros(native.train_vector_len(), dtype=torch.float32, device=dev)
        else:
            rvec = native.train_grads_soilnet(sub["features"], sub["edges"], sub["node_off"].to(torch.int32), sub["edge_off"], lab,
                                              sub["T"], moments, cw, global_batch=global_batch, n_nodes_total=sub["n_nodes_total"])
        dp.all_reduce_sum(rvec)
        native.train_apply(rvec, moments, lr)
    elif kind == "cml":
        empty = sub["anom"].shape[0] == 0
        moments = torch.zeros(6, dtype=torch.float64, device=dev) if empty else native.train_moments_cml(sub["features"])
        dp.all_reduce_sum(moments)
        if empty:
            rvec = torch.zeros(native.train_vector_len(), dtype=torch.float32, device=dev)
        else:
            rvec = native.train_grads_cml(sub["features"], sub["anom"], sub["edges"], sub["node_off"].to(torch.int32), sub["edge_off"],
                                          lab, moments, cw, global_batch=global_batch, n_nodes_total=sub["n_nodes_total"])
        dp.all_reduce_sum(rvec)
        native.train_apply(rvec, moments, lr)
    else:
        xs = sub["x"]
        if xs.shape[0] == 0:
            rvec = torch.zeros(native.train_vector_len(), dtype=torch.float32, device=dev)
        else:
            rvec = native.train_grads_seq(xs.to(dtype=torch.float32), lab, cw, global_batch=global_batch)
        dp.all_reduce_sum(rvec)
        native.train_apply(rvec, None, lr)
    return rvec[-1:]


def train_step(native, inputs, labels, class_weights, lr):
    """One optimisation step on one global batch; returns the (global) loss.  Works for one process
    or for every rank of a torch.distributed group (each rank passes the same global batch and
    works on its own shard)."""
    kind, sub, lab, global_batch = shard_for_rank(native, inputs, labels)
    sub, lab = upload_shard(native, sub, lab)
    return float(step_on_shard(native, kind, sub, lab, global_batch, class_weights, lr).item())


class _Uploads:
    """The shard of batch i + 1 travels to the device on a copy stream while the step on batch i runs (what
    dataset.prefetch does for the reference's model.fit).  Host tensors are pinned once so the copies are asynchronous."""

    def __init__(self, native, batches):
        self.native, self.batches = native, batches
        self.dev = _device_of(native)
        self.cuda = self.dev.type == "cuda"
        self.stream = torch.cuda.Stream(self.dev) if self.cuda else None
        self.pinned = {}

    def _pin(self, i, sub, lab):
        if i not in self.pinned:
            pin = lambda v: v.pin_memory() if isinstance(v, torch.Tensor) and not v.is_cuda and not v.is_pinned() and v.numel() > 0 else v
            self.pinned[i] = ({k: pin(v) for k, v in sub.items()}, pin(lab))
        return self.pinned[i]

    def _start(self, i):
        inputs, labels = self.batches[i]
        kind, sub, lab, gb = shard_for_rank(self.native, inputs, labels)
        if not self.cuda:
            return kind, sub, lab, gb, None
        sub, lab = self._pin(i, sub, lab)
        with torch.cuda.stream(self.stream):
            sub, lab = upload_shard(self.native, sub, lab)
            ready = torch.cuda.Event()
            ready.record(self.stream)
        return kind, sub, lab, gb, ready

    def __iter__(self):
        nxt = self._start(0) if self.batches else None
        for i in range(len(self.batches)):
            kind, sub, lab, gb, ready = nxt
            nxt = self._start(i + 1) if i + 1 < len(self.batches) else None
            if ready is not None:
                cur = torch.cuda.current_stream(self.dev)
                cur.wait_event(ready)
                for v in list(sub.values()) + [lab]:       # allocated on the copy stream, used on this one
                    if isinstance(v, torch.Tensor) and v.is_cuda:
                        v.record_stream(cur)
            yield kind, sub, lab, gb


def train_model(model, model_config, preproc_config, train_dataset_batched, val_dataset_batched=None, baseline=False,
                classes_weights=None, labels=None, checkpoint_path=None):
    """Returns (history, model).  history: dict of per-epoch lists (loss, val_loss, lr, MCC_score_*)."""
    if model_config.optimizer not in ("adam", "sgd", "rmsprop"):
        raise ValueError("optimizer must be 'adam', 'sgd' or 'rmsprop' (libs/fit_model.py:71-74)")
    model.native.set_optimizer(model_config.optimizer)
    classes_weights = calculate_weights(model_config, train_dataset_batched)
    train_batches = list(train_dataset_batched)
    val_batches = list(val_dataset_batched) if val_dataset_batched is not None else None
    history = {"loss": [], "val_loss": [], "lr": []}
    lr = float(model_config.learning_rate)
    best, best_params, since_best = np.inf, None, 0
    mcc = MCC_custom(train_batches, val_batches) if model_config.get("mcc_callback", False) else None
    uploads = _Uploads(model.native, train_batches)
    for epoch in range(int(model_config.epochs)):
        lr = learning_rate_for_epoch(model_config, epoch, lr)
        losses = [step_on_shard(model.native, kind, sub, lab, gb, classes_weights, lr) for kind, sub, lab, gb in uploads]
        logs = {"loss": float(torch.cat(losses).mean().item()) if losses else float("nan"), "lr": lr}   # one read per epoch
        if val_batches:
            p = model.predict(val_batches)
            y = np.concatenate([np.asarray(torch.as_tensor(b[1]).cpu()).reshape(-1) for b in val_batches])
            logs["val_loss"] = _bce_from_scores(p, y, None)
        if mcc:
            mcc.on_epoch_end(model, logs)
        for k, v in logs.items():
            history.setdefault(k, []).append(v)
        print("Epoch %d/%d - %s" % (epoch + 1, int(model_config.epochs), " - ".join("%s: %.5g" % kv for kv in logs.items())))
        # EarlyStopping(monitor='val_loss', patience, restore_best_weights=True) + ModelCheckpoint(save_best_only)
        monitored = logs.get("val_loss")
        if monitored is not None:
            if monitored < best:
                best, since_best = monitored, 0
                best_params = copy.deepcopy(model.native.get_params())
                if checkpoint_path and dp.world()[0] == 0:
                    model.save(checkpoint_path)
            else:
                since_best += 1
                if since_best >= int(model_config.es_patience):
                    print("Early stopping at epoch %d" % (epoch + 1))
                    break
    if best_params is not None and since_best > 0:
        model.native.set_params(best_params)
    return history, model
