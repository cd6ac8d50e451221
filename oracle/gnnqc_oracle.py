"""CPU ORACLE - TEST INFRASTRUCTURE ONLY.  Not imported by the product (gnnqc/, libs/, xai/).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / --impl reference legs may
use this module, and only as the checker / the timed CPU baseline.

PARITY UNPINNED: the reference (timohouben/GNN-XAI-TimeSeries-QualityControl) is TensorFlow
2.11 / Keras 2.11 / spektral 1.3.0 code with no tests and no golden vectors, TensorFlow is not
installable in the build image, and the example datasets are missing from the mount.  This file
is therefore a restatement of the reference's arithmetic from its source and from the layer
configs recorded in model_*/keras_metadata.pb; it has never been compared with TF output.
What *is* pinned (tests/test_reference_pins.py): the weights (parsed from the reference's own
checkpoints by oracle/ckpt_reader.py, independently of the product's reader, and hashed against the
original bytes), every layer hyper-parameter and the layer order / inter-layer shapes against the
Keras-serialised configs in model_*/keras_metadata.pb (tests/golden/keras_layer_configs.json), and
internal consistency (fp32 vs fp64, a second independent NumPy implementation, finite differences, the
IG completeness axiom) - see tests/test_oracle.py.

Restated (reference file:line):
  general_conv          spektral 1.3.0 GeneralConv.call as configured at libs/create_model.py:185-189
  timeseries_pooling    libs/create_model.py:8-41   (type='pool', aggregation 'mean')
  graph_reshape         libs/create_model.py:242-258
  time_layer            libs/create_model.py:103-123 (layers built :61-79), Keras LSTM / MaxPooling1D
  head                  libs/create_model.py:209-213,235-239 / :337-341,372-376
  gcn_forward           libs/create_model.py:215-240
  baseline_forward      libs/create_model.py:343-377
  prepare_batch_cml     libs/preprocessing_functions.py:638-666 (layout only)
  integrated_gradients  xai/libs/integrated_gradients.py:898-942,955-1012,1185-1191
  train_step            libs/fit_model.py:61-112 (+ Keras binary_crossentropy / Adam semantics)

All functions take torch CPU tensors (float32 or float64) and a `weights` dict keyed by the
checkpoint's object-graph paths ("gcn_layer/kernel", "time_layer/time1/cell/kernel", ...).
"""
import numpy as np
import torch

BN_EPS = 1e-3      # keras_metadata.pb: BatchNormalization epsilon 0.001
LEAKY = 0.3        # model_config.yml dense.alpha / keras_metadata.pb LeakyReLU alpha
POOL = 3           # MaxPooling1D pool_size = strides = 3, padding 'valid'


def to_torch_weights(named, dtype=torch.float32):
    return {k: torch.as_tensor(np.asarray(v)).to(dtype) for k, v in named.items()}


def load_named(savedmodel_dir):
    """Weights of a reference SavedModel directory as {attribute path: float32 ndarray}, read by the
    oracle's OWN checkpoint parser (oracle/ckpt_reader.py) - never through libgnnqc's reader, so a
    mapping error in either reader shows up as a parity failure instead of cancelling out."""
    import ckpt_reader

    return ckpt_reader.load_named(savedmodel_dir)[0]


def load_weights(savedmodel_dir, dtype=torch.float32):
    return to_torch_weights(load_named(savedmodel_dir), dtype)


# ---- GeneralConv (spektral 1.3.0), inference mode ------------------------------------------------
def general_conv(features, edge_index, w):
    """features [Ntot,F]; edge_index int64 [E,2] = SparseTensor indices (row, col).

    x = PReLU(Dropout_0(BN(x W + b)));  out_i = mean_{(i,j) in E} x_j  (0 if i has no edge).
    spektral MessagePassing: sources = indices[:,1], targets = indices[:,0], aggregate 'mean'
    = unsorted_segment_mean over targets."""
    z = features @ w["gcn_layer/kernel"] + w["gcn_layer/bias"]
    z = (z - w["gcn_layer/batch_norm/moving_mean"]) / torch.sqrt(w["gcn_layer/batch_norm/moving_variance"] + BN_EPS)
    z = z * w["gcn_layer/batch_norm/gamma"] + w["gcn_layer/batch_norm/beta"]
    alpha = w["gcn_layer/activation/alpha"]
    z = torch.clamp(z, min=0) + alpha * torch.clamp(z, max=0)  # Keras PReLU: relu(x) - alpha*relu(-x)
    tgt, src = edge_index[:, 0], edge_index[:, 1]
    n = features.shape[0]
    summed = torch.zeros((n, z.shape[1]), dtype=z.dtype).index_add_(0, tgt, z[src])
    count = torch.zeros(n, dtype=z.dtype).index_add_(0, tgt, torch.ones(tgt.shape[0], dtype=z.dtype))
    return summed / torch.clamp(count, min=1).unsqueeze(1)


def timeseries_pooling(x, sample_indices, batch_size, seq_len, keep_all=True):
    """Per sample: rows -> [T, N_b, C] -> mean over nodes -> [T, C]; stack -> [B, T, C].
    keep_all=False additionally drops samples whose pooled sum is 0 / non-finite, as the
    reference does (create_model.py:30-36)."""
    out = []
    for i in range(batch_size):
        part = x[sample_indices == i]
        out.append(part.reshape(seq_len, -1, x.shape[1]).mean(dim=1))
    out = torch.stack(out)
    if not keep_all:
        s = out.sum(dim=(1, 2))
        out = out[(s != 0) & torch.isfinite(s)]
    return out


def graph_reshape(x, sample_indices, batch_size, seq_len, keep_all=True):
    """Per sample: rows -> [T, N_b, C] -> [N_b, T, C]; concat over samples."""
    parts = []
    for i in range(batch_size):
        part = x[sample_indices == i]
        parts.append(part.reshape(seq_len, -1, x.shape[1]).permute(1, 0, 2))
    out = torch.cat(parts, dim=0)
    if not keep_all:
        s = out.sum(dim=(1, 2))
        out = out[(s != 0) & torch.isfinite(s)]
    return out


# ---- TimeLayer --------------------------------------------------------------------------------
def lstm(x, W, U, b, return_sequences=True, return_gates=False):
    """Keras LSTM, gate column order i | f | c | o, zero initial state."""
    B, T, _ = x.shape
    H = U.shape[0]
    h = torch.zeros((B, H), dtype=x.dtype)
    c = torch.zeros((B, H), dtype=x.dtype)
    xw = x @ W + b
    hs = []
    for t in range(T):
        z = xw[:, t] + h @ U
        i = torch.sigmoid(z[:, :H])
        f = torch.sigmoid(z[:, H:2 * H])
        g = torch.tanh(z[:, 2 * H:3 * H])
        o = torch.sigmoid(z[:, 3 * H:])
        c = f * c + i * g
        h = o * torch.tanh(c)
        hs.append(h)
    return torch.stack(hs, dim=1) if return_sequences else h


def max_pool3(x):
    B, T, C = x.shape
    n = T // POOL
    return x[:, : n * POOL].reshape(B, n, POOL, C).max(dim=2).values


LSTM_ORDER = ["time1", "time2", "time_layers/0", "time_layers/1", "time_layers/2", "time_layers/3", "time4"]


def time_layer(x, w, prefix):
    """time1, time2, pool, (stack, stack, pool) x2, time4 -> last state [B,128]."""
    def cell(name):
        base = prefix + name + "/cell/"
        return w[base + "kernel"], w[base + "recurrent_kernel"], w[base + "bias"]

    x = lstm(x, *cell("time1"))
    x = lstm(x, *cell("time2"))
    x = max_pool3(x)
    for i in range(2):
        x = lstm(x, *cell("time_layers/%d" % (2 * i)))
        x = lstm(x, *cell("time_layers/%d" % (2 * i + 1)))
        x = max_pool3(x)
    return lstm(x, *cell("time4"), return_sequences=False)


def head(x, w, first):
    x = torch.nn.functional.leaky_relu(x @ w[first + "/kernel"] + w[first + "/bias"], LEAKY)
    x = torch.nn.functional.leaky_relu(x @ w["dense2/kernel"] + w["dense2/bias"], LEAKY)
    return torch.sigmoid(x @ w["dense_out/kernel"] + w["dense_out/bias"])


# ---- whole models -----------------------------------------------------------------------------
def baseline_forward(x, w):
    """BaselineClassifier.call for [S,T,F] sequences (CML directly; SoilNet after graph_reshape)."""
    return head(time_layer(x, w, ""), w, "dense1")


def gcn_forward_cml(features, anom, edge_index, sample_indices, w, keep_all=True):
    """GCNClassifier.call, ds_type 'cml'. features [Ntot,2] rows (b,t,n); anom [B,T,2];
    edge_index = joint (tiled, block-diagonal) adjacency indices; sample_indices [Ntot]."""
    B, T = anom.shape[0], anom.shape[1]
    x = general_conv(features, edge_index, w)
    x = timeseries_pooling(x, sample_indices, B, T, keep_all)
    x = torch.cat([anom, x], dim=-1)
    return head(time_layer(x, w, "time_layer/"), w, "dense")


def gcn_forward_soilnet(features, edge_index, sample_indices, batch_size, seq_len, w, keep_all=True):
    """GCNClassifier.call, ds_type 'soilnet': one score per (sample, node)."""
    x = general_conv(features, edge_index, w)
    x = torch.cat([x, features], dim=-1)
    x = graph_reshape(x, sample_indices, batch_size, seq_len, keep_all)
    return head(time_layer(x, w, "time_layer/"), w, "dense")


def baseline_forward_soilnet(features, sample_indices, batch_size, seq_len, w, keep_all=True):
    x = graph_reshape(features, sample_indices, batch_size, seq_len, keep_all)
    return baseline_forward(x, w)


# ---- batch layout (prepare_batch_cml / prepare_batch_soilnet) ------------------------------------
def reference_batch_layout(windows, n_features):
    """windows: list of dicts {x: [T,N,F] array, edges: [E,2] local (row, col)}.
    Returns features [sum T*N, F] in (b,t,n) order, joint tiled adjacency indices [sum T*E, 2],
    sample_indices [sum T*N] - the tensors prepare_batch_* hand to the model."""
    feats, idx, samp = [], [], []
    base = 0
    for b, wdw in enumerate(windows):
        x = np.asarray(wdw["x"])
        T, N, F = x.shape
        assert F == n_features
        feats.append(x.reshape(T * N, F))
        e = np.asarray(wdw["edges"], dtype=np.int64).reshape(-1, 2)
        for t in range(T):
            idx.append(e + base + t * N)
        samp.append(np.full(T * N, b, dtype=np.int64))
        base += T * N
    return (np.concatenate(feats).astype(np.float32), np.concatenate(idx).astype(np.int64),
            np.concatenate(samp))


# ---- integrated gradients ---------------------------------------------------------------------
def integrated_gradients_cml(features, anom, edge_index, sample_indices, w, m_steps=50, scale_gradients=True):
    """Zero baseline, alphas = linspace(0,1,m_steps+1), both inputs interpolated with the same
    alpha, gradient of sum_b p_b per alpha, trapezoid mean, scaled by (x - 0).
    Returns (attr_features [Ntot,2], attr_anom [B,T,2], scores [B])."""
    alphas = torch.linspace(0.0, 1.0, m_steps + 1, dtype=features.dtype)
    gf, ga = [], []
    for a in alphas:
        f = (a * features).clone().requires_grad_(True)
        an = (a * anom).clone().requires_grad_(True)
        p = gcn_forward_cml(f, an, edge_index, sample_indices, w)
        g1, g2 = torch.autograd.grad(p.sum(), [f, an])
        gf.append(g1)
        ga.append(g2)
    gf, ga = torch.stack(gf), torch.stack(ga)
    igf = ((gf[:-1] + gf[1:]) / 2.0).mean(dim=0)
    iga = ((ga[:-1] + ga[1:]) / 2.0).mean(dim=0)
    with torch.no_grad():
        scores = gcn_forward_cml(features, anom, edge_index, sample_indices, w).reshape(-1)
    if not scale_gradients:        # xai_config scale_gradients: False -> the trapezoid mean itself (:1006-1012)
        return igf, iga, scores
    return features * igf, anom * iga, scores


# ---- independent NumPy forward (cross-check of the torch restatement) ----------------------------
def np_lstm_stack_head(x, named, prefix, first_dense):
    """Plain NumPy float64 loops over time; written separately from the torch path on purpose."""
    def sig(v):
        return 1.0 / (1.0 + np.exp(-v))

    def run(x, name, seq=True):
        W = named[prefix + name + "/cell/kernel"].astype(np.float64)
        U = named[prefix + name + "/cell/recurrent_kernel"].astype(np.float64)
        b = named[prefix + name + "/cell/bias"].astype(np.float64)
        H = U.shape[0]
        h = np.zeros((x.shape[0], H))
        c = np.zeros((x.shape[0], H))
        out = np.zeros((x.shape[0], x.shape[1], H))
        for t in range(x.shape[1]):
            z = x[:, t] @ W + h @ U + b
            c = sig(z[:, H:2 * H]) * c + sig(z[:, :H]) * np.tanh(z[:, 2 * H:3 * H])
            h = sig(z[:, 3 * H:]) * np.tanh(c)
            out[:, t] = h
        return out if seq else h

    def pool(x):
        n = x.shape[1] // 3
        return x[:, : 3 * n].reshape(x.shape[0], n, 3, x.shape[2]).max(axis=2)

    x = np.asarray(x, dtype=np.float64)
    x = pool(run(run(x, "time1"), "time2"))
    x = pool(run(run(x, "time_layers/0"), "time_layers/1"))
    x = pool(run(run(x, "time_layers/2"), "time_layers/3"))
    x = run(x, "time4", seq=False)

    def lrelu(v):
        return np.where(v > 0, v, LEAKY * v)

    x = lrelu(x @ named[first_dense + "/kernel"].astype(np.float64) + named[first_dense + "/bias"])
    x = lrelu(x @ named["dense2/kernel"].astype(np.float64) + named["dense2/bias"])
    return sig(x @ named["dense_out/kernel"].astype(np.float64) + named["dense_out/bias"]).reshape(-1)


# ---- training step ---------------------------------------------------------------------------------
# libs/fit_model.py:61-112: model.compile(loss='binary_crossentropy', Adam(lr)), model.fit(class_weight=...).
# Keras semantics restated: the sigmoid output keeps its logits (keras/activations.py sigmoid sets
# _keras_logits), so binary_crossentropy is sigmoid_cross_entropy_with_logits; class_weight becomes a
# per-sample weight and the loss is sum(w_b * bce_b) / batch_size; BatchNormalization in training mode
# normalises with the (biased) batch statistics over ALL node rows of the batch and moves the running
# statistics with momentum 0.99; Adam is the Keras 2.11 update with epsilon 1e-7.
BN_MOMENTUM = 0.99
ADAM_B1, ADAM_B2, ADAM_EPS = 0.9, 0.999, 1e-7
TRAINABLE_EXCLUDE = ("gcn_layer/batch_norm/moving_mean", "gcn_layer/batch_norm/moving_variance")


def _general_conv_train(features, edge_index, w):
    z = features @ w["gcn_layer/kernel"] + w["gcn_layer/bias"]
    mu = z.mean(dim=0)
    var = z.var(dim=0, unbiased=False)
    zn = (z - mu) / torch.sqrt(var + BN_EPS)
    z = zn * w["gcn_layer/batch_norm/gamma"] + w["gcn_layer/batch_norm/beta"]
    alpha = w["gcn_layer/activation/alpha"]
    z = torch.clamp(z, min=0) + alpha * torch.clamp(z, max=0)
    tgt, src = edge_index[:, 0], edge_index[:, 1]
    n = features.shape[0]
    summed = torch.zeros((n, z.shape[1]), dtype=z.dtype).index_add_(0, tgt, z[src])
    count = torch.zeros(n, dtype=z.dtype).index_add_(0, tgt, torch.ones(tgt.shape[0], dtype=z.dtype))
    return summed / torch.clamp(count, min=1).unsqueeze(1), mu.detach(), var.detach()


def _logits_head(x, w, first):
    x = torch.nn.functional.leaky_relu(x @ w[first + "/kernel"] + w[first + "/bias"], LEAKY)
    x = torch.nn.functional.leaky_relu(x @ w["dense2/kernel"] + w["dense2/bias"], LEAKY)
    return x @ w["dense_out/kernel"] + w["dense_out/bias"]


def loss_and_grads(weights, labels, class_weights, forward_logits, global_batch=None):
    """weights: dict of leaf tensors (requires_grad on the trainable ones).  Returns (loss, grads)."""
    logits = forward_logits(weights).reshape(-1)
    y = labels.to(logits.dtype)
    sw = torch.where(labels > 0, torch.as_tensor(class_weights[1], dtype=logits.dtype), torch.as_tensor(class_weights[0], dtype=logits.dtype))
    bce = torch.nn.functional.binary_cross_entropy_with_logits(logits, y, reduction="none")
    n = global_batch if global_batch is not None else labels.numel()
    loss = (sw * bce).sum() / n
    names = [k for k in weights if k not in TRAINABLE_EXCLUDE]
    grads = torch.autograd.grad(loss, [weights[k] for k in names], allow_unused=True)
    return loss.detach(), {k: (g if g is not None else torch.zeros_like(weights[k])) for k, g in zip(names, grads)}


def train_grads_cml(features, anom, edge_index, sample_indices, labels, named, class_weights=(1.0, 5.0), dtype=torch.float64):
    """One training-mode forward/backward of the CML GCN model.  Returns (loss, grads dict,
    batch mean, batch variance of the GeneralConv pre-normalisation activations)."""
    w = {k: torch.as_tensor(np.asarray(v)).to(dtype).clone().requires_grad_(k not in TRAINABLE_EXCLUDE) for k, v in named.items()}
    stats = {}

    def fwd(w):
        B, T = anom.shape[0], anom.shape[1]
        x, mu, var = _general_conv_train(features.to(dtype), edge_index, w)
        stats["mean"], stats["var"] = mu, var
        x = timeseries_pooling(x, sample_indices, B, T)
        x = torch.cat([anom.to(dtype), x], dim=-1)
        return _logits_head(time_layer(x, w, "time_layer/"), w, "dense")

    loss, grads = loss_and_grads(w, labels, class_weights, fwd)
    return loss, grads, stats["mean"], stats["var"]


def train_grads_soilnet(features, edge_index, sample_indices, batch_size, seq_len, labels, named, class_weights=(1.0, 5.0), dtype=torch.float64):
    """One training-mode forward/backward of the SoilNet GCN model (create_model.py:224-231): one label and one loss
    term per (window, sensor), BatchNorm statistics over all node rows.  Returns (loss, grads, batch mean, batch variance)."""
    w = {k: torch.as_tensor(np.asarray(v)).to(dtype).clone().requires_grad_(k not in TRAINABLE_EXCLUDE) for k, v in named.items()}
    stats = {}

    def fwd(w):
        f = features.to(dtype)
        x, mu, var = _general_conv_train(f, edge_index, w)
        stats["mean"], stats["var"] = mu, var
        x = torch.cat([x, f], dim=-1)
        x = graph_reshape(x, sample_indices, batch_size, seq_len)
        return _logits_head(time_layer(x, w, "time_layer/"), w, "dense")

    loss, grads = loss_and_grads(w, labels, class_weights, fwd)
    return loss, grads, stats["mean"], stats["var"]


def train_grads_baseline(x, labels, named, class_weights=(1.0, 5.0), dtype=torch.float64):
    w = {k: torch.as_tensor(np.asarray(v)).to(dtype).clone().requires_grad_(True) for k, v in named.items()}
    return loss_and_grads(w, labels, class_weights, lambda w: _logits_head(time_layer(x.to(dtype), w, ""), w, "dense1"))


def adam_update(param, grad, m, v, step, lr):
    """Keras 2.11 Adam: step counts from 1."""
    m = ADAM_B1 * m + (1 - ADAM_B1) * grad
    v = ADAM_B2 * v + (1 - ADAM_B2) * grad * grad
    lr_t = lr * np.sqrt(1 - ADAM_B2 ** step) / (1 - ADAM_B1 ** step)
    return param - lr_t * m / (torch.sqrt(v) + ADAM_EPS), m, v
