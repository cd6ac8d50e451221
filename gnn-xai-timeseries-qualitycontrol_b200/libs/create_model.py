"""Drop-in for the reference's libs/create_model.py: same class names, constructor arguments
and call conventions; every forward pass runs in libgnnqc.so on a B200.

    GCNClassifier(model_config, preproc_config)        reference :140-240
    BaselineClassifier(model_config, preproc_config)   reference :261-377
    load_model(model_path)                             tf.keras.models.load_model, pipeline.ipynb:202

Tensors are torch tensors (CPU tensors / numpy arrays are copied to the model's GPU).  Only the
configuration the shipped models use is supported (GeneralConv + 'lstm' TimeLayer, mean
aggregation); anything else raises NotImplementedError rather than falling back.
"""
import math

import numpy as np
import torch

from gnnqc import NativeModel
from gnnqc import layout as _layout


def _as_device(t, device, dtype=None):
    if isinstance(t, np.ndarray):
        t = torch.from_numpy(t)
    if not isinstance(t, torch.Tensor):
        t = torch.as_tensor(t)
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    if not t.is_cuda:
        t = t.cuda(device, non_blocking=True)
    return t


def _keras_init(spec_in, spec_hid, has_gcn, n_features, dense_units, seed):
    """Keras default initialisers: glorot_uniform kernels, orthogonal recurrent kernels, zero
    biases with unit forget bias, BN gamma 1 / beta 0 / mean 0 / var 1, PReLU alpha 0."""
    g = torch.Generator().manual_seed(int(seed))
    parts = []

    def glorot(fan_in, fan_out, shape):
        lim = math.sqrt(6.0 / (fan_in + fan_out))
        return (torch.rand(shape, generator=g) * 2 - 1) * lim

    if has_gcn:
        parts += [glorot(n_features, 16, (n_features, 16)), torch.zeros(16), torch.zeros(16), torch.ones(16), torch.zeros(16),
                  torch.zeros(16), torch.ones(16)]
    for i, h in zip(spec_in, spec_hid):
        parts.append(glorot(i, 4 * h, (i, 4 * h)))
        q, r = torch.linalg.qr(torch.randn((4 * h, h), generator=g))
        q = q * torch.sign(torch.diagonal(r))[None, :]
        parts.append(q.t().contiguous())          # [h, 4h] with orthonormal rows
        b = torch.zeros(4 * h)
        b[h:2 * h] = 1.0
        parts.append(b)
    last = spec_hid[-1]
    parts += [glorot(last, dense_units, (last, dense_units)), torch.zeros(dense_units),
              glorot(dense_units, dense_units, (dense_units, dense_units)), torch.zeros(dense_units),
              glorot(dense_units, 1, (dense_units, 1)), torch.zeros(1)]
    return torch.cat([p.reshape(-1) for p in parts]).numpy().astype(np.float32)


class _ModelInfo:
    """model.model_info[i].numpy() as the reference uses it (libs/test_model.py:22-25)."""

    def __init__(self, values):
        self._v = torch.tensor(list(values), dtype=torch.int32)

    def __getitem__(self, i):
        return self._v[i]

    def __len__(self):
        return self._v.numel()

    def numpy(self):
        return self._v.numpy()


class _Classifier:
    def __init__(self, native):
        self.native = native
        self.device = native.device
        self.ds_type = native.ds_type
        self.model_info = _ModelInfo(native.model_info)
        self.sequence_length = native.seq_len

    # -- Keras-like surface ---------------------------------------------------------------
    def __call__(self, inputs, training=False):
        if training:
            raise NotImplementedError("training-mode forward is only available through libs.fit_model.train_model")
        return self._forward(inputs).reshape(-1, 1)

    def predict(self, dataset, verbose=0):
        """Iterates `dataset` (batches of `inputs` or `(inputs, labels)`), returns float32 [n, 1]
        on the host like tf.keras.Model.predict."""
        outs = []
        if isinstance(dataset, (torch.Tensor, np.ndarray)) or (isinstance(dataset, tuple) and not self._is_labelled(dataset)):
            dataset = [dataset]
        for batch in dataset:
            inputs = batch[0] if self._is_labelled(batch) else batch
            outs.append(self._forward(inputs))
        if not outs:
            return np.zeros((0, 1), dtype=np.float32)
        return torch.cat(outs).cpu().numpy().reshape(-1, 1)

    def _is_labelled(self, batch):
        raise NotImplementedError

    def save(self, path):
        self.native.save(path)

    def get_weights(self):
        return self.native.named_params()

    def count_params(self):
        return self.native.n_params


def _require(cfg, key, allowed, what):
    """The native kernels hard-wire the hyper-parameters of the shipped checkpoints (csrc/model_spec.h, checked against
    keras_metadata.pb in tests/test_reference_pins.py); a config that asks for anything else would silently build a
    different model than the reference does, so it is refused."""
    v = cfg.get(key) if hasattr(cfg, "get") else getattr(cfg, key, None)
    if isinstance(v, float) and any(isinstance(a, float) and abs(v - a) < 1e-12 for a in allowed):
        return
    if v not in allowed:
        raise NotImplementedError("%s.%s = %r is not built (supported: %s)" % (what, key, v, ", ".join(repr(a) for a in allowed)))


class GCNClassifier(_Classifier):
    def __init__(self, model_config=None, preprocessing_config=None, device=0, seed=0, _native=None):
        if _native is None:
            mc, pc = model_config, preprocessing_config
            if mc.graph_convolution.layer != "GeneralConv" or mc.sequence_layer.algorithm != "lstm":
                raise NotImplementedError("only graph_convolution.layer='GeneralConv' with sequence_layer.algorithm='lstm' is built")
            if mc.graph_convolution.aggregation_type != "mean" or mc.pooling.aggregation_type != "mean":
                raise NotImplementedError("only mean aggregation is built")
            if mc.graph_convolution.units != 16:
                raise NotImplementedError("graph_convolution.units must be 16")
            _require(mc.graph_convolution, "activation", ("prelu",), "graph_convolution")
            _require(mc.graph_convolution, "regularizer", (None,), "graph_convolution")
            _require(mc.graph_convolution, "dropout_rate", (0, 0.0, None), "graph_convolution")
            _require(mc.sequence_layer, "activation", ("tanh",), "sequence_layer")
            _require(mc.sequence_layer, "pool_size", (3,), "sequence_layer")
            _require(mc.sequence_layer, "regularizer", (None,), "sequence_layer")
            _require(mc.dense, "alpha", (0.3,), "dense")
            _require(mc.dense, "regularizer", (None,), "dense")
            soil = pc.ds_type == "soilnet"
            freq = 15 if soil else 1
            probe = NativeModel.create(True, soil, None, mc.sequence_layer.filter_1_size, mc.sequence_layer.n_stacks,
                                       mc.dense.units, device=None)
            params = _keras_init(probe.lstm_in, probe.lstm_hid, True, 3 if soil else 2, mc.dense.units, seed)
            info = [pc.timestep_before, pc.timestep_after, pc.batch_size, freq]
            _native = NativeModel.create(True, soil, params, mc.sequence_layer.filter_1_size, mc.sequence_layer.n_stacks,
                                         mc.dense.units, model_info=info, normalization=pc.get("normalization", "") or "",
                                         device=device)
        super().__init__(_native)

    def _is_labelled(self, batch):
        return isinstance(batch, (tuple, list)) and len(batch) == 2 and isinstance(batch[0], (tuple, list, dict))

    def _forward(self, inputs):
        dev = self.device
        if self.ds_type == "soilnet":
            T = self.sequence_length
            if isinstance(inputs, dict):
                return self.native.score_soilnet(_as_device(inputs["features"], dev, torch.float32), _as_device(inputs["edges"], dev, torch.int64),
                                                 _as_device(inputs["node_off"], dev, torch.int32), _as_device(inputs["edge_off"], dev, torch.int64),
                                                 inputs.get("T", T), inputs.get("n_nodes_total"))
            # reference layout: (features, adjacency, sample_indices)
            features = _as_device(inputs[0], dev, torch.float32)
            idx = _as_device(_layout.adjacency_indices(inputs[1]), dev, torch.int64)
            samp = _as_device(inputs[2], dev)
            n_windows = max(int(self.model_info[2]), 1)     # the reference's partition count; no device read for it
            edges, node_off, edge_off, n_nodes = _layout.disjoint_to_offsets(idx, samp, n_windows, T)
            return self.native.score_soilnet(features, edges, node_off, edge_off, T, n_nodes)
        if isinstance(inputs, dict):     # offsets layout, see include/gnnqc.h
            return self.native.score_cml(_as_device(inputs["features"], dev, torch.float32), _as_device(inputs["anom"], dev, torch.float32),
                                         _as_device(inputs["edges"], dev, torch.int64), _as_device(inputs["node_off"], dev, torch.int32),
                                         _as_device(inputs["edge_off"], dev, torch.int64), inputs.get("n_nodes_total"))
        # reference layout: (features, anomalous_sensor_features, adjacency, sample_indices, sensor_ind)
        features, anom, adjacency, sample_indices = inputs[0], inputs[1], inputs[2], inputs[3]
        features = _as_device(features, dev, torch.float32)
        anom = _as_device(anom, dev, torch.float32)
        idx = _as_device(_layout.adjacency_indices(adjacency), dev, torch.int64)
        samp = _as_device(sample_indices, dev)
        edges, node_off, edge_off, n_nodes = _layout.disjoint_to_offsets(idx, samp, anom.shape[0], anom.shape[1])
        return self.native.score_cml(features, anom, edges, node_off, edge_off, n_nodes)


class BaselineClassifier(_Classifier):
    def __init__(self, model_config=None, preprocessing_config=None, device=0, seed=0, _native=None):
        if _native is None:
            mc, pc = model_config, preprocessing_config
            bm = mc.baseline_model
            if bm.type == "cnn":
                raise NotImplementedError("only baseline_model.type='lstm' is built")
            if bm.regularizer is not None:
                raise NotImplementedError("baseline_model.regularizer must be null (the reference's branch is broken, create_model.py:288-291)")
            _require(bm, "activation", ("tanh",), "baseline_model")
            _require(bm, "pool_size", (3,), "baseline_model")
            _require(bm, "alpha", (0.3,), "baseline_model")
            soil = pc.ds_type == "soilnet"
            freq = 15 if soil else 1
            probe = NativeModel.create(False, soil, None, bm.filter_1_size, bm.n_stacks, bm.dense_layer_units, device=None)
            params = _keras_init(probe.lstm_in, probe.lstm_hid, False, 3 if soil else 2, bm.dense_layer_units, seed)
            info = [pc.timestep_before, pc.timestep_after, pc.batch_size, freq]
            _native = NativeModel.create(False, soil, params, bm.filter_1_size, bm.n_stacks, bm.dense_layer_units,
                                         model_info=info, normalization=pc.get("normalization", "") or "", device=device)
        super().__init__(_native)

    def _is_labelled(self, batch):
        if not isinstance(batch, (tuple, list)) or len(batch) != 2:
            return False
        if self.ds_type == "cml":
            return True                      # (x, labels)
        return isinstance(batch[0], (tuple, list))   # ((features, sample_indicator), labels)

    def _forward(self, inputs):
        dev = self.device
        if self.ds_type == "cml":
            return self.native.score_seq(_as_device(inputs, dev, torch.float32))
        features, sample_indices = inputs
        features = _as_device(features, dev, torch.float32)
        samp = _as_device(sample_indices, dev, torch.int64)
        T = self.sequence_length
        # graph_reshape (reference :242-258) happens inside gnnqc_score_rows; only the per-window node counts are formed
        # here, on the device and without a host read: the reference partitions into `batch_size` windows (model_info[2]),
        # short last batches leave the trailing ones empty
        n_windows = max(int(self.model_info[2]), 1)
        node_off = _layout.node_offsets(samp, n_windows, T)
        return self.native.score_rows(features, node_off, T, features.shape[0] // T)


def load_model(model_path, device=0, compile=False):
    """tf.keras.models.load_model(model_path, compile=False) for the four SavedModel
    directories the reference ships (and anything saved by `.save`)."""
    native = NativeModel.load(model_path, device=device)
    cls = GCNClassifier if native.has_gcn else BaselineClassifier
    return cls(_native=native)
