"""Handle-level Python wrapper over the C ABI.  torch is used only to own device buffers and
streams; all arithmetic happens inside libgnnqc.so."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check, lib


def _torch():
    import torch

    return torch


class NativeModel:
    """One gnnqc_model handle (one GPU, or host-only when device is None / < 0)."""

    def __init__(self, handle):
        self._h = C.c_void_p(handle)
        d = _lib.ModelDesc()
        check(lib.gnnqc_model_describe(self._h, C.byref(d)))
        self.desc = d
        self.has_gcn = bool(d.has_gcn)
        self.soilnet = bool(d.soilnet)
        self.ds_type = d.ds_type.decode()
        self.normalization = d.normalization.decode()
        self.n_features = d.n_features
        self.n_params = d.n_params
        self.n_trainable = d.n_trainable
        self.lstm_in = list(d.lstm_in[: d.n_lstm])
        self.lstm_hid = list(d.lstm_hid[: d.n_lstm])
        self.model_info = np.array(list(d.model_info[: d.n_model_info]), dtype=np.int32)
        self.seq_len = d.seq_len

    # ---- construction ------------------------------------------------------------------
    @classmethod
    def load(cls, savedmodel_dir, device=0):
        h = C.c_void_p()
        dev = -1 if device is None else int(device)
        check(lib.gnnqc_model_load(str(savedmodel_dir).encode(), dev, C.byref(h)))
        m = cls(h.value)
        m.device = dev
        return m

    @classmethod
    def create(cls, has_gcn, soilnet, params=None, filter_1_size=16, n_stacks=2, dense_units=64, model_info=None,
               normalization="", device=0):
        h = C.c_void_p()
        dev = -1 if device is None else int(device)
        p = None
        n = 0
        if params is not None:
            params = np.ascontiguousarray(params, dtype=np.float32)
            p = params.ctypes.data_as(C.POINTER(C.c_float))
            n = params.size
        mi = None
        n_mi = 0
        if model_info is not None:
            mi_arr = np.ascontiguousarray(model_info, dtype=np.int32)
            mi = mi_arr.ctypes.data_as(C.POINTER(C.c_int32))
            n_mi = mi_arr.size
        check(lib.gnnqc_model_create(int(has_gcn), int(soilnet), filter_1_size, n_stacks, dense_units, p, n, mi, n_mi,
                                     normalization.encode(), dev, C.byref(h)))
        m = cls(h.value)
        m.device = dev
        return m

    def close(self):
        if self._h:
            lib.gnnqc_model_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters --------------------------------------------------------------------
    def get_params(self):
        out = np.empty(self.n_params, dtype=np.float32)
        check(lib.gnnqc_model_get_params(self._h, out.ctypes.data_as(C.POINTER(C.c_float)), out.size))
        return out

    def set_params(self, params):
        params = np.ascontiguousarray(params, dtype=np.float32)
        check(lib.gnnqc_model_set_params(self._h, params.ctypes.data_as(C.POINTER(C.c_float)), params.size))

    def segments(self):
        """[(path, offset, shape, trainable)] of the flat parameter vector."""
        out = []
        idx = 0
        while True:
            path = (C.c_char * 128)()
            off = C.c_int64()
            shape = (C.c_int64 * 4)()
            rank = C.c_int32()
            tr = C.c_int32()
            rc = lib.gnnqc_model_segment(self._h, idx, path, C.byref(off), shape, C.byref(rank), C.byref(tr))
            if rc != _lib.OK:
                break
            out.append((path.value.decode(), off.value, tuple(shape[: rank.value]), bool(tr.value)))
            idx += 1
        return out

    def named_params(self):
        flat = self.get_params()
        return {p: flat[o : o + int(np.prod(s))].reshape(s) for p, o, s, _ in self.segments()}

    def save(self, directory):
        check(lib.gnnqc_model_save(self._h, str(directory).encode()))

    # ---- device entry points -----------------------------------------------------------
    def _stream(self):
        torch = _torch()
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _dev_tensor(self, t, dtype, name):
        torch = _torch()
        if not isinstance(t, torch.Tensor) or not t.is_cuda or t.device.index != self.device:
            raise ValueError("%s must be a CUDA tensor on device %d" % (name, self.device))
        if t.dtype != dtype:
            raise ValueError("%s must have dtype %s, got %s" % (name, dtype, t.dtype))
        return t.contiguous()

    def score_seq(self, x):
        """x: cuda f32 [n_seq, T, C] -> cuda f32 [n_seq]."""
        torch = _torch()
        x = self._dev_tensor(x, torch.float32, "x")
        if x.dim() != 3 or x.shape[2] != self.lstm_in[0]:
            raise ValueError("x must be [n_seq, T, %d], got %s" % (self.lstm_in[0], tuple(x.shape)))
        out = torch.empty(x.shape[0], dtype=torch.float32, device=x.device)
        check(lib.gnnqc_score_seq(self._h, x.data_ptr(), x.shape[0], x.shape[1], out.data_ptr(), self._stream()))
        return out

    def score_rows(self, features, node_off, T, n_nodes_total=None):
        """gnnqc_score_rows: baseline scoring from the reference's (window, t, sensor) row layout; features cuda f32
        [sum N_b * T, C], node_off cuda i32 [B+1] -> cuda f32 [sum N_b]."""
        torch = _torch()
        features = self._dev_tensor(features, torch.float32, "features")
        node_off = self._dev_tensor(node_off, torch.int32, "node_off")
        if n_nodes_total is None:
            n_nodes_total = features.shape[0] // int(T)
        out = torch.empty(n_nodes_total, dtype=torch.float32, device=features.device)
        check(lib.gnnqc_score_rows(self._h, features.data_ptr(), node_off.data_ptr(), node_off.numel() - 1, int(T), n_nodes_total,
                                   out.data_ptr(), self._stream()))
        return out

    def score_seq_noalloc(self, x, out):
        """gnnqc_score_seq_noalloc: like score_seq, but the library allocates nothing, so the
        caller also supplies `out` (cuda f32 [n_seq]).  For callers that must not allocate on
        the scoring path; see include/gnnqc.h for the size limit and the sharing rules."""
        torch = _torch()
        x = self._dev_tensor(x, torch.float32, "x")
        out = self._dev_tensor(out, torch.float32, "out")
        if x.dim() != 3 or x.shape[2] != self.lstm_in[0] or out.numel() != x.shape[0]:
            raise ValueError("x must be [n_seq, T, %d] and out [n_seq]" % self.lstm_in[0])
        check(lib.gnnqc_score_seq_noalloc(self._h, x.data_ptr(), x.shape[0], x.shape[1], out.data_ptr(), self._stream()))
        return out

    def score_cml(self, features, anom, edges, node_off, edge_off, n_nodes_total=None):
        """Arguments as documented for gnnqc_score_cml in include/gnnqc.h (all CUDA tensors)."""
        torch = _torch()
        features = self._dev_tensor(features, torch.float32, "features")
        anom = self._dev_tensor(anom, torch.float32, "anom")
        edges = self._dev_tensor(edges, torch.int64, "edges")
        node_off = self._dev_tensor(node_off, torch.int32, "node_off")
        edge_off = self._dev_tensor(edge_off, torch.int64, "edge_off")
        B, T = anom.shape[0], anom.shape[1]
        if anom.dim() != 3 or anom.shape[2] != 2 or features.dim() != 2 or features.shape[1] != 2:
            raise ValueError("expected features [Ntot,2] and anom [B,T,2]")
        if node_off.numel() != B + 1 or edge_off.numel() != B + 1:
            raise ValueError("node_off / edge_off must have B+1 entries")
        if n_nodes_total is None:
            n_nodes_total = int(node_off[-1].item())
        if features.shape[0] != n_nodes_total * T:
            raise ValueError("features has %d rows, expected sum(N_b)*T = %d" % (features.shape[0], n_nodes_total * T))
        out = torch.empty(B, dtype=torch.float32, device=anom.device)
        check(lib.gnnqc_score_cml(self._h, features.data_ptr(), anom.data_ptr(), edges.data_ptr(), node_off.data_ptr(),
                                  edge_off.data_ptr(), B, T, n_nodes_total, out.data_ptr(), self._stream()))
        return out

    def score_cml_host(self, features, anom, edges, node_off, edge_off, chunk_windows=8192, out=None):
        """gnnqc_score_cml_host: the same batch as score_cml but in HOST tensors (pin them for full PCIe
        speed).  Returns (scores [B] pinned CPU float32, bytes copied host-to-device)."""
        torch = _torch()
        import ctypes as C

        def host(t, dtype, name):
            t = torch.as_tensor(t)
            if t.is_cuda or t.dtype != dtype or not t.is_contiguous():
                raise ValueError("%s must be a contiguous CPU %s tensor" % (name, dtype))
            return t
        features, anom = host(features, torch.float32, "features"), host(anom, torch.float32, "anom")
        edges, node_off, edge_off = host(edges, torch.int64, "edges"), host(node_off, torch.int32, "node_off"), host(edge_off, torch.int64, "edge_off")
        B, T = anom.shape[0], anom.shape[1]
        if anom.dim() != 3 or anom.shape[2] != 2 or features.dim() != 2 or features.shape[1] != 2:
            raise ValueError("expected features [Ntot,2] and anom [B,T,2]")
        if node_off.numel() != B + 1 or edge_off.numel() != B + 1:
            raise ValueError("node_off / edge_off must have B+1 entries")
        if features.shape[0] != int(node_off[-1]) * T or edges.shape[0] != int(edge_off[-1]):
            raise ValueError("features / edges do not match node_off / edge_off")
        if out is None:
            out = torch.empty(B, dtype=torch.float32).pin_memory()
        moved = C.c_int64(0)
        check(lib.gnnqc_score_cml_host(self._h, features.data_ptr(), anom.data_ptr(), edges.data_ptr(), node_off.data_ptr(), edge_off.data_ptr(),
                                       B, T, int(chunk_windows), out.data_ptr(), C.addressof(moved)))
        return out, int(moved.value)

    def score_soilnet(self, features, edges, node_off, edge_off, T, n_nodes_total=None):
        """SoilNet GCN: features cuda f32 [sum N_b*T, 3] -> scores cuda f32 [sum N_b]."""
        torch = _torch()
        features = self._dev_tensor(features, torch.float32, "features")
        edges = self._dev_tensor(edges, torch.int64, "edges")
        node_off = self._dev_tensor(node_off, torch.int32, "node_off")
        edge_off = self._dev_tensor(edge_off, torch.int64, "edge_off")
        B = node_off.numel() - 1
        if n_nodes_total is None:
            n_nodes_total = int(node_off[-1].item())
        if features.dim() != 2 or features.shape[1] != 3 or features.shape[0] != n_nodes_total * T:
            raise ValueError("features must be [sum(N_b)*T, 3]")
        out = torch.empty(n_nodes_total, dtype=torch.float32, device=features.device)
        check(lib.gnnqc_score_soilnet(self._h, features.data_ptr(), edges.data_ptr(), node_off.data_ptr(), edge_off.data_ptr(),
                                      B, int(T), n_nodes_total, edges.shape[0], out.data_ptr(), self._stream()))
        return out

    def score_cml_series(self, series, target, t_anom, timestep_before=120, timestep_after=60, return_counts=False):
        """Scores windows named by (target sensor, time index) straight from a resident CmlSeries.
        target, t_anom: cuda int32 [B].  Windows the reference would skip come back as NaN."""
        torch = _torch()
        target = self._dev_tensor(target, torch.int32, "target")
        t_anom = self._dev_tensor(t_anom, torch.int32, "t_anom")
        B = target.numel()
        out = torch.empty(B, dtype=torch.float32, device=target.device)
        counts = torch.empty(B, dtype=torch.int32, device=target.device) if return_counts else None
        check(lib.gnnqc_score_cml_series(self._h, series._h, target.data_ptr(), t_anom.data_ptr(), B, int(timestep_before), int(timestep_after),
                                         out.data_ptr(), counts.data_ptr() if return_counts else None, self._stream()))
        return (out, counts) if return_counts else out

    def ig_cml_series(self, series, target, t_anom, timestep_before=120, timestep_after=60, m_steps=50):
        """Integrated gradients for windows named by (target sensor, time index) on a resident CmlSeries.  Returns
        (attr_nodes [B,T,32,2], attr_anom [B,T,2], scores [B], node_sensor [B,32], n_nodes [B]) as CUDA tensors."""
        torch = _torch()
        target = self._dev_tensor(target, torch.int32, "target")
        t_anom = self._dev_tensor(t_anom, torch.int32, "t_anom")
        B, T = target.numel(), int(timestep_before) + int(timestep_after) + 1
        dev = target.device
        attr_nodes = torch.empty((B, T, 32, 2), dtype=torch.float32, device=dev)
        attr_anom = torch.empty((B, T, 2), dtype=torch.float32, device=dev)
        scores = torch.empty(B, dtype=torch.float32, device=dev)
        node_sensor = torch.empty((B, 32), dtype=torch.int32, device=dev)
        n_nodes = torch.empty(B, dtype=torch.int32, device=dev)
        check(lib.gnnqc_ig_cml_series(self._h, series._h, target.data_ptr(), t_anom.data_ptr(), B, int(timestep_before), int(timestep_after),
                                      int(m_steps), attr_nodes.data_ptr(), attr_anom.data_ptr(), scores.data_ptr(), node_sensor.data_ptr(),
                                      n_nodes.data_ptr(), self._stream()))
        return attr_nodes, attr_anom, scores, node_sensor, n_nodes

    def ig_cml(self, features, anom, edges, node_off, edge_off, m_steps=50, n_nodes_total=None, scale_by_input=True):
        """Integrated gradients (zero baseline, trapezoid, scaled by the input unless scale_by_input=False) for the CML GCN
        model.  Returns (attr_features [Ntot,2], attr_anom [B,T,2], scores [B]) as CUDA tensors."""
        torch = _torch()
        features = self._dev_tensor(features, torch.float32, "features")
        anom = self._dev_tensor(anom, torch.float32, "anom")
        edges = self._dev_tensor(edges, torch.int64, "edges")
        node_off = self._dev_tensor(node_off, torch.int32, "node_off")
        edge_off = self._dev_tensor(edge_off, torch.int64, "edge_off")
        B, T = anom.shape[0], anom.shape[1]
        if n_nodes_total is None:
            n_nodes_total = int(node_off[-1].item())
        if features.shape[0] != n_nodes_total * T:
            raise ValueError("features has %d rows, expected sum(N_b)*T = %d" % (features.shape[0], n_nodes_total * T))
        attr_f = torch.empty_like(features)
        attr_a = torch.empty_like(anom)
        scores = torch.empty(B, dtype=torch.float32, device=anom.device)
        fn = lib.gnnqc_ig_cml if scale_by_input else lib.gnnqc_ig_cml_raw
        check(fn(self._h, features.data_ptr(), anom.data_ptr(), edges.data_ptr(), node_off.data_ptr(),
                               edge_off.data_ptr(), B, T, n_nodes_total, int(m_steps), attr_f.data_ptr(), attr_a.data_ptr(),
                               scores.data_ptr(), self._stream()))
        return attr_f, attr_a, scores

    # ---- training ------------------------------------------------------------------------
    def train_vector_len(self):
        return int(lib.gnnqc_train_vector_len(self._h))

    def train_moments_cml(self, features):
        """Local feature moments (cuda f64 [6]); all-reduce(sum) them across ranks before use."""
        torch = _torch()
        features = self._dev_tensor(features, torch.float32, "features")
        out = torch.empty(6, dtype=torch.float64, device=features.device)
        check(lib.gnnqc_train_moments_cml(self._h, features.data_ptr(), features.shape[0], out.data_ptr(), self._stream()))
        return out

    def train_grads_cml(self, features, anom, edges, node_off, edge_off, labels, moments, class_weights=(1.0, 5.0),
                        global_batch=None, n_nodes_total=None):
        """This rank's reduction vector (cuda f32 [train_vector_len]): gradients, raw GCN sums, loss."""
        torch = _torch()
        features = self._dev_tensor(features, torch.float32, "features")
        anom = self._dev_tensor(anom, torch.float32, "anom")
        edges = self._dev_tensor(edges, torch.int64, "edges")
        node_off = self._dev_tensor(node_off, torch.int32, "node_off")
        edge_off = self._dev_tensor(edge_off, torch.int64, "edge_off")
        labels = self._dev_tensor(labels, torch.int64, "labels")
        moments = self._dev_tensor(moments, torch.float64, "moments")
        B, T = anom.shape[0], anom.shape[1]
        if n_nodes_total is None:
            n_nodes_total = int(node_off[-1].item())
        rvec = torch.empty(self.train_vector_len(), dtype=torch.float32, device=anom.device)
        check(lib.gnnqc_train_grads_cml(self._h, features.data_ptr(), anom.data_ptr(), edges.data_ptr(), node_off.data_ptr(),
                                        edge_off.data_ptr(), labels.data_ptr(), B, T, n_nodes_total, moments.data_ptr(),
                                        float(class_weights[0]), float(class_weights[1]), int(global_batch or B), rvec.data_ptr(),
                                        self._stream()))
        return rvec

    def train_moments_soilnet(self, features):
        """Local feature moments of a SoilNet batch (cuda f64 [10]); all-reduce(sum) them across ranks before use."""
        torch = _torch()
        features = self._dev_tensor(features, torch.float32, "features")
        out = torch.empty(10, dtype=torch.float64, device=features.device)
        check(lib.gnnqc_train_moments_soilnet(self._h, features.data_ptr(), features.shape[0], out.data_ptr(), self._stream()))
        return out

    def train_grads_soilnet(self, features, edges, node_off, edge_off, labels, T, moments, class_weights=(1.0, 5.0), global_batch=None,
                            n_nodes_total=None):
        """This rank's reduction vector for a SoilNet GCN batch; labels: cuda i64 [sum N_b], one per (window, sensor)."""
        torch = _torch()
        features = self._dev_tensor(features, torch.float32, "features")
        edges = self._dev_tensor(edges, torch.int64, "edges")
        node_off = self._dev_tensor(node_off, torch.int32, "node_off")
        edge_off = self._dev_tensor(edge_off, torch.int64, "edge_off")
        labels = self._dev_tensor(labels, torch.int64, "labels")
        moments = self._dev_tensor(moments, torch.float64, "moments")
        B = node_off.numel() - 1
        if n_nodes_total is None:
            n_nodes_total = int(node_off[-1].item())
        if labels.numel() != n_nodes_total or features.shape[0] != n_nodes_total * int(T):
            raise ValueError("expected features [sum(N_b)*T, 3] and one label per (window, sensor)")
        rvec = torch.empty(self.train_vector_len(), dtype=torch.float32, device=features.device)
        check(lib.gnnqc_train_grads_soilnet(self._h, features.data_ptr(), edges.data_ptr(), node_off.data_ptr(), edge_off.data_ptr(),
                                            labels.data_ptr(), B, int(T), n_nodes_total, edges.shape[0], moments.data_ptr(),
                                            float(class_weights[0]), float(class_weights[1]), int(global_batch or n_nodes_total),
                                            rvec.data_ptr(), self._stream()))
        return rvec

    def train_grads_seq(self, x, labels, class_weights=(1.0, 5.0), global_batch=None):
        torch = _torch()
        x = self._dev_tensor(x, torch.float32, "x")
        labels = self._dev_tensor(labels, torch.int64, "labels")
        rvec = torch.empty(self.train_vector_len(), dtype=torch.float32, device=x.device)
        check(lib.gnnqc_train_grads_seq(self._h, x.data_ptr(), labels.data_ptr(), x.shape[0], x.shape[1], float(class_weights[0]),
                                        float(class_weights[1]), int(global_batch or x.shape[0]), rvec.data_ptr(), self._stream()))
        return rvec

    def set_optimizer(self, name):
        """'adam' (default), 'sgd' or 'rmsprop' - the rules train_model accepts (libs/fit_model.py:71-74), Keras defaults."""
        kinds = {"adam": 0, "sgd": 1, "rmsprop": 2}
        if name not in kinds:
            raise ValueError("optimizer must be one of %s" % sorted(kinds))
        check(lib.gnnqc_train_set_optimizer(self._h, kinds[name]))

    def train_iterations(self):
        return int(lib.gnnqc_train_iterations(self._h))

    def train_apply(self, rvec, moments=None, learning_rate=1e-4):
        torch = _torch()
        rvec = self._dev_tensor(rvec, torch.float32, "rvec")
        mp = None
        if moments is not None:
            moments = self._dev_tensor(moments, torch.float64, "moments")
            mp = moments.data_ptr()
        check(lib.gnnqc_train_apply(self._h, rvec.data_ptr(), mp, float(learning_rate), self._stream()))

    # ---- measurement ---------------------------------------------------------------------
    def launch_count(self):
        return int(lib.gnnqc_launch_count(self._h))

    def profile(self, on=True):
        check(lib.gnnqc_profile_enable(self._h, int(on)))

    def profile_read(self, max_entries=64):
        """{kernel name: (total ms, launches)} since the last read; synchronises the device."""
        names = ((C.c_char * 32) * max_entries)()
        ms = (C.c_double * max_entries)()
        cnt = (C.c_int64 * max_entries)()
        n = C.c_int()
        check(lib.gnnqc_profile_read(self._h, max_entries, C.cast(names, C.c_void_p), ms, cnt, C.byref(n)))
        return {names[i].value.decode(): (ms[i], cnt[i]) for i in range(n.value)}

    def scratch_bytes(self, n_seq, T):
        return int(lib.gnnqc_score_scratch_bytes(self._h, n_seq, T))
