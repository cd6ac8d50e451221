"""Scoring from HOST buffers: chunked, double-buffered host->device staging on a copy stream
overlapped with the scoring kernels on a compute stream.  This is the path behind
`model.predict` for large host-resident batches and the one bench.py times as `e2e`."""
import numpy as np
import torch


def pin(t):
    t = torch.as_tensor(t)
    return t if t.is_pinned() else t.pin_memory()


class CmlHostScorer:
    """Scores CML GCN windows that live in (pinned) host memory, in the offsets layout of
    gnnqc_score_cml.  Inputs are never required to fit on the GPU at once."""

    def __init__(self, native, chunk_windows=8192):
        self.m = native
        self.dev = torch.device("cuda", native.device)
        self.chunk = int(chunk_windows)
        self.copy_stream = torch.cuda.Stream(self.dev)
        self.compute_stream = torch.cuda.Stream(self.dev)
        self.slots = [None, None]

    def _slot(self, i, n_feat_rows, n_win, n_edges, T):
        s = self.slots[i]
        if s is None or s["feat"].shape[0] < n_feat_rows or s["anom"].shape[0] < n_win or s["edges"].shape[0] < n_edges:
            cap_f = max(n_feat_rows, 0 if s is None else s["feat"].shape[0])
            cap_w = max(n_win, 0 if s is None else s["anom"].shape[0])
            cap_e = max(n_edges, 0 if s is None else s["edges"].shape[0])
            s = {
                "feat": torch.empty((cap_f, 2), dtype=torch.float32, device=self.dev),
                "anom": torch.empty((cap_w, T, 2), dtype=torch.float32, device=self.dev),
                "edges": torch.empty((cap_e, 2), dtype=torch.int64, device=self.dev),
                "node_off": torch.empty(cap_w + 1, dtype=torch.int32, device=self.dev),
                "edge_off": torch.empty(cap_w + 1, dtype=torch.int64, device=self.dev),
                "ready": torch.cuda.Event(), "free": torch.cuda.Event(),
            }
            s["free"].record(self.compute_stream)
            self.slots[i] = s
        return s

    def score(self, features, anom, edges, node_off, edge_off, out=None):
        """All arguments are CPU tensors (pinned for full PCIe speed).  Returns a pinned CPU
        float32 [B] tensor; the call returns after the last score has landed."""
        B, T = anom.shape[0], anom.shape[1]
        no = node_off.numpy().astype(np.int64)
        eo = edge_off.numpy().astype(np.int64)
        if out is None:
            out = torch.empty(B, dtype=torch.float32).pin_memory()
        h2d = 0
        n_chunks = (B + self.chunk - 1) // self.chunk
        staged = []   # small per-chunk pinned offset tensors must outlive their async copies
        for c in range(n_chunks):
            w0, w1 = c * self.chunk, min(B, (c + 1) * self.chunk)
            r0, r1 = no[w0] * T, no[w1] * T
            e0, e1 = eo[w0], eo[w1]
            s = self._slot(c % 2, r1 - r0, w1 - w0, e1 - e0, T)
            rel_no = torch.from_numpy((no[w0:w1 + 1] - no[w0]).astype(np.int32)).pin_memory()
            rel_eo = torch.from_numpy(eo[w0:w1 + 1] - eo[w0]).pin_memory()
            staged.append((rel_no, rel_eo))
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(s["free"])
                s["feat"][: r1 - r0].copy_(features[r0:r1], non_blocking=True)
                s["anom"][: w1 - w0].copy_(anom[w0:w1], non_blocking=True)
                s["edges"][: e1 - e0].copy_(edges[e0:e1], non_blocking=True)
                s["node_off"][: w1 - w0 + 1].copy_(rel_no, non_blocking=True)
                s["edge_off"][: w1 - w0 + 1].copy_(rel_eo, non_blocking=True)
                s["ready"].record(self.copy_stream)
            h2d += (r1 - r0) * 8 + (w1 - w0) * T * 8 + (e1 - e0) * 16 + (w1 - w0 + 1) * 12
            with torch.cuda.stream(self.compute_stream):
                self.compute_stream.wait_event(s["ready"])
                scores = self.m.score_cml(s["feat"][: r1 - r0], s["anom"][: w1 - w0], s["edges"][: e1 - e0],
                                          s["node_off"][: w1 - w0 + 1], s["edge_off"][: w1 - w0 + 1], int(no[w1] - no[w0]))
                out[w0:w1].copy_(scores, non_blocking=True)
                s["free"].record(self.compute_stream)
        self.compute_stream.synchronize()
        self.last_h2d_bytes = int(h2d)
        self.last_d2h_bytes = int(B * 4)
        return out


class SeqHostScorer:
    """Same for [n_seq, T, C] host sequences (baseline models)."""

    def __init__(self, native, chunk_seqs=None):
        self.m = native
        self.dev = torch.device("cuda", native.device)
        # default: one wave of the LSTM kernels (SMs x 4 window groups x 128 sequences) - the recurrences are serial
        # chains per group, so a smaller chunk takes the same kernel time and only adds passes
        self.chunk = int(chunk_seqs) if chunk_seqs else torch.cuda.get_device_properties(self.dev).multi_processor_count * 512
        self.copy_stream = torch.cuda.Stream(self.dev)
        self.compute_stream = torch.cuda.Stream(self.dev)
        self.slots = [None, None]

    def score(self, x, out=None):
        n, T, C = x.shape
        if out is None:
            out = torch.empty(n, dtype=torch.float32).pin_memory()
        for c in range((n + self.chunk - 1) // self.chunk):
            w0, w1 = c * self.chunk, min(n, (c + 1) * self.chunk)
            s = self.slots[c % 2]
            if s is None or s["x"].shape[0] < w1 - w0 or s["x"].shape[1] != T:
                s = {"x": torch.empty((max(self.chunk if n > self.chunk else n, w1 - w0), T, C), dtype=torch.float32, device=self.dev),
                     "ready": torch.cuda.Event(), "free": torch.cuda.Event()}
                s["free"].record(self.compute_stream)
                self.slots[c % 2] = s
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(s["free"])
                s["x"][: w1 - w0].copy_(x[w0:w1], non_blocking=True)
                s["ready"].record(self.copy_stream)
            with torch.cuda.stream(self.compute_stream):
                self.compute_stream.wait_event(s["ready"])
                out[w0:w1].copy_(self.m.score_seq(s["x"][: w1 - w0]), non_blocking=True)
                s["free"].record(self.compute_stream)
        self.compute_stream.synchronize()
        self.last_h2d_bytes = int(n * T * C * 4)
        self.last_d2h_bytes = int(n * 4)
        return out
