"""Batch layouts either side of the C ABI.

The reference hands its model a *disjoint-graph* batch (libs/preprocessing_functions.py:638-666):
features [sum_b T*N_b, F] in (window, timestep, node) order, one block-diagonal COO adjacency in
which every window's edge list is repeated for each of its T timesteps, and sample_indices
[rows] -> window.  libgnnqc takes the same features plus the *untiled* per-window edge list and
prefix offsets.  `disjoint_to_offsets` converts the former into the latter with torch ops on
whatever device the tensors live on (no host round trip except two scalars).
"""
import torch


def adjacency_indices(adjacency):
    """Accepts a torch sparse COO tensor, an (indices, values, dense_shape) triple as
    tf.sparse.SparseTensor exposes it, or a plain [E,2] index tensor; returns int64 [E,2]."""
    if isinstance(adjacency, torch.Tensor) and adjacency.is_sparse:
        return adjacency.coalesce().indices().t().contiguous() if not adjacency.is_coalesced() else adjacency.indices().t().contiguous()
    if isinstance(adjacency, (tuple, list)):
        adjacency = adjacency[0]
    idx = torch.as_tensor(adjacency)
    if idx.dim() != 2 or idx.shape[1] != 2:
        raise ValueError("adjacency indices must be [E,2], got %s" % (tuple(idx.shape),))
    return idx.to(torch.int64)


def _rows_per_window(sample_indices, n_windows):
    """Counts without torch.bincount (which reads the maximum back to the host to size its output)."""
    sample_indices = sample_indices.to(torch.int64)
    return torch.zeros(n_windows, dtype=torch.int64, device=sample_indices.device).scatter_add_(0, sample_indices, torch.ones_like(sample_indices))


def node_offsets(sample_indices, n_windows, T):
    """sample_indices [rows] -> node_off i32 [n_windows + 1]; windows past the last one present stay empty."""
    nodes = _rows_per_window(sample_indices, n_windows) // T
    node_off = torch.zeros(n_windows + 1, dtype=torch.int64, device=sample_indices.device)
    node_off[1:] = torch.cumsum(nodes, 0)
    return node_off.to(torch.int32)


def disjoint_to_offsets(indices, sample_indices, n_windows, T, validate=False):
    """indices int64 [E_tiled,2], sample_indices [rows] -> (edges [E,2] local, node_off i32
    [B+1], edge_off i64 [B+1], n_nodes_total).  Runs on the tensors' device without reading anything back
    (sizes follow from the shapes: rows / T nodes, E_tiled / T edges); validate=True adds the one check that needs a
    host read (rows per window divisible by T)."""
    sample_indices = sample_indices.to(torch.int64)
    rows_per_window = _rows_per_window(sample_indices, n_windows)
    if validate and int((rows_per_window % T).abs().sum()) != 0:
        raise ValueError("rows per window are not a multiple of the sequence length %d" % T)
    if sample_indices.numel() % T or indices.shape[0] % T:
        raise ValueError("rows / tiled edges are not a multiple of the sequence length %d" % T)
    nodes = rows_per_window // T
    node_off = torch.zeros(n_windows + 1, dtype=torch.int64, device=indices.device)
    node_off[1:] = torch.cumsum(nodes, 0)
    base = node_off[:-1] * T                                   # first row of each window
    win = sample_indices[indices[:, 0]]                        # window of every tiled edge
    local_row = indices[:, 0] - base[win]
    first_step = local_row < nodes[win]                        # the t == 0 copy of the edge list
    n_edges = indices.shape[0] // T
    # compaction without a dynamic-size op: stable sort by (not first step, window), keep the first E entries
    key = (~first_step).to(torch.int64) * (n_windows + 1) + win
    order = torch.argsort(key, stable=True)[:n_edges]
    win0 = win[order]
    edges = torch.stack([local_row[order], indices[order, 1] - base[win0]], dim=1).contiguous()
    edge_off = torch.zeros(n_windows + 1, dtype=torch.int64, device=indices.device)
    edge_off[1:] = torch.cumsum(_rows_per_window(win0, n_windows), 0)
    return edges, node_off.to(torch.int32), edge_off, sample_indices.numel() // T


def offsets_to_disjoint(edges, node_off, edge_off, T):
    """Inverse of disjoint_to_offsets (what prepare_batch_cml builds): tiled joint adjacency
    indices and sample_indices."""
    node_off = node_off.to(torch.int64)
    nodes = node_off[1:] - node_off[:-1]
    n_edges = edge_off[1:] - edge_off[:-1]
    B = nodes.numel()
    win = torch.repeat_interleave(torch.arange(B, device=edges.device), n_edges)
    base = node_off[:-1] * T
    t = torch.arange(T, device=edges.device)
    # (window, t, edge) order
    out = []
    for b in range(B):
        e = edges[edge_off[b]: edge_off[b + 1]]
        out.append((e[None, :, :] + (base[b] + t * nodes[b])[:, None, None]).reshape(-1, 2))
    sample_indices = torch.repeat_interleave(torch.arange(B, device=edges.device), nodes * T)
    del win
    return torch.cat(out), sample_indices
