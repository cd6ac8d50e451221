"""Data-parallel plumbing for the training step: batch sharding and the two all-reduces.

One process per GPU (torch.distributed, NCCL over NVLink); on CPU the same code runs over gloo
(tests/test_dp_gloo.py).  Scoring and attribution shard by window with no collective at all; the
training step has exactly two exchange points (gnnqc.h, "training").
"""
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n, rank, world_size):
    """Contiguous, balanced [lo, hi) of n items for `rank`."""
    base, rem = divmod(n, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_cml_batch(batch, labels, rank, world_size):
    """Slice an offsets-layout CML batch (see include/gnnqc.h) to this rank's windows."""
    B = batch["anom"].shape[0]
    T = batch["anom"].shape[1]
    lo, hi = shard_range(B, rank, world_size)
    no, eo = batch["node_off"], batch["edge_off"]
    n0, n1 = int(no[lo]), int(no[hi])
    e0, e1 = int(eo[lo]), int(eo[hi])
    sub = {
        "features": batch["features"][n0 * T: n1 * T],
        "anom": batch["anom"][lo:hi],
        "edges": batch["edges"][e0:e1],
        "node_off": (no[lo: hi + 1] - no[lo]).contiguous(),
        "edge_off": (eo[lo: hi + 1] - eo[lo]).contiguous(),
        "n_nodes_total": n1 - n0,
    }
    return sub, labels[lo:hi]


def shard_soil_batch(batch, labels, rank, world_size):
    """Slice an offsets-layout SoilNet batch to this rank's WINDOWS (a window's sensors share one graph, so windows
    stay whole); labels are per (window, sensor)."""
    no, eo = batch["node_off"], batch["edge_off"]
    B = no.numel() - 1
    T = int(batch["T"])
    lo, hi = shard_range(B, rank, world_size)
    n0, n1 = int(no[lo]), int(no[hi])
    e0, e1 = int(eo[lo]), int(eo[hi])
    sub = {
        "features": batch["features"][n0 * T: n1 * T],
        "edges": batch["edges"][e0:e1],
        "node_off": (no[lo: hi + 1] - no[lo]).contiguous(),
        "edge_off": (eo[lo: hi + 1] - eo[lo]).contiguous(),
        "n_nodes_total": n1 - n0,
        "T": T,
    }
    return sub, labels[n0:n1]


def shard_seq_batch(x, labels, rank, world_size):
    lo, hi = shard_range(x.shape[0], rank, world_size)
    return x[lo:hi], labels[lo:hi]


def all_reduce_sum(t):
    """In-place sum over ranks (no-op for a single process)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def global_count(n, device=None):
    t = torch.tensor([n], dtype=torch.int64, device=device)
    all_reduce_sum(t)
    return int(t.item())
