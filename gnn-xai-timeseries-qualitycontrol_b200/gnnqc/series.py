"""Resident raw sensor network + sliding-window scoring (gnnqc_series_* in include/gnnqc.h)."""
import ctypes as C

import numpy as np

from ._lib import check, lib


def neighbourhoods_from_distances(dist, max_distance):
    """Per-target neighbour lists and adjacency, as the reference derives them from the distance
    matrix (get_neighbors, adjacency = distances < max_sample_distance incl. the zero diagonal,
    libs/preprocessing_functions.py:346,408).  Returns (nbr_off, nbr_idx, edge_off, edges)."""
    dist = np.asarray(dist)
    S = dist.shape[0]
    nbr_off, nbr_idx, edge_off, edges = [0], [], [0], []
    for s in range(S):
        nb = np.nonzero(dist[s] < max_distance)[0]          # includes s itself (distance 0)
        sub = dist[np.ix_(nb, nb)] < max_distance
        r, c = np.nonzero(sub)
        nbr_idx.append(nb)
        nbr_off.append(nbr_off[-1] + len(nb))
        edges.append(np.stack([r, c], axis=1))
        edge_off.append(edge_off[-1] + len(r))
    return (np.asarray(nbr_off, np.int32), np.concatenate(nbr_idx).astype(np.int32), np.asarray(edge_off, np.int64),
            np.concatenate(edges).astype(np.int64).reshape(-1, 2))


def shard_network_by_link(net, rank, world_size):
    """Shards a network (dict with raw, rmed [S,Tt,2], nbr_off, nbr_idx, edge_off, edges as CmlSeries takes them) by
    TARGET link for one rank of a multi-GPU job (scoring / IG shard by link with no collective, the
    reference's analogue is the SLURM file split, xai/libs/integrated_gradients.py:432-448).  The rank keeps the series of
    its own links plus the halo - every link that appears in one of their neighbourhoods - renumbered densely; halo links
    get an empty neighbourhood (they are never targets here).  Returns the sharded dict plus
      'owned_global'  global ids of this rank's target links (contiguous range)
      'owned_local'   their row in the sharded arrays
      'global_ids'    global id of every local row (targets + halo)"""
    nbr_off, nbr_idx = np.asarray(net["nbr_off"]), np.asarray(net["nbr_idx"])
    edge_off, edges = np.asarray(net["edge_off"]), np.asarray(net["edges"]).reshape(-1, 2)
    S = nbr_off.size - 1
    base, rem = divmod(S, world_size)
    lo = rank * base + min(rank, rem)
    hi = lo + base + (1 if rank < rem else 0)
    owned = np.arange(lo, hi)
    needed = np.unique(np.concatenate([nbr_idx[nbr_off[lo]: nbr_off[hi]], owned])).astype(np.int64)
    remap = np.full(S, -1, dtype=np.int64)
    remap[needed] = np.arange(needed.size)
    is_owned = (needed >= lo) & (needed < hi)
    n_nbr = np.where(is_owned, nbr_off[needed + 1] - nbr_off[needed], 0)
    n_edge = np.where(is_owned, edge_off[needed + 1] - edge_off[needed], 0)
    new_nbr_off = np.concatenate([[0], np.cumsum(n_nbr)]).astype(np.int32)
    new_edge_off = np.concatenate([[0], np.cumsum(n_edge)]).astype(np.int64)
    # the owned range is contiguous in the global arrays, and `needed` is sorted, so its owned rows keep their order
    new_nbr_idx = remap[nbr_idx[nbr_off[lo]: nbr_off[hi]]].astype(np.int32)
    new_edges = edges[edge_off[lo]: edge_off[hi]].astype(np.int64)
    out = {"raw": np.ascontiguousarray(np.asarray(net["raw"])[needed]), "rmed": np.ascontiguousarray(np.asarray(net["rmed"])[needed]),
           "nbr_off": new_nbr_off, "nbr_idx": new_nbr_idx, "edge_off": new_edge_off, "edges": new_edges,
           "owned_global": owned, "owned_local": remap[owned].astype(np.int32), "global_ids": needed}
    return out


class CmlSeries:
    """raw, rmed: float32 [n_sensors, n_times, 2] (NaN = missing)."""

    def __init__(self, raw, rmed, nbr_off, nbr_idx, edge_off, edges, device=0):
        raw = np.ascontiguousarray(raw, dtype=np.float32)
        rmed = np.ascontiguousarray(rmed, dtype=np.float32)
        if raw.ndim != 3 or raw.shape[2] != 2 or rmed.shape != raw.shape:
            raise ValueError("raw and rmed must both be [n_sensors, n_times, 2]")
        self.n_sensors, self.n_times = raw.shape[0], raw.shape[1]
        self.device = int(device)
        nbr_off = np.ascontiguousarray(nbr_off, np.int32)
        nbr_idx = np.ascontiguousarray(nbr_idx, np.int32)
        edge_off = np.ascontiguousarray(edge_off, np.int64)
        edges = np.ascontiguousarray(edges, np.int64)
        h = C.c_void_p()
        check(lib.gnnqc_series_create(self.device, raw.ctypes.data, rmed.ctypes.data, self.n_sensors, self.n_times, nbr_off.ctypes.data,
                                      nbr_idx.ctypes.data, edge_off.ctypes.data, edges.ctypes.data, C.byref(h)))
        self._h = h

    def update(self, t0, raw_slab, rmed_slab, stream=None):
        """Overwrites [t0, t0 + n_t) from host slabs [n_sensors, n_t, 2] (numpy arrays, or pinned torch tensors for
        copies that do not block the host).  stream: a raw CUDA stream handle; the update is ordered on it."""
        def host(a):
            if hasattr(a, "data_ptr"):          # torch CPU tensor (pinned memory keeps the copy asynchronous)
                if a.dtype.is_floating_point and a.element_size() == 4 and a.is_contiguous():
                    return a, a.data_ptr(), a.shape[1]
                a = a.numpy()
            a = np.ascontiguousarray(a, dtype=np.float32)
            return a, a.ctypes.data, a.shape[1]
        raw_keep, raw_ptr, n_t = host(raw_slab)
        med_keep, med_ptr, _ = host(rmed_slab)
        if stream is None:
            check(lib.gnnqc_series_update(self._h, int(t0), n_t, raw_ptr, med_ptr))
        else:
            check(lib.gnnqc_series_update_async(self._h, int(t0), n_t, raw_ptr, med_ptr, C.c_void_p(int(stream))))
            self._keep = (raw_keep, med_keep)    # the host slabs must outlive the asynchronous copies

    def close(self):
        if self._h:
            lib.gnnqc_series_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def soilnet_adjacency(dist, depth_diff, max_distance, max_depth):
    """The reference's SoilNet adjacency over the whole network (libs/preprocessing_functions.py:474-477)."""
    dist, depth_diff = np.asarray(dist), np.asarray(depth_diff)
    return (((dist <= max_distance) & (depth_diff == 0)) | ((dist == 0) & (depth_diff <= max_depth))).astype(np.uint8)


class SoilSeries:
    """Resident SoilNet network: raw float32 [n_sensors, n_times, 3] (moisture, temp, battv; NaN = missing),
    target float32 [n_sensors, n_times] (NaN = no label), adjacency uint8 [n_sensors, n_sensors]."""

    def __init__(self, raw, target, adjacency, device=0):
        raw = np.ascontiguousarray(raw, dtype=np.float32)
        target = np.ascontiguousarray(target, dtype=np.float32)
        adjacency = np.ascontiguousarray(adjacency, dtype=np.uint8)
        if raw.ndim != 3 or raw.shape[2] != 3 or target.shape != raw.shape[:2] or adjacency.shape != (raw.shape[0], raw.shape[0]):
            raise ValueError("expected raw [S, Tt, 3], target [S, Tt], adjacency [S, S]")
        self.n_sensors, self.n_times, self.device = raw.shape[0], raw.shape[1], int(device)
        h = C.c_void_p()
        check(lib.gnnqc_soil_series_create(self.device, raw.ctypes.data, target.ctypes.data, self.n_sensors, self.n_times,
                                           adjacency.ctypes.data, C.byref(h)))
        self._h = h

    def score(self, model, t_anom, steps_before=240, steps_after=96, norm=None):
        """t_anom: cuda int32 [B].  Returns (scores [B, n_sensors] with NaN for dropped sensors, valid uint8 [B, n_sensors])."""
        import torch
        t_anom = t_anom.to(torch.int32).contiguous()
        B = t_anom.numel()
        scores = torch.empty((B, self.n_sensors), dtype=torch.float32, device=t_anom.device)
        valid = torch.empty((B, self.n_sensors), dtype=torch.uint8, device=t_anom.device)
        sub = div = None
        if norm is not None:
            sub = np.ascontiguousarray(norm[0], np.float32)
            div = np.ascontiguousarray(norm[1], np.float32)
        check(lib.gnnqc_score_soilnet_series(model._h, self._h, t_anom.data_ptr(), B, int(steps_before), int(steps_after),
                                             sub.ctypes.data if sub is not None else None, div.ctypes.data if div is not None else None,
                                             scores.data_ptr(), valid.data_ptr(), torch.cuda.current_stream(t_anom.device).cuda_stream))
        return scores, valid

    def close(self):
        if self._h:
            lib.gnnqc_soil_series_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
