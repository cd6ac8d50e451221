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

    def update(self, t0, raw_slab, rmed_slab):
        raw_slab = np.ascontiguousarray(raw_slab, dtype=np.float32)
        rmed_slab = np.ascontiguousarray(rmed_slab, dtype=np.float32)
        check(lib.gnnqc_series_update(self._h, int(t0), raw_slab.shape[1], raw_slab.ctypes.data, rmed_slab.ctypes.data))

    def close(self):
        if self._h:
            lib.gnnqc_series_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
