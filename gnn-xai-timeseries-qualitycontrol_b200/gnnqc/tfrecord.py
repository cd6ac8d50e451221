"""Readers for the reference's TFRecord window files (gnnqc_tfrecord_read_cml / _soilnet)."""
import ctypes as C

import numpy as np
import torch

from ._lib import check, lib


def read_cml_tfrecord(path, normalization="rolling_median"):
    """One file -> (batch dict in the offsets layout with torch CPU tensors, labels, ids, dates).
    Mirrors dataset.map(parse_cml_tfrecord_fn).batch(n).map(prepare_batch_cml) of the reference."""
    h = C.c_void_p()
    check(lib.gnnqc_tfrecord_read_cml(str(path).encode(), normalization.encode(), C.byref(h)))
    try:
        n, T, nodes, edges = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        check(lib.gnnqc_cml_records_info(h, C.byref(n), C.byref(T), C.byref(nodes), C.byref(edges)))
        n, T, nodes, edges = n.value, T.value, nodes.value, edges.value
        feat = np.empty((nodes * T, 2), np.float32)
        anom = np.empty((n, T, 2), np.float32)
        e = np.empty((edges, 2), np.int64)
        no = np.empty(n + 1, np.int32)
        eo = np.empty(n + 1, np.int64)
        lab = np.empty(n, np.int64)
        check(lib.gnnqc_cml_records_copy(h, feat.ctypes.data, anom.ctypes.data, e.ctypes.data, no.ctypes.data, eo.ctypes.data, lab.ctypes.data))
        ids = [lib.gnnqc_cml_records_id(h, i).decode() for i in range(n)]
        dates = [lib.gnnqc_cml_records_dates(h, i).decode().split(",") for i in range(n)]
    finally:
        lib.gnnqc_cml_records_free(h)
    batch = {"features": torch.from_numpy(feat), "anom": torch.from_numpy(anom), "edges": torch.from_numpy(e),
             "node_off": torch.from_numpy(no), "edge_off": torch.from_numpy(eo), "T": T, "n_nodes_total": nodes}
    return batch, torch.from_numpy(lab), ids, dates


def read_soilnet_tfrecord(path, normalization="scale_range"):
    """One file -> (batch dict in the layout of gnnqc_score_soilnet with torch CPU tensors, labels [sum N_b],
    sensor_ids [sum N_b], dates).  Mirrors dataset.map(parse_soilnet_tfrecord_fn).batch(n).map(prepare_batch_soilnet)."""
    h = C.c_void_p()
    check(lib.gnnqc_tfrecord_read_soilnet(str(path).encode(), normalization.encode(), C.byref(h)))
    try:
        n, T, nodes, edges = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        check(lib.gnnqc_soilnet_records_info(h, C.byref(n), C.byref(T), C.byref(nodes), C.byref(edges)))
        n, T, nodes, edges = n.value, T.value, nodes.value, edges.value
        feat = np.empty((nodes * T, 3), np.float32)
        e = np.empty((edges, 2), np.int64)
        no = np.empty(n + 1, np.int32)
        eo = np.empty(n + 1, np.int64)
        lab = np.empty(nodes, np.int64)
        sid = np.empty(nodes, np.int64)
        check(lib.gnnqc_soilnet_records_copy(h, feat.ctypes.data, e.ctypes.data, no.ctypes.data, eo.ctypes.data, lab.ctypes.data, sid.ctypes.data))
        dates = [lib.gnnqc_soilnet_records_dates(h, i).decode().split(",") for i in range(n)]
    finally:
        lib.gnnqc_soilnet_records_free(h)
    batch = {"features": torch.from_numpy(feat), "edges": torch.from_numpy(e), "node_off": torch.from_numpy(no),
             "edge_off": torch.from_numpy(eo), "T": T, "n_nodes_total": nodes}
    return batch, torch.from_numpy(lab), torch.from_numpy(sid), dates
