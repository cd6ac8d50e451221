"""CPU ORACLE (test infrastructure only): the reference's window materialisation restated in NumPy.

Follows libs/preprocessing_functions.py:388-412 (window slicing, NaN rules, adjacency among the
sensors that remain), :259-262 and :620-622 (rolling median at the processed date subtracted from
every sensor's series) and :638-666 (row order).  Produces the offsets-layout batch that
gnnqc_score_cml / the forward oracle take, so the on-device window builder can be checked against
"materialise on the host, then score"."""
import numpy as np


def rolling_median(raw, window_length):
    """xarray's .rolling(time=window_length, min_periods=1).median(): trailing window, NaNs ignored."""
    import pandas as pd
    S, Tt, F = raw.shape
    out = np.empty_like(raw)
    for s in range(S):
        out[s] = pd.DataFrame(raw[s]).rolling(window_length, min_periods=1).median().to_numpy()
    return out


def materialise_windows(raw, rmed, nbr_off, nbr_idx, edge_off, edges, targets, t_anom, before, after):
    """Returns (batch dict in offsets layout for the kept windows, kept mask [B], node counts [B])."""
    T = before + after + 1
    feats, anoms, all_edges, node_off, eoff, kept, counts = [], [], [], [0], [0], [], []
    for s, tau in zip(targets, t_anom):
        t0, t1 = tau - before, tau + after
        ok = 0 <= s < raw.shape[0] and t0 >= 0 and t1 < raw.shape[1]
        if ok:
            nb = nbr_idx[nbr_off[s]: nbr_off[s + 1]]
            win = raw[nb, t0: t1 + 1]                                    # [N, T, 2]
            tgt_pos = int(np.nonzero(nb == s)[0][0])
            ok = not np.isnan(win[tgt_pos]).any()
        if not ok:
            kept.append(False)
            counts.append(0)
            continue
        valid = ~np.isnan(win).any(axis=(1, 2))
        slot = np.cumsum(valid) - 1
        e = edges[edge_off[s]: edge_off[s + 1]]
        e = e[valid[e[:, 0]] & valid[e[:, 1]]]
        e = np.stack([slot[e[:, 0]], slot[e[:, 1]]], axis=1)
        x = win[valid] - rmed[nb[valid], tau][:, None, :]                # subtract the median at the target time
        feats.append(np.transpose(x, (1, 0, 2)).reshape(-1, 2))          # rows (t, n)
        anoms.append(x[slot[tgt_pos]])
        all_edges.append(e)
        node_off.append(node_off[-1] + int(valid.sum()))
        eoff.append(eoff[-1] + len(e))
        kept.append(True)
        counts.append(int(valid.sum()))
    batch = None
    if feats:
        batch = {"features": np.concatenate(feats).astype(np.float32), "anom": np.stack(anoms).astype(np.float32),
                 "edges": np.concatenate(all_edges).astype(np.int64).reshape(-1, 2), "node_off": np.asarray(node_off, np.int32),
                 "edge_off": np.asarray(eoff, np.int64), "T": T}
    return batch, np.asarray(kept), np.asarray(counts)


def materialise_soilnet_windows(raw, target, adjacency, t_anom, before, after, sub=(0.0, -20.0, 2800.0), div=(60.0, 60.0, 800.0)):
    """SoilNet windows as create_tfrecords_dataset builds them (reference libs/preprocessing_functions.py:449-480)
    and parse_soilnet_tfrecord_fn normalises them (:825-832, 'scale_range' by default), in the offsets layout of
    gnnqc_score_soilnet.  raw [S, Tt, 3], target [S, Tt] (NaN = no label), adjacency [S, S] bool.
    Returns (batch dict, kept): kept[i] = global sensor ids of window i (None for a window that does not fit)."""
    raw = np.asarray(raw, np.float32)
    S, Tt, _ = raw.shape
    sub, div = np.asarray(sub, np.float32), np.asarray(div, np.float32)
    feats, edges, node_off, edge_off, kept = [], [], [0], [0], []
    for ta in t_anom:
        t0, t1 = int(ta) - before, int(ta) + after
        if t0 < 0 or t1 >= Tt:
            kept.append(None)
            node_off.append(node_off[-1])
            edge_off.append(edge_off[-1])
            continue
        win = raw[:, t0:t1 + 1]                                            # [S, T, 3]
        ok = ~np.isnan(target[:, int(ta)]) & ~np.isnan(win).any(axis=(1, 2))
        ids = np.nonzero(ok)[0]
        kept.append(ids)
        x = (win[ids] - sub) / div                                         # float32 arithmetic, like the tf ops
        feats.append(np.transpose(x, (1, 0, 2)).reshape(-1, 3))            # rows in (timestep, node) order
        r, c = np.nonzero(np.asarray(adjacency, bool)[np.ix_(ids, ids)])
        edges.append(np.stack([r, c], axis=1))
        node_off.append(node_off[-1] + len(ids))
        edge_off.append(edge_off[-1] + len(r))
    batch = {"features": np.concatenate(feats).astype(np.float32) if feats else np.zeros((0, 3), np.float32),
             "edges": np.concatenate(edges).astype(np.int64).reshape(-1, 2) if edges else np.zeros((0, 2), np.int64),
             "node_off": np.asarray(node_off, np.int32), "edge_off": np.asarray(edge_off, np.int64)}
    return batch, kept
