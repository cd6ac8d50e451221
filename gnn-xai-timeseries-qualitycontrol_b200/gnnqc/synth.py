"""Seeded synthetic batches with the shapes and statistics of the reference's example data
(SURVEY.md section 8(d)); the real example_data/*.nc files are not available.

CML: T=181 steps of 2 channels (TRSL1/2 minus rolling median, dB), a target link plus the links
within 20 km of it, edges between links closer than 20 km incl. self-loops
(reference: libs/preprocessing_functions.py:346,408; N_b <= 23 in the example data).
SoilNet: T=337 steps of 3 scale_range-normalised channels (moisture, temperature, battery),
sensors in boxes x depths, edges (d <= 20 m & same depth) | (d == 0 & |ddepth| <= 0.1)
(libs/preprocessing_functions.py:475-478).

Everything is generated with torch on the requested device so large bench batches do not
have to cross PCIe; results are deterministic per (seed, device type).
"""
import math

import torch


def _ar1(noise, rho):
    """AR(1) along dim 1: y_t = rho*y_{t-1} + e_t via the closed form (T is a few hundred)."""
    T = noise.shape[1]
    t = torch.arange(T, device=noise.device, dtype=torch.float64)
    shape = [1, T] + [1] * (noise.dim() - 2)
    up = torch.pow(torch.tensor(rho, dtype=torch.float64, device=noise.device), -t).reshape(shape)
    down = torch.pow(torch.tensor(rho, dtype=torch.float64, device=noise.device), t).reshape(shape)
    return (torch.cumsum(noise.double() * up, dim=1) * down).float()


def make_cml_batch(n_windows, seed=44, n_nodes=23, ragged=False, T=181, device="cpu", min_nodes=4):
    """Returns a dict in the layout gnnqc_score_cml takes (see include/gnnqc.h):
    features [sum N_b*T, 2] rows (window, t, node), anom [B,T,2], edges [sum E_b, 2] (row, col)
    local to each window, node_off i32 [B+1], edge_off i64 [B+1], labels i64 [B],
    target_index i64 [B]."""
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    B, N = int(n_windows), int(n_nodes)
    dev = torch.device(device)
    if ragged:
        counts = torch.randint(min_nodes, N + 1, (B,), generator=g, device=dev)
    else:
        counts = torch.full((B,), N, dtype=torch.int64, device=dev)
    node_mask = torch.arange(N, device=dev)[None, :] < counts[:, None]            # [B,N]

    # Geometry: target at the origin, neighbours uniform in the 20 km disc around it.
    r = 20.0 * torch.sqrt(torch.rand((B, N), generator=g, device=dev))
    phi = 2 * math.pi * torch.rand((B, N), generator=g, device=dev)
    xy = torch.stack([r * torch.cos(phi), r * torch.sin(phi)], dim=-1)
    target = (torch.rand((B,), generator=g, device=dev) * counts).long().clamp_(max=N - 1)
    xy[torch.arange(B, device=dev), target] = 0.0
    dist = torch.cdist(xy, xy)
    adj = (dist < 20.0) & node_mask[:, :, None] & node_mask[:, None, :]
    bij = adj.nonzero()                                                           # sorted (b, i, j)
    edges = bij[:, 1:].contiguous()
    edge_counts = torch.bincount(bij[:, 0], minlength=B)

    # Signals.
    x = _ar1(0.5 * math.sqrt(1 - 0.95 ** 2) * torch.randn((B, T, N, 2), generator=g, device=dev), 0.95)
    rain = torch.rand((B,), generator=g, device=dev) < 0.10
    amp = 1.0 + 7.0 * torch.rand((B,), generator=g, device=dev)
    centre = torch.randint(min(20, T // 4), max(T - 20, T // 4 + 1), (B,), generator=g, device=dev)
    width = 5.0 + 25.0 * torch.rand((B,), generator=g, device=dev)
    tt = torch.arange(T, device=dev, dtype=torch.float32)[None, :]
    bump = amp[:, None] * torch.exp(-0.5 * ((tt - centre[:, None]) / width[:, None]) ** 2)
    x = x + (rain[:, None] * bump)[:, :, None, None]
    anomalous = torch.rand((B,), generator=g, device=dev) < 0.05
    jump_at = torch.randint(T // 2, max(T // 2 + 1, (5 * T) // 6), (B,), generator=g, device=dev)
    jump = (2.0 + 8.0 * torch.rand((B,), generator=g, device=dev))[:, None] * (tt >= jump_at[:, None])
    sel = torch.arange(B, device=dev)
    x[sel, :, target] = x[sel, :, target] + (anomalous[:, None] * jump)[:, :, None]

    anom = x[sel, :, target].contiguous()                                         # [B,T,2]
    if ragged:
        features = x[node_mask[:, None, :].expand(B, T, N)].reshape(-1, 2).contiguous()
    else:
        features = x.reshape(-1, 2)
    node_off = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    node_off[1:] = torch.cumsum(counts, 0).to(torch.int32)
    edge_off = torch.zeros(B + 1, dtype=torch.int64, device=dev)
    edge_off[1:] = torch.cumsum(edge_counts, 0)
    return {
        "features": features, "anom": anom, "edges": edges, "node_off": node_off, "edge_off": edge_off,
        "labels": anomalous.long(), "target_index": target, "T": T,
    }


def cml_batch_to_windows(batch):
    """Per-window view (numpy) used by tests to build the reference's disjoint-graph layout."""
    T = batch["T"]
    f = batch["features"].cpu().numpy()
    e = batch["edges"].cpu().numpy()
    no = batch["node_off"].cpu().numpy()
    eo = batch["edge_off"].cpu().numpy()
    out = []
    for b in range(len(no) - 1):
        n = int(no[b + 1] - no[b])
        out.append({"x": f[no[b] * T: no[b + 1] * T].reshape(T, n, f.shape[1]), "edges": e[eo[b]: eo[b + 1]]})
    return out


def make_soilnet_batch(n_windows, seed=44, n_nodes=210, ragged=False, T=337, device="cpu", min_nodes=20):
    """SoilNet-shaped batch in the layout gnnqc_score_soilnet takes: features [sum N_b*T, 3] rows
    (window, t, node), edges [sum E_b, 2] window-local (row = receiving node), node_off, edge_off.
    Sensors sit in boxes of up to 6 depths; edges follow libs/preprocessing_functions.py:475-478:
    (distance <= 20 m and same depth) or (same box and depth difference <= 0.1 m)."""
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    B, N = int(n_windows), int(n_nodes)
    dev = torch.device(device)
    if ragged:
        counts = torch.randint(min(min_nodes, N), N + 1, (B,), generator=g, device=dev)
    else:
        counts = torch.full((B,), N, dtype=torch.int64, device=dev)
    node_mask = torch.arange(N, device=dev)[None, :] < counts[:, None]
    depths = torch.tensor([0.05, 0.1, 0.2, 0.3, 0.4, 0.5], device=dev)
    idx = torch.arange(N, device=dev)
    box, depth = idx // 6, depths[idx % 6]
    n_box = (N + 5) // 6
    box_xy = 120.0 * torch.rand((B, n_box, 2), generator=g, device=dev)
    xy = box_xy[:, box]                                                 # [B,N,2]
    dist = torch.cdist(xy, xy)
    same_depth = (depth[:, None] == depth[None, :])[None]
    same_box = (box[:, None] == box[None, :])[None]
    near_depth = ((depth[:, None] - depth[None, :]).abs() <= 0.1 + 1e-6)[None]
    adj = ((dist <= 20.0) & same_depth) | (same_box & near_depth)
    adj = adj & node_mask[:, :, None] & node_mask[:, None, :]
    bij = adj.nonzero()
    edges = bij[:, 1:].contiguous()
    edge_counts = torch.bincount(bij[:, 0], minlength=B)

    t = torch.arange(T, device=dev, dtype=torch.float32)[None, :, None]
    moist = 0.45 + 0.1 * torch.randn((B, 1, N), generator=g, device=dev) + _ar1(0.01 * torch.randn((B, T, N), generator=g, device=dev), 0.99)
    temp = 0.5 + 0.1 * torch.sin(2 * math.pi * t / 96.0 + 6.28 * torch.rand((B, 1, 1), generator=g, device=dev)) \
        + 0.01 * torch.randn((B, T, N), generator=g, device=dev)
    batt = 0.8 - 0.2 * torch.rand((B, 1, N), generator=g, device=dev) * (t / T)
    x = torch.stack([moist.clamp(0.1, 0.8), temp, batt.expand(B, T, N)], dim=-1)      # [B,T,N,3]
    # a few sensors with a step anomaly in the moisture channel
    anomalous = torch.rand((B, N), generator=g, device=dev) < 0.03
    step = (t[:, :, 0] >= (T * 0.6)).float()[:, :, None] * 0.25
    x[..., 0] = x[..., 0] + anomalous[:, None, :] * step
    if ragged:
        features = x[node_mask[:, None, :].expand(B, T, N)].reshape(-1, 3).contiguous()
        labels = anomalous[node_mask]
    else:
        features = x.reshape(-1, 3)
        labels = anomalous.reshape(-1)
    node_off = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    node_off[1:] = torch.cumsum(counts, 0).to(torch.int32)
    edge_off = torch.zeros(B + 1, dtype=torch.int64, device=dev)
    edge_off[1:] = torch.cumsum(edge_counts, 0)
    return {"features": features, "edges": edges, "node_off": node_off, "edge_off": edge_off, "labels": labels.long(), "T": T}


def make_cml_network(n_sensors, n_times, seed=42, mean_neighbours=20, max_neighbours=32, nan_frac=0.0005):
    """Synthetic raw CML network for the window-builder path (SURVEY.md 8(d) config 4 shape, scaled):
    returns numpy raw [S,Tt,2] (dB levels with NaN gaps), rolling medians [S,Tt,2] (here: a slow
    moving baseline, exact rolling medians are an offline product in the reference), and the
    per-target neighbourhoods / adjacency for a 20 km radius (nearest `max_neighbours` kept)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    side = math.sqrt(n_sensors * math.pi * 400.0 / mean_neighbours)
    xy = rng.uniform(0, side, size=(n_sensors, 2)).astype(np.float32)
    base = rng.uniform(40, 80, size=(n_sensors, 1, 2)).astype(np.float32)
    drift = np.cumsum(rng.normal(0, 0.01, size=(n_sensors, n_times, 2)).astype(np.float32), axis=1)
    rmed = base + drift
    raw = rmed + rng.normal(0, 0.3, size=(n_sensors, n_times, 2)).astype(np.float32)
    raw[rng.random((n_sensors, n_times)) < nan_frac] = np.nan
    nbr_off, nbr_idx, edge_off, edges = [0], [], [0], []
    for s0 in range(0, n_sensors, 256):          # blocked distance computation
        d = np.linalg.norm(xy[s0:s0 + 256, None, :] - xy[None, :, :], axis=-1)
        for k in range(d.shape[0]):
            nb = np.nonzero(d[k] < 20.0)[0]
            if len(nb) > max_neighbours:
                nb = np.sort(nb[np.argsort(d[k][nb], kind="stable")[:max_neighbours]])
            sub = np.linalg.norm(xy[nb, None, :] - xy[None, nb, :], axis=-1) < 20.0
            r, c = np.nonzero(sub)
            nbr_idx.append(nb)
            nbr_off.append(nbr_off[-1] + len(nb))
            edges.append(np.stack([r, c], axis=1))
            edge_off.append(edge_off[-1] + len(r))
    return {"raw": raw, "rmed": rmed.astype(np.float32), "nbr_off": np.asarray(nbr_off, np.int32),
            "nbr_idx": np.concatenate(nbr_idx).astype(np.int32), "edge_off": np.asarray(edge_off, np.int64),
            "edges": np.concatenate(edges).astype(np.int64).reshape(-1, 2)}


def make_seq_batch(n_seq, n_features, T, seed=44, device="cpu"):
    """Plain [n_seq, T, F] sequences for the baseline models (CML: the target series)."""
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    dev = torch.device(device)
    if n_features == 2:
        b = make_cml_batch(n_seq, seed=seed, n_nodes=4, T=T, device=device)
        return b["anom"]
    # SoilNet-like channels: moisture/60 in [0.1,0.8], (temp+20)/60 diurnal, battery in [0,1].
    t = torch.arange(T, device=dev, dtype=torch.float32)[None, :]
    moist = 0.45 + _ar1(0.02 * torch.randn((n_seq, T), generator=g, device=dev), 0.99)
    temp = 0.5 + 0.1 * torch.sin(2 * math.pi * t / 96.0 + 6.28 * torch.rand((n_seq, 1), generator=g, device=dev))
    batt = 0.8 - 0.2 * torch.rand((n_seq, 1), generator=g, device=dev) * (t / T)
    return torch.stack([moist.clamp(0.1, 0.8), temp, batt.expand(n_seq, T)], dim=-1).contiguous()
