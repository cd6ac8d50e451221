"""On-device window builder (gnnqc_score_cml_series) == materialise the windows on the host the way
the reference does, then score them (bit-identical to gnnqc_score_cml, within 1e-4 of the oracle)."""
import os

import numpy as np
import pytest
import torch

import gnnqc_oracle as O
import window_builder as WB
from conftest import MODELS
from gnnqc import NativeModel, synth
from gnnqc.series import CmlSeries, neighbourhoods_from_distances

pytestmark = pytest.mark.gpu


def make_network(S=14, Tt=900, seed=0, nan_frac=0.002):
    rng = np.random.default_rng(seed)
    xy = rng.uniform(0, 45, size=(S, 2))
    dist = np.linalg.norm(xy[:, None] - xy[None], axis=-1)
    raw = 60.0 + np.cumsum(rng.normal(0, 0.05, size=(S, Tt, 2)), axis=1) + rng.normal(0, 0.3, size=(S, Tt, 2))
    raw[3, 400:] += 6.0                                   # a jump on one link
    holes = rng.random((S, Tt)) < nan_frac
    raw[holes] = np.nan
    raw[5, 200:260] = np.nan                              # a long gap on one neighbour
    raw = raw.astype(np.float32)
    rmed = WB.rolling_median(raw, 240).astype(np.float32)
    return raw, rmed, dist


def test_series_scoring_matches_materialised_windows():
    raw, rmed, dist = make_network()
    nbr_off, nbr_idx, edge_off, edges = neighbourhoods_from_distances(dist, 20.0)
    assert (np.diff(nbr_off) <= 32).all()
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    series = CmlSeries(raw, rmed, nbr_off, nbr_idx, edge_off, edges, device=0)
    try:
        rng = np.random.default_rng(1)
        targets = rng.integers(0, raw.shape[0], size=300).astype(np.int32)
        t_anom = rng.integers(100, raw.shape[1] - 30, size=300).astype(np.int32)   # some windows run off either end
        got, counts = m.score_cml_series(series, torch.tensor(targets).cuda(), torch.tensor(t_anom).cuda(), 120, 60, return_counts=True)
        got, counts = got.cpu().numpy(), counts.cpu().numpy()
        batch, kept, want_counts = WB.materialise_windows(raw, rmed, nbr_off, nbr_idx, edge_off, edges, targets, t_anom, 120, 60)
        assert kept.sum() > 100 and (~kept).sum() > 5
        assert np.array_equal(np.isnan(got), ~kept)
        assert np.array_equal(counts, want_counts)
        assert (want_counts[kept] < np.diff(nbr_off)[targets[kept]]).any()        # some neighbours were dropped for NaNs
        b = {k: torch.as_tensor(v).cuda() for k, v in batch.items() if k != "T"}
        ref = m.score_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"]).cpu().numpy()
        assert np.array_equal(got[kept], ref)                                      # same arithmetic, same order
        tb = {k: (torch.as_tensor(v) if k != "T" else v) for k, v in batch.items()}
        feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(tb), 2)
        w = O.load_weights(os.path.join(MODELS, "model_cml"), torch.float32)
        oracle = O.gcn_forward_cml(torch.tensor(feat), tb["anom"], torch.tensor(idx), torch.tensor(samp), w).reshape(-1).numpy()
        assert np.all(np.abs(got[kept] - oracle) <= 1e-4 + 1e-4 * np.abs(oracle))
        # streaming update: overwrite a range, scores of windows that do not touch it are unchanged
        new = raw[:, 700:760].copy()
        new[:, :, 0] += 1.0
        series.update(700, new, rmed[:, 700:760])
        again = m.score_cml_series(series, torch.tensor(targets).cuda(), torch.tensor(t_anom).cuda(), 120, 60).cpu().numpy()
        untouched = kept & (t_anom + 60 < 700)
        assert np.array_equal(again[untouched], got[untouched])
        assert not np.array_equal(again[kept & (t_anom + 60 >= 700) & (t_anom - 120 < 760)], got[kept & (t_anom + 60 >= 700) & (t_anom - 120 < 760)])
    finally:
        series.close()
        m.close()


def test_soilnet_series_scoring_matches_materialised_windows():
    """gnnqc_score_soilnet_series == build the windows on the host as the reference does, then gnnqc_score_soilnet."""
    from gnnqc.series import SoilSeries, soilnet_adjacency
    rng = np.random.default_rng(3)
    S, Tt, before, after = 23, 700, 240, 96                           # T = 337 samples, as model_soilnet expects
    xy = rng.uniform(0, 60, size=(S // 2 + 1, 2))
    place = rng.integers(0, len(xy), size=S)                           # several sensors (depths) per location
    depth = rng.choice([5.0, 20.0, 50.0], size=S)
    dist = np.linalg.norm(xy[place][:, None] - xy[place][None], axis=-1)
    adjacency = soilnet_adjacency(dist, np.abs(depth[:, None] - depth[None]), 25.0, 45.0)
    raw = np.stack([rng.uniform(5, 55, size=(S, Tt)), rng.uniform(-10, 35, size=(S, Tt)), rng.uniform(2900, 3500, size=(S, Tt))], axis=-1)
    raw = raw.astype(np.float32)
    raw[rng.random((S, Tt)) < 0.0015] = np.nan                          # sporadic gaps
    raw[4, 300:420, 1] = np.nan                                        # a long gap in one feature of one sensor
    target = (rng.random((S, Tt)) < 0.1).astype(np.float32)
    target[rng.random((S, Tt)) < 0.02] = np.nan                        # unlabelled samples
    target[7, :] = np.nan                                              # a sensor that never has a label
    t_anom = np.concatenate([rng.integers(before, Tt - after, size=40), [before - 1, Tt - after, before, Tt - after - 1]]).astype(np.int32)
    m = NativeModel.load(os.path.join(MODELS, "model_soilnet"), device=0)
    series = SoilSeries(raw, target, adjacency, device=0)
    try:
        scores, valid = series.score(m, torch.from_numpy(t_anom).cuda(), before, after)
        scores, valid = scores.cpu().numpy(), valid.cpu().numpy().astype(bool)
        batch, kept = WB.materialise_soilnet_windows(raw, target, adjacency, t_anom, before, after)
        b = {k: torch.from_numpy(v).cuda() for k, v in batch.items()}
        want = m.score_soilnet(b["features"], b["edges"], b["node_off"], b["edge_off"], before + after + 1).cpu().numpy()
        no = batch["node_off"]
        dropped_some = False
        for i, ids in enumerate(kept):
            if ids is None:                                            # window does not fit the series
                assert not valid[i].any() and np.isnan(scores[i]).all()
                continue
            mask = np.zeros(S, bool)
            mask[ids] = True
            dropped_some |= not mask.all()
            np.testing.assert_array_equal(valid[i], mask)
            np.testing.assert_array_equal(scores[i][ids], want[no[i]:no[i + 1]])   # same arithmetic, same order: bit-identical
            assert np.isnan(scores[i][~mask]).all()
        assert dropped_some and sum(k is None for k in kept) == 2
    finally:
        series.close()
        m.close()


def test_series_integrated_gradients_equal_materialised_windows():
    """gnnqc_ig_cml_series (windows named by (sensor, time) on the resident network) == materialise the windows on the
    host, then gnnqc_ig_cml: same arithmetic, bit for bit; skipped windows get NaN and zero attributions."""
    raw, rmed, dist = make_network()
    nbr_off, nbr_idx, edge_off, edges = neighbourhoods_from_distances(dist, 20.0)
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    series = CmlSeries(raw, rmed, nbr_off, nbr_idx, edge_off, edges, device=0)
    try:
        rng = np.random.default_rng(2)
        n, m_steps, T = 150, 4, 181
        targets = rng.integers(0, raw.shape[0], size=n).astype(np.int32)
        t_anom = rng.integers(100, raw.shape[1] - 30, size=n).astype(np.int32)
        an, aa, p, sens, cnt = m.ig_cml_series(series, torch.tensor(targets).cuda(), torch.tensor(t_anom).cuda(), 120, 60, m_steps=m_steps)
        an, aa, p, sens, cnt = an.cpu().numpy(), aa.cpu().numpy(), p.cpu().numpy(), sens.cpu().numpy(), cnt.cpu().numpy()
        batch, kept, want_counts = WB.materialise_windows(raw, rmed, nbr_off, nbr_idx, edge_off, edges, targets, t_anom, 120, 60)
        assert kept.sum() > 50 and (~kept).sum() > 3
        assert np.array_equal(cnt, want_counts) and np.array_equal(np.isnan(p), ~kept)
        assert not an[~kept].any() and not aa[~kept].any()
        b = {k: torch.as_tensor(v).cuda() for k, v in batch.items() if k != "T"}
        rf, ra, rp = m.ig_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], m_steps=m_steps)
        rf, ra, rp = rf.cpu().numpy(), ra.cpu().numpy(), rp.cpu().numpy()
        assert np.array_equal(p[kept], rp) and np.array_equal(aa[kept], ra)
        no = batch["node_off"]
        for i, w in enumerate(np.nonzero(kept)[0]):
            nn = int(no[i + 1] - no[i])
            want = rf[no[i] * T: no[i + 1] * T].reshape(T, nn, 2)
            assert np.array_equal(an[w, :, :nn], want), "window %d" % w
            assert not an[w, :, nn:].any() and (sens[w, nn:] == -1).all() and (sens[w, :nn] >= 0).all()
        assert np.abs(an).max() > 0
    finally:
        series.close()
        m.close()


def test_link_sharded_series_scores_equal_the_whole_network():
    """A rank that holds only its links plus the neighbour halo scores its targets exactly as the whole network does."""
    from gnnqc.series import shard_network_by_link
    net = synth.make_cml_network(48, 400, seed=9, mean_neighbours=7)
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    full = CmlSeries(net["raw"], net["rmed"], net["nbr_off"], net["nbr_idx"], net["edge_off"], net["edges"], device=0)
    try:
        t = torch.full((48,), 250, dtype=torch.int32).cuda()
        want = m.score_cml_series(full, torch.arange(48, dtype=torch.int32).cuda(), t, 120, 60).cpu().numpy()
        got = np.full(48, np.nan, dtype=np.float32)
        for rank in range(3):
            sh = shard_network_by_link(net, rank, 3)
            part = CmlSeries(sh["raw"], sh["rmed"], sh["nbr_off"], sh["nbr_idx"], sh["edge_off"], sh["edges"], device=0)
            try:
                tgt = torch.tensor(sh["owned_local"]).cuda()
                got[sh["owned_global"]] = m.score_cml_series(part, tgt, t[: tgt.numel()], 120, 60).cpu().numpy()
            finally:
                part.close()
        assert np.array_equal(got, want, equal_nan=True) and np.isfinite(want).sum() > 30
    finally:
        full.close()
        m.close()
