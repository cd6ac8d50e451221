"""GPU parity: CUDA scoring path (through the C ABI) vs the CPU oracle and the golden vectors.
Tolerance (BASELINE.json north_star): fp32 abs/rel <= 1e-4 on scores."""
import os

import numpy as np
import pytest
import torch

import gnnqc_oracle as O
from conftest import GOLDEN, MODELS, oracle_named
from gnnqc import NativeModel, synth

pytestmark = pytest.mark.gpu

ATOL = RTOL = 1e-4


def close(got, want):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    err = np.abs(got - want)
    assert err.shape == want.shape
    bad = err > ATOL + RTOL * np.abs(want)
    assert not bad.any(), "max abs err %.3e at %d (got %r want %r)" % (err.max(), err.argmax(), got.flat[err.argmax()], want.flat[err.argmax()])


@pytest.fixture(scope="module")
def models():
    out = {n: NativeModel.load(os.path.join(MODELS, n), device=0)
           for n in ("model_cml", "model_cml_baseline", "model_soilnet", "model_soilnet_baseline")}
    yield out
    for m in out.values():
        m.close()


def cuda(batch):
    return {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}


def score_cml(model, batch):
    b = cuda(batch)
    return model.score_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"]).cpu().numpy()


def oracle_cml(named, batch, dtype=torch.float32):
    feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
    w = O.to_torch_weights(named, dtype)
    p = O.gcn_forward_cml(torch.tensor(feat).to(dtype), batch["anom"].to(dtype), torch.tensor(idx), torch.tensor(samp), w)
    return p.reshape(-1).numpy()


# ---- golden vectors ----------------------------------------------------------------------------
@pytest.mark.parametrize("fixture,model", [("cml_baseline_scores.npz", "model_cml_baseline"),
                                           ("soilnet_baseline_scores.npz", "model_soilnet_baseline")])
def test_baseline_golden(models, fixture, model):
    g = np.load(os.path.join(GOLDEN, fixture))
    got = models[model].score_seq(torch.tensor(g["x"]).cuda()).cpu().numpy()
    close(got, g["scores"])


def test_cml_gcn_golden(models):
    g = np.load(os.path.join(GOLDEN, "cml_gcn_scores.npz"))
    batch = {k: torch.as_tensor(g[k]) for k in ("features", "anom", "edges", "node_off", "edge_off")}
    close(score_cml(models["model_cml"], batch), g["scores"])


# ---- seeded batches vs the oracle ---------------------------------------------------------------
@pytest.mark.parametrize("n,ragged", [(1, False), (63, True), (65, False), (200, True)])
def test_cml_gcn_vs_oracle(models, n, ragged):
    batch = synth.make_cml_batch(n, seed=100 + n, ragged=ragged)
    want = oracle_cml(oracle_named("model_cml"), batch)
    close(score_cml(models["model_cml"], batch), want)


@pytest.mark.parametrize("model,F,T,n", [("model_cml_baseline", 2, 181, 130), ("model_soilnet_baseline", 3, 337, 70)])
def test_baseline_vs_oracle(models, model, F, T, n):
    x = synth.make_seq_batch(n, F, T, seed=5)
    w = O.to_torch_weights(oracle_named(model), torch.float32)
    want = O.baseline_forward(x, w).reshape(-1).numpy()
    close(models[model].score_seq(x.cuda()).cpu().numpy(), want)


def oracle_soilnet(named, batch, dtype=torch.float32):
    feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 3)
    w = O.to_torch_weights(named, dtype)
    B = batch["node_off"].numel() - 1
    p = O.gcn_forward_soilnet(torch.tensor(feat).to(dtype), torch.tensor(idx), torch.tensor(samp), B, batch["T"], w)
    return p.reshape(-1).numpy()


@pytest.mark.parametrize("n,nodes,ragged", [(2, 14, False), (5, 40, True)])
def test_soilnet_gcn_vs_oracle(models, n, nodes, ragged):
    batch = synth.make_soilnet_batch(n, seed=31 + n, n_nodes=nodes, ragged=ragged, min_nodes=7)
    b = cuda(batch)
    got = models["model_soilnet"].score_soilnet(b["features"], b["edges"], b["node_off"], b["edge_off"], batch["T"]).cpu().numpy()
    close(got, oracle_soilnet(oracle_named("model_soilnet"), batch))


def test_soilnet_edge_order_does_not_matter(models):
    """The CSR build sorts neighbour lists: shuffling the caller's edge list changes nothing."""
    batch = synth.make_soilnet_batch(3, seed=5, n_nodes=30)
    b = cuda(batch)
    m = models["model_soilnet"]
    base = m.score_soilnet(b["features"], b["edges"], b["node_off"], b["edge_off"], batch["T"])
    eo = batch["edge_off"].numpy()
    g = torch.Generator().manual_seed(0)
    parts = []
    for w in range(3):
        e = batch["edges"][eo[w]: eo[w + 1]]
        parts.append(e[torch.randperm(e.shape[0], generator=g)])
    shuffled = torch.cat(parts).cuda()
    assert torch.equal(m.score_soilnet(b["features"], shuffled, b["node_off"], b["edge_off"], batch["T"]), base)


def test_other_sequence_lengths(models):
    """T is a runtime argument (the reference bakes it into the graph)."""
    for T in (27, 100, 243):
        x = synth.make_seq_batch(9, 2, T, seed=T)
        w = O.to_torch_weights(oracle_named("model_cml_baseline"), torch.float32)
        close(models["model_cml_baseline"].score_seq(x.cuda()).cpu().numpy(), O.baseline_forward(x, w).reshape(-1).numpy())
    from gnnqc import GnnqcError
    with pytest.raises(GnnqcError, match="too short"):
        models["model_cml_baseline"].score_seq(torch.zeros(4, 26, 2).cuda())


# ---- size-independent properties at larger sizes ---------------------------------------------------
def test_empty_batch(models):
    out = models["model_cml_baseline"].score_seq(torch.zeros(0, 181, 2).cuda())
    assert out.shape == (0,)


def test_permutation_and_split_invariance(models):
    """Windows are independent: scoring a permuted / split batch gives bit-identical scores."""
    m = models["model_cml_baseline"]
    x = synth.make_seq_batch(5000, 2, 181, seed=9).cuda()
    base = m.score_seq(x)
    perm = torch.randperm(x.shape[0], generator=torch.Generator().manual_seed(1)).cuda()
    assert torch.equal(m.score_seq(x[perm]), base[perm])
    halves = torch.cat([m.score_seq(x[:2777]), m.score_seq(x[2777:])])
    assert torch.equal(halves, base)
    assert torch.isfinite(base).all() and (base >= 0).all() and (base <= 1).all()


def test_cml_permutation_invariance_ragged(models):
    m = models["model_cml"]
    batch = synth.make_cml_batch(700, seed=3, ragged=True)
    base = score_cml(m, batch)
    # rebuild the batch with the windows reversed
    T = batch["T"]
    no, eo = batch["node_off"].numpy(), batch["edge_off"].numpy()
    order = np.arange(len(no) - 1)[::-1]
    feats = [batch["features"][no[b] * T: no[b + 1] * T] for b in order]
    edges = [batch["edges"][eo[b]: eo[b + 1]] for b in order]
    nn = np.array([no[b + 1] - no[b] for b in order])
    ne = np.array([eo[b + 1] - eo[b] for b in order])
    rev = {"features": torch.cat(feats), "anom": batch["anom"][torch.tensor(order.copy())], "edges": torch.cat(edges),
           "node_off": torch.tensor(np.concatenate([[0], np.cumsum(nn)]), dtype=torch.int32),
           "edge_off": torch.tensor(np.concatenate([[0], np.cumsum(ne)]), dtype=torch.int64), "T": T}
    assert np.array_equal(score_cml(m, rev), base[order])


def test_dead_steps_are_dead(models):
    """Steps past 162 (= 27*6) of a 181-step window cannot change the score."""
    m = models["model_cml_baseline"]
    x = synth.make_seq_batch(256, 2, 181, seed=11).cuda()
    y = x.clone()
    y[:, 162:] = 7.5
    assert torch.equal(m.score_seq(x), m.score_seq(y))


def test_isolated_node_and_duplicate_edges(models):
    """A node without incoming edges contributes 0 to the mean; duplicate edges count twice
    (unsorted_segment_mean semantics)."""
    m = models["model_cml"]
    T, n = 181, 4
    g = torch.Generator().manual_seed(0)
    x = 0.3 * torch.randn((1, T, n, 2), generator=g)
    edges = torch.tensor([[0, 0], [0, 1], [0, 1], [1, 0], [2, 3], [2, 2]], dtype=torch.int64)  # node 3 receives nothing
    batch = {"features": x.reshape(-1, 2), "anom": x[:, :, 0].contiguous(), "edges": edges,
             "node_off": torch.tensor([0, n], dtype=torch.int32), "edge_off": torch.tensor([0, len(edges)], dtype=torch.int64), "T": T}
    close(score_cml(m, batch), oracle_cml(oracle_named("model_cml"), batch))


def test_large_graph_uses_general_coefficient_path(models):
    """More than 32 nodes per window: the node-coefficient kernel's global-scratch path."""
    batch = synth.make_cml_batch(5, seed=77, n_nodes=40)
    close(score_cml(models["model_cml"], batch), oracle_cml(oracle_named("model_cml"), batch))


def test_roc_auc_matches_oracle_to_3_decimals(models):
    from sklearn.metrics import roc_auc_score
    batch = synth.make_cml_batch(400, seed=21, ragged=True)
    labels = batch["labels"].numpy()
    if labels.sum() == 0:
        labels[::7] = 1
    got = score_cml(models["model_cml"], batch)
    want = oracle_cml(oracle_named("model_cml"), batch)
    assert round(roc_auc_score(labels, got), 3) == round(roc_auc_score(labels, want), 3)


@pytest.mark.gpu
@pytest.mark.parametrize("ragged", [False, True])
def test_host_entry_point_matches_device_call(ragged):
    """gnnqc_score_cml_host (chunked staging, live timesteps only) == gnnqc_score_cml on the same batch, bit for bit."""
    from gnnqc import NativeModel
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    batch = synth.make_cml_batch(700, seed=77, ragged=ragged, n_nodes=11)
    dev = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
    want = m.score_cml(dev["features"], dev["anom"], dev["edges"], dev["node_off"], dev["edge_off"]).cpu()
    host = {k: batch[k].contiguous().pin_memory() for k in ("features", "anom", "edges", "node_off", "edge_off")}
    for chunk in (256, 8192):
        got, moved = m.score_cml_host(host["features"], host["anom"], host["edges"], host["node_off"], host["edge_off"], chunk_windows=chunk)
        assert torch.equal(got, want)
        full = host["features"].numel() * 4 + host["anom"].numel() * 4
        assert moved < full + host["edges"].numel() * 8 + 12 * 701 * 4
        if not ragged:
            assert moved < 0.93 * full + host["edges"].numel() * 8 + 12 * 701 * 4   # 162 of 181 steps cross PCIe
    m.close()


def _slice_cml(batch, w0, w1, T):
    """Windows [w0, w1) of an offsets-layout CML batch (device tensors), re-based to start at 0."""
    no, eo = batch["node_off"], batch["edge_off"]
    n0, n1, e0, e1 = int(no[w0]), int(no[w1]), int(eo[w0]), int(eo[w1])
    return {"features": batch["features"][n0 * T:n1 * T], "anom": batch["anom"][w0:w1], "edges": batch["edges"][e0:e1],
            "node_off": (no[w0:w1 + 1] - n0).to(torch.int32), "edge_off": eo[w0:w1 + 1] - e0}


@pytest.mark.gpu
def test_large_batch_equals_its_chunks():
    """More than 8 tiles per SM: the persistent kernels run several rounds per CTA.  The score of a window
    must not depend on how tiles are dealt, so one 80,037-window call == the same windows in 4,096-window calls."""
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    T = 181
    batch = synth.make_cml_batch(80037, seed=5, n_nodes=5, device="cuda")
    full = m.score_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"])
    parts = []
    for w0 in range(0, 80037, 4096):
        s = _slice_cml(batch, w0, min(80037, w0 + 4096), T)
        parts.append(m.score_cml(s["features"], s["anom"], s["edges"], s["node_off"], s["edge_off"]))
    assert torch.equal(full, torch.cat(parts))
    assert torch.isfinite(full).all()
    m.close()
