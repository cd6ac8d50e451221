"""End-to-end use of the reference-named API on a GPU: construct, train a few epochs, score, save,
reload (libs/create_model.py, libs/fit_model.py, libs/test_model.py mirrors)."""
import numpy as np
import pytest
import torch

from gnnqc import synth
from gnnqc.config import default_model_config, default_preproc_config
from libs import fit_model, test_model
from libs.create_model import BaselineClassifier, GCNClassifier, load_model

pytestmark = pytest.mark.gpu


def cml_batches(n_batches, size, seed):
    out = []
    for i in range(n_batches):
        b = synth.make_cml_batch(size, seed=seed + i, n_nodes=6)
        labels = b.pop("labels")
        labels[0] = 1
        out.append((b, labels))
    return out


def test_train_model_gcn_round_trip(tmp_path):
    mc, pc = default_model_config(), default_preproc_config("cml")
    mc.epochs, mc.learning_rate, mc.es_patience = 4, 2e-3, 10
    model = GCNClassifier(mc, pc, device=0, seed=1)
    assert list(model.model_info.numpy()) == [120, 60, 128, 1]
    train, val = cml_batches(3, 64, 100), cml_batches(1, 64, 200)
    history, model = fit_model.train_model(model, mc, pc, train, val, checkpoint_path=str(tmp_path / "ckpt"))
    assert len(history["loss"]) == 4 and history["loss"][-1] < history["loss"][0]
    assert np.isfinite(history["val_loss"]).all()
    p = model.predict(val)
    assert p.shape == (64, 1) and np.isfinite(p).all()
    model.save(str(tmp_path / "saved"))
    again = load_model(str(tmp_path / "saved"), device=0)
    np.testing.assert_array_equal(again.predict(val), p)
    thr, idx = test_model.calculate_threshold(mc, pc, val, model)
    assert 0.0 <= thr <= 1.0 and idx == 120


def test_train_model_baseline():
    mc, pc = default_model_config(), default_preproc_config("cml")
    mc.epochs, mc.learning_rate = 3, 2e-3
    model = BaselineClassifier(mc, pc, device=0, seed=2)
    x = synth.make_seq_batch(128, 2, 181, seed=3)
    y = (x[:, 120:, 0].mean(dim=1) > 0.2).long()
    y[0] = 1
    history, model = fit_model.train_model(model, mc, pc, [(x, y)], [(x, y)], baseline=True)
    assert history["loss"][-1] < history["loss"][0]


def test_reference_layout_inputs_give_the_same_scores(models_dir):
    """model(inputs) with the reference's disjoint-graph tuple == the offsets-layout dict."""
    import os

    from gnnqc import layout
    model = load_model(os.path.join(models_dir, "model_cml"), device=0)
    b = synth.make_cml_batch(20, seed=6, ragged=True)
    idx, samp = layout.offsets_to_disjoint(b["edges"], b["node_off"], b["edge_off"], b["T"])
    ref_inputs = (b["features"], b["anom"], (idx, torch.ones(idx.shape[0]), None), samp, None)
    a = model(ref_inputs).cpu().numpy()
    c = model(b).cpu().numpy()
    np.testing.assert_array_equal(a, c)
    assert a.shape == (20, 1)


def test_soilnet_reference_layouts_and_training(models_dir):
    """SoilNet GCN with the reference's (features, adjacency, sample_indices) tuple == the offsets-layout dict; the
    SoilNet baseline from (features, sample_indicator) rows == scoring the graph_reshape'd sequences; and a few
    training epochs of a freshly initialised SoilNet GCN (one label per (window, sensor))."""
    import os

    from gnnqc import layout
    b = synth.make_soilnet_batch(5, seed=8, n_nodes=17, ragged=True, min_nodes=6)
    labels = b.pop("labels")
    idx, samp = layout.offsets_to_disjoint(b["edges"], b["node_off"], b["edge_off"], b["T"])
    model = load_model(os.path.join(models_dir, "model_soilnet"), device=0)
    a = model((b["features"], (idx, torch.ones(idx.shape[0]), None), samp)).cpu().numpy()
    c = model(b).cpu().numpy()
    np.testing.assert_array_equal(a, c)
    assert a.shape == (int(b["node_off"][-1]), 1)

    base = load_model(os.path.join(models_dir, "model_soilnet_baseline"), device=0)
    rows = base((b["features"], samp)).cpu().numpy()
    T, seqs, no = b["T"], [], b["node_off"].tolist()
    for w in range(5):
        n = no[w + 1] - no[w]
        seqs.append(b["features"][no[w] * T: no[w + 1] * T].reshape(T, n, 3).permute(1, 0, 2))
    direct = base.native.score_seq(torch.cat(seqs).contiguous().cuda()).cpu().numpy().reshape(-1, 1)
    np.testing.assert_array_equal(rows, direct)

    mc, pc = default_model_config(), default_preproc_config("soilnet")
    mc.epochs, mc.learning_rate, mc.es_patience = 3, 2e-3, 10
    fresh = GCNClassifier(mc, pc, device=0, seed=3)
    labels[0] = 1
    history, fresh = fit_model.train_model(fresh, mc, pc, [(b, labels)], [(b, labels)])
    assert len(history["loss"]) == 3 and np.isfinite(history["loss"]).all() and history["loss"][-1] < history["loss"][0]
