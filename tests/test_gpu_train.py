"""GPU parity of the training step (gnnqc_train_*) against the CPU oracle (torch autograd through
the training-mode restatement of the model, float64).  Gradients are compared per parameter tensor
with |err| <= 2e-4 * max|grad of that tensor| + 1e-3 * |grad| (fp32 accumulation over up to 162
steps x 64 windows); the loss with 1e-5 relative."""
import os

import numpy as np
import pytest
import torch

import gnnqc_oracle as O
from conftest import MODELS
from gnnqc import NativeModel, synth

pytestmark = pytest.mark.gpu


def cuda(batch):
    return {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}


def split_rvec(model, rvec):
    flat = rvec[: model.n_params].cpu().numpy()
    grads = {p: flat[o: o + int(np.prod(s))].reshape(s) for p, o, s, _ in model.segments()}
    return grads, float(rvec[-1].item())


def check_grads(got, want, skip=()):
    for name, g in want.items():
        if name in skip:
            continue
        w = g.numpy() if isinstance(g, torch.Tensor) else g
        err = np.abs(got[name].astype(np.float64) - w)
        tol = 2e-4 * np.abs(w).max() + 1e-3 * np.abs(w) + 1e-9
        assert (err <= tol).all(), "%s: max err %.3e vs max |g| %.3e" % (name, err.max(), np.abs(w).max())


def make_labels(n, seed):
    g = torch.Generator().manual_seed(seed)
    lab = (torch.rand(n, generator=g) < 0.3).long()
    lab[0], lab[-1] = 1, 0
    return lab


@pytest.mark.parametrize("n,ragged", [(6, False), (70, True)])
def test_gcn_gradients_match_oracle(n, ragged):
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    try:
        batch = synth.make_cml_batch(n, seed=40 + n, n_nodes=6, ragged=ragged, min_nodes=3)
        labels = make_labels(n, n)
        b = cuda(batch)
        mom = m.train_moments_cml(b["features"])
        rvec = m.train_grads_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], labels.cuda(), mom)
        fin = rvec.clone()
        m.train_apply(fin, mom, 1e-4)      # closes the BN backward: `fin` now holds the GeneralConv gradients too
        got, loss = split_rvec(m, fin)
        feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
        named = O.load_named(os.path.join(MODELS, "model_cml"))
        oloss, ograds, mu, var = O.train_grads_cml(torch.tensor(feat), batch["anom"], torch.tensor(idx), torch.tensor(samp), labels, named)
        assert abs(loss - oloss.item()) <= 1e-5 * abs(oloss.item()) + 1e-7
        check_grads(got, ograds, skip=("gcn_layer/bias",))
        assert np.abs(got["gcn_layer/bias"]).max() == 0.0     # exactly zero in front of batch-stat BN
        # BN moving statistics moved towards the batch statistics with momentum 0.99
        new = m.named_params()
        np.testing.assert_allclose(new["gcn_layer/batch_norm/moving_mean"], 0.99 * named["gcn_layer/batch_norm/moving_mean"] + 0.01 * mu.numpy(), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(new["gcn_layer/batch_norm/moving_variance"], 0.99 * named["gcn_layer/batch_norm/moving_variance"] + 0.01 * var.numpy(), rtol=1e-4, atol=1e-7)
    finally:
        m.close()


def test_baseline_gradients_and_adam_step():
    m = NativeModel.load(os.path.join(MODELS, "model_cml_baseline"), device=0)
    try:
        n = 90
        x = synth.make_seq_batch(n, 2, 181, seed=8)
        labels = make_labels(n, 1)
        before = m.get_params()
        rvec = m.train_grads_seq(x.cuda(), labels.cuda())
        got, loss = split_rvec(m, rvec)
        named = O.load_named(os.path.join(MODELS, "model_cml_baseline"))
        oloss, ograds = O.train_grads_baseline(x, labels, named)
        assert abs(loss - oloss.item()) <= 1e-5 * abs(oloss.item()) + 1e-7
        check_grads(got, ograds)
        # Adam, first step, from the GPU's own gradients
        g = rvec[: m.n_params].cpu().numpy().astype(np.float64)
        m.train_apply(rvec, None, 1e-4)
        after = m.get_params()
        mm, vv = 0.1 * g, 0.001 * g * g
        lr_t = 1e-4 * np.sqrt(1 - 0.999) / (1 - 0.9)
        want = before - lr_t * mm / (np.sqrt(vv) + 1e-7)
        np.testing.assert_allclose(after, want, rtol=0, atol=2e-7)
    finally:
        m.close()


def test_data_parallel_decomposition_is_exact_enough():
    """Two 'ranks' with half the batch each: summed moments and summed reduction vectors give the
    full-batch gradients (this is what the NCCL all-reduce does across GPUs)."""
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    try:
        n = 128
        full = synth.make_cml_batch(n, seed=3, n_nodes=5)
        labels = make_labels(n, 2)
        b = cuda(full)
        mom = m.train_moments_cml(b["features"])
        ref = m.train_grads_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], labels.cuda(), mom)
        T, N = 181, 5
        parts, moms = [], []
        for lo, hi in ((0, 64), (64, 128)):
            eo = full["edge_off"]
            sub = {"features": b["features"][lo * T * N: hi * T * N], "anom": b["anom"][lo:hi],
                   "edges": b["edges"][eo[lo]: eo[hi]], "node_off": (b["node_off"][lo: hi + 1] - b["node_off"][lo]).contiguous(),
                   "edge_off": (b["edge_off"][lo: hi + 1] - b["edge_off"][lo]).contiguous()}
            parts.append(sub)
            moms.append(m.train_moments_cml(sub["features"]))
        gmom = moms[0] + moms[1]
        torch.testing.assert_close(gmom, mom, rtol=1e-12, atol=1e-9)
        total = None
        for (lo, hi), sub in zip(((0, 64), (64, 128)), parts):
            r = m.train_grads_cml(sub["features"], sub["anom"], sub["edges"], sub["node_off"], sub["edge_off"], labels[lo:hi].cuda(), gmom,
                                  global_batch=n)
            total = r if total is None else total + r
        scale = ref.abs().max().item()
        assert (total - ref).abs().max().item() <= 2e-5 * scale
    finally:
        m.close()


def test_loss_goes_down_on_a_fixed_batch():
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    try:
        batch = cuda(synth.make_cml_batch(64, seed=12, n_nodes=8))
        labels = make_labels(64, 5).cuda()
        losses = []
        for _ in range(12):
            mom = m.train_moments_cml(batch["features"])
            rvec = m.train_grads_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"], labels, mom)
            losses.append(float(rvec[-1].item()))
            m.train_apply(rvec, mom, 5e-4)
        assert losses[-1] < losses[0]
        s = m.score_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"])
        assert torch.isfinite(s).all()
    finally:
        m.close()


@pytest.mark.parametrize("B,part", [(19200, 6400), (40960, 8192)])
def test_large_batch_gradients_equal_the_sum_of_their_parts(B, part):
    """With the same global batch size and BN moments the reduction vectors are additive, so one call and the sum of
    smaller calls must agree to rounding.  19,200 windows (150 tile pairs) and its 6,400-window parts all run the
    two-threads-per-window shape of the narrow LSTM kernels; 40,960 windows (320 pairs, more than one wave of that
    shape) run the four-group shape while their 8,192-window parts do not, and the weight-gradient kernel goes from 8
    to 45 items per slice."""
    from test_gpu_scoring import _slice_cml
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    try:
        T = 181
        batch = synth.make_cml_batch(B, seed=8, n_nodes=4, device="cuda")
        labels = make_labels(B, 8).cuda()
        mom = m.train_moments_cml(batch["features"])
        full = m.train_grads_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"], labels, mom,
                                 global_batch=B).double()
        parts = torch.zeros_like(full)
        for w0 in range(0, B, part):
            s = _slice_cml(batch, w0, w0 + part, T)
            parts += m.train_grads_cml(s["features"], s["anom"], s["edges"], s["node_off"], s["edge_off"], labels[w0:w0 + part], mom,
                                       global_batch=B).double()
        for name, off, shape, _ in m.segments():
            n = int(np.prod(shape))
            a, b = full[off:off + n].cpu().numpy(), parts[off:off + n].cpu().numpy()
            tol = 2e-4 * np.abs(b).max() + 1e-3 * np.abs(b) + 1e-9
            assert (np.abs(a - b) <= tol).all(), "%s: max err %.3e vs max |g| %.3e" % (name, np.abs(a - b).max(), np.abs(b).max())
        tail_a, tail_b = full[m.n_params:].cpu().numpy(), parts[m.n_params:].cpu().numpy()
        assert np.all(np.abs(tail_a - tail_b) <= 1e-3 * np.abs(tail_b) + 2e-4 * np.abs(tail_b).max())
    finally:
        m.close()


def test_device_refresh_of_tensor_core_operands_equals_host_packing():
    """After an optimizer step the bf16 operand tiles are re-derived on the device; re-uploading the same
    parameters from the host (host-side packing) must give bit-identical scores."""
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    try:
        batch = cuda(synth.make_cml_batch(300, seed=3, n_nodes=5))
        labels = make_labels(300, 3).cuda()
        mom = m.train_moments_cml(batch["features"])
        rvec = m.train_grads_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"], labels, mom)
        m.train_apply(rvec, mom, 1e-3)
        after_step = m.score_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"]).clone()
        m.set_params(m.get_params())
        from_host = m.score_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"])
        assert torch.equal(after_step, from_host)
    finally:
        m.close()


# ---- SoilNet (create_model.py:224-231; one label per (window, sensor), preprocessing_functions.py:861-897) -------------
@pytest.mark.parametrize("n_win,nodes,ragged", [(2, 9, False), (3, 14, True)])
def test_soilnet_gcn_gradients_match_oracle(n_win, nodes, ragged):
    m = NativeModel.load(os.path.join(MODELS, "model_soilnet"), device=0)
    try:
        batch = synth.make_soilnet_batch(n_win, seed=70 + n_win, n_nodes=nodes, ragged=ragged, min_nodes=5)
        n_seq = int(batch["node_off"][-1])
        labels = make_labels(n_seq, n_win)
        b = cuda(batch)
        mom = m.train_moments_soilnet(b["features"])
        rvec = m.train_grads_soilnet(b["features"], b["edges"], b["node_off"], b["edge_off"], labels.cuda(), batch["T"], mom)
        fin = rvec.clone()
        m.train_apply(fin, mom, 1e-4)      # closes the BN backward: `fin` now holds the GeneralConv gradients too
        got, loss = split_rvec(m, fin)
        feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 3)
        named = O.load_named(os.path.join(MODELS, "model_soilnet"))
        oloss, ograds, mu, var = O.train_grads_soilnet(torch.tensor(feat), torch.tensor(idx), torch.tensor(samp), n_win, batch["T"], labels, named)
        assert abs(loss - oloss.item()) <= 1e-5 * abs(oloss.item()) + 1e-7
        check_grads(got, ograds, skip=("gcn_layer/bias",))
        assert np.abs(got["gcn_layer/bias"]).max() == 0.0
        new = m.named_params()
        np.testing.assert_allclose(new["gcn_layer/batch_norm/moving_mean"], 0.99 * named["gcn_layer/batch_norm/moving_mean"] + 0.01 * mu.numpy(), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(new["gcn_layer/batch_norm/moving_variance"], 0.99 * named["gcn_layer/batch_norm/moving_variance"] + 0.01 * var.numpy(), rtol=1e-4, atol=1e-7)
    finally:
        m.close()


def test_soilnet_baseline_gradients_match_oracle():
    m = NativeModel.load(os.path.join(MODELS, "model_soilnet_baseline"), device=0)
    try:
        n = 40
        x = synth.make_seq_batch(n, 3, 337, seed=9)
        labels = make_labels(n, 4)
        rvec = m.train_grads_seq(x.cuda(), labels.cuda())
        got, loss = split_rvec(m, rvec)
        oloss, ograds = O.train_grads_baseline(x, labels, O.load_named(os.path.join(MODELS, "model_soilnet_baseline")))
        assert abs(loss - oloss.item()) <= 1e-5 * abs(oloss.item()) + 1e-7
        check_grads(got, ograds)
    finally:
        m.close()


def test_soilnet_loss_goes_down_and_windows_shard():
    """A few optimiser steps on a fixed SoilNet batch, and the data-parallel decomposition by whole windows."""
    from gnnqc import dp
    m = NativeModel.load(os.path.join(MODELS, "model_soilnet"), device=0)
    try:
        batch = synth.make_soilnet_batch(4, seed=13, n_nodes=12)
        n_seq = int(batch["node_off"][-1])
        labels = make_labels(n_seq, 6)
        b = cuda(batch)
        mom = m.train_moments_soilnet(b["features"])
        ref = m.train_grads_soilnet(b["features"], b["edges"], b["node_off"], b["edge_off"], labels.cuda(), batch["T"], mom)
        total, gmom = None, None
        for rank in range(2):
            sub, lab = dp.shard_soil_batch(batch, labels, rank, 2)
            sub = cuda(sub)
            mm = m.train_moments_soilnet(sub["features"])
            gmom = mm if gmom is None else gmom + mm
        torch.testing.assert_close(gmom, mom, rtol=1e-12, atol=1e-9)
        for rank in range(2):
            sub, lab = dp.shard_soil_batch(batch, labels, rank, 2)
            sub = cuda(sub)
            r = m.train_grads_soilnet(sub["features"], sub["edges"], sub["node_off"], sub["edge_off"], lab.cuda(), batch["T"], gmom,
                                      global_batch=n_seq, n_nodes_total=sub["n_nodes_total"])
            total = r if total is None else total + r
        assert (total - ref).abs().max().item() <= 2e-5 * ref.abs().max().item()
        losses = []
        for _ in range(8):
            mom = m.train_moments_soilnet(b["features"])
            rvec = m.train_grads_soilnet(b["features"], b["edges"], b["node_off"], b["edge_off"], labels.cuda(), batch["T"], mom)
            losses.append(float(rvec[-1].item()))
            m.train_apply(rvec, mom, 5e-4)
        assert losses[-1] < losses[0]
        s = m.score_soilnet(b["features"], b["edges"], b["node_off"], b["edge_off"], batch["T"])
        assert torch.isfinite(s).all()
    finally:
        m.close()


def test_optimizer_state_round_trips_and_other_rules(tmp_path):
    """model.save keeps the Adam slots and the iteration counter: a run resumed from the saved
    directory takes exactly the step the uninterrupted run takes.  SGD and RMSprop follow the Keras update rules."""
    x = synth.make_seq_batch(64, 2, 181, seed=21).cuda()
    labels = make_labels(64, 9).cuda()

    def step(model, lr=1e-3):
        rvec = model.train_grads_seq(x, labels)
        g = rvec[: model.n_params].cpu().numpy().astype(np.float64)
        model.train_apply(rvec, None, lr)
        return g

    a = NativeModel.load(os.path.join(MODELS, "model_cml_baseline"), device=0)
    try:
        step(a); step(a)
        a.save(tmp_path / "mid")
        assert a.train_iterations() == 2
        step(a)
        want = a.get_params()
    finally:
        a.close()
    import ckpt_reader as R
    bundle = R.read_bundle(str(tmp_path / "mid" / "variables" / "variables"))
    assert int(bundle["optimizer/_iterations/.ATTRIBUTES/VARIABLE_VALUE"]) == 2
    assert sum(k.startswith("optimizer/_variables/") for k in bundle) == 2 * 27     # m and v of the 27 trainable tensors
    names = {key: full for key, _, full in R.object_graph(bundle)}
    assert names["optimizer/_variables/1/.ATTRIBUTES/VARIABLE_VALUE"].startswith("Adam/m/")
    assert names["optimizer/_variables/2/.ATTRIBUTES/VARIABLE_VALUE"].startswith("Adam/v/")
    b = NativeModel.load(tmp_path / "mid", device=0)
    try:
        assert b.train_iterations() == 2
        step(b)
        assert np.array_equal(b.get_params(), want)
    finally:
        b.close()

    for name in ("sgd", "rmsprop"):
        m = NativeModel.load(os.path.join(MODELS, "model_cml_baseline"), device=0)
        try:
            m.set_optimizer(name)
            before = m.get_params().astype(np.float64)
            g = step(m, 1e-3)
            after = m.get_params()
            want = before - 1e-3 * g if name == "sgd" else before - 1e-3 * g / (np.sqrt(0.1 * g * g) + 1e-7)
            np.testing.assert_allclose(after, want, rtol=0, atol=3e-7)
            if name == "rmsprop":
                g2 = step(m, 1e-3)
                v = 0.9 * 0.1 * g * g + 0.1 * g2 * g2
                np.testing.assert_allclose(m.get_params(), want - 1e-3 * g2 / (np.sqrt(v) + 1e-7), rtol=0, atol=5e-7)
                m.save(tmp_path / "rms")
                again = NativeModel.load(tmp_path / "rms", device=0)
                assert again.train_iterations() == 2
                again.close()
        finally:
            m.close()
