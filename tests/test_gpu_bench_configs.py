"""GPU parity at the shapes bench.py times (BASELINE.json configs 1-5), against the float64 oracle:
SoilNet GCN windows of up to 210 sensors at T = 337, integrated gradients at 50 and 100 interpolation
steps with 23-node windows, the tensor-core backward of the training step, and inputs that drive the LSTM
gates into saturation (where the split-bf16 contraction and the ex2/rcp activations have the least margin).
Each test prints the observed maximum error so the margin against the 1e-4 budget is on record."""
import os

import numpy as np
import pytest
import torch

import gnnqc_oracle as O
from conftest import MODELS, oracle_named
from gnnqc import NativeModel, synth

pytestmark = pytest.mark.gpu

ATOL = RTOL = 1e-4     # north_star: fp32 abs/rel <= 1e-4 on scores


def cuda(batch):
    return {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}


def close(got, want, what, atol=ATOL, rtol=RTOL):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    err = np.abs(got - want)
    print("%s: max abs err %.3e, max rel err %.3e over %d values" % (what, err.max(), (err / np.maximum(np.abs(want), 1e-12)).max(), err.size))
    bad = err > atol + rtol * np.abs(want)
    assert not bad.any(), "%s: max abs err %.3e at %d (got %r want %r)" % (what, err.max(), err.argmax(), got.flat[err.argmax()], want.flat[err.argmax()])


def test_soilnet_gcn_210_sensors_T337():
    """Config 3: one full 210-sensor window plus ragged ones (20..210 sensors), 337 steps each."""
    m = NativeModel.load(os.path.join(MODELS, "model_soilnet"), device=0)
    try:
        w = O.to_torch_weights(oracle_named("model_soilnet"), torch.float64)
        for batch in (synth.make_soilnet_batch(1, seed=61, n_nodes=210), synth.make_soilnet_batch(3, seed=62, n_nodes=210, ragged=True, min_nodes=20)):
            b = cuda(batch)
            got = m.score_soilnet(b["features"], b["edges"], b["node_off"], b["edge_off"], batch["T"]).cpu().numpy()
            feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 3)
            B = batch["node_off"].numel() - 1
            want = O.gcn_forward_soilnet(torch.tensor(feat).double(), torch.tensor(idx), torch.tensor(samp), B, batch["T"], w).reshape(-1).numpy()
            assert got.shape == want.shape == (int(batch["node_off"][-1]),)
            close(got, want, "soilnet gcn, %d windows, %d sequences" % (B, got.size))
    finally:
        m.close()


@pytest.mark.parametrize("m_steps,n", [(50, 3), (100, 2)])
def test_ig_50_and_100_steps_23_nodes(m_steps, n):
    """Config 4 (50 steps) and the reference's own default (100, xai_config_20240318.yml:12), N_b = 23."""
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    try:
        batch = synth.make_cml_batch(n, seed=900 + m_steps, n_nodes=23)
        b = cuda(batch)
        af, aa, p = m.ig_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], m_steps=m_steps)
        feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
        w = O.to_torch_weights(oracle_named("model_cml"), torch.float64)
        of, oa, op = O.integrated_gradients_cml(torch.tensor(feat).double(), batch["anom"].double(), torch.tensor(idx), torch.tensor(samp), w, m_steps=m_steps)
        close(p.cpu().numpy(), op.numpy(), "ig %d steps: scores" % m_steps)
        # attributions are means of ~m_steps gradient products: 1e-4 absolute + 1e-3 relative (tests/test_gpu_ig.py)
        close(aa.cpu().numpy(), oa.numpy(), "ig %d steps: target-series attributions" % m_steps, 1e-4, 1e-3)
        close(af.cpu().numpy(), of.numpy(), "ig %d steps: node-feature attributions" % m_steps, 1e-4, 1e-3)
        # completeness holds for the oracle to the discretisation error, and the GPU agrees with it
        tot_g = af.cpu().double().reshape(n, -1).sum(1) + aa.cpu().double().reshape(n, -1).sum(1)
        tot_o = of.reshape(n, -1).sum(1) + oa.reshape(n, -1).sum(1)
        assert torch.allclose(tot_g, tot_o, atol=2e-4, rtol=1e-3)
    finally:
        m.close()


def _grads(model, rvec):
    flat = rvec[: model.n_params].cpu().numpy()
    return {p: flat[o: o + int(np.prod(s))].reshape(s) for p, o, s, _ in model.segments()}, float(rvec[-1].item())


def _check_grads(got, want, skip=()):
    worst = 0.0
    for name, g in want.items():
        if name in skip:
            continue
        w = g.numpy()
        err = np.abs(got[name].astype(np.float64) - w)
        tol = 2e-4 * np.abs(w).max() + 1e-3 * np.abs(w) + 1e-9       # the tolerance of tests/test_gpu_train.py
        worst = max(worst, float((err / tol).max()))
        assert (err <= tol).all(), "%s: max err %.3e vs max |g| %.3e" % (name, err.max(), np.abs(w).max())
    print("worst gradient error / tolerance: %.3f" % worst)


@pytest.mark.parametrize("kind", ["gcn", "baseline"])
def test_tensor_core_backward_matches_oracle(kind, monkeypatch):
    """Training steps below 2 tiles per SM use the fp32 backward; GNNQC_TRAIN_TC_BWD_MIN_TILES=1 sends this
    small batch through lstm_tc_bwd_kernel (+ its dz export for the weight gradients) instead, so the
    tensor-core backward is compared with the float64 oracle directly, not only with the fp32 kernels."""
    monkeypatch.setenv("GNNQC_TRAIN_TC_BWD_MIN_TILES", "1")
    n = 150
    g = torch.Generator().manual_seed(5)
    labels = (torch.rand(n, generator=g) < 0.3).long()
    if kind == "gcn":
        m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
        try:
            m.profile(True)
            batch = synth.make_cml_batch(n, seed=77, n_nodes=6, ragged=True, min_nodes=3)
            b = cuda(batch)
            mom = m.train_moments_cml(b["features"])
            rvec = m.train_grads_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], labels.cuda(), mom)
            fin = rvec.clone()
            m.train_apply(fin, mom, 1e-4)
            got, loss = _grads(m, fin)
            assert any("bwd" in k for k in m.profile_read())
            feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
            oloss, ograds, _, _ = O.train_grads_cml(torch.tensor(feat), batch["anom"], torch.tensor(idx), torch.tensor(samp), labels, oracle_named("model_cml"))
            assert abs(loss - oloss.item()) <= 1e-5 * abs(oloss.item()) + 1e-7
            _check_grads(got, ograds, skip=("gcn_layer/bias",))
        finally:
            m.close()
    else:
        m = NativeModel.load(os.path.join(MODELS, "model_cml_baseline"), device=0)
        try:
            x = synth.make_seq_batch(n, 2, 181, seed=78)
            rvec = m.train_grads_seq(x.cuda(), labels.cuda())
            got, loss = _grads(m, rvec)
            oloss, ograds = O.train_grads_baseline(x, labels, oracle_named("model_cml_baseline"))
            assert abs(loss - oloss.item()) <= 1e-5 * abs(oloss.item()) + 1e-7
            _check_grads(got, ograds)
        finally:
            m.close()


@pytest.mark.parametrize("name,F,T,scale", [("model_cml_baseline", 2, 181, 1.0), ("model_cml_baseline", 2, 181, 25.0),
                                            ("model_soilnet_baseline", 3, 337, 1.0), ("model_soilnet_baseline", 3, 337, 8.0)])
def test_margin_on_long_and_saturated_sequences(name, F, T, scale):
    """Where the split-bf16 contraction has the least margin: T = 337 (324 dependent steps in layers 1-2)
    and inputs scaled until most gates of the first layers saturate.  Checked against float64."""
    m = NativeModel.load(os.path.join(MODELS, name), device=0)
    try:
        x = synth.make_seq_batch(512, F, T, seed=17) * scale
        want = O.baseline_forward(x.double(), O.to_torch_weights(oracle_named(name), torch.float64)).reshape(-1).numpy()
        got = m.score_seq(x.cuda()).cpu().numpy()
        close(got, want, "%s T=%d input x%g" % (name, T, scale))
    finally:
        m.close()


def test_cml_gcn_saturated_windows():
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    try:
        batch = synth.make_cml_batch(130, seed=19, n_nodes=23)
        batch["features"] = batch["features"] * 20.0
        batch["anom"] = batch["anom"] * 20.0
        b = cuda(batch)
        got = m.score_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"]).cpu().numpy()
        feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
        w = O.to_torch_weights(oracle_named("model_cml"), torch.float64)
        want = O.gcn_forward_cml(torch.tensor(feat).double(), batch["anom"].double(), torch.tensor(idx), torch.tensor(samp), w).reshape(-1).numpy()
        close(got, want, "cml gcn, inputs x20")
    finally:
        m.close()
