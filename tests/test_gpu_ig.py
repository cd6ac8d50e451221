"""GPU parity of integrated gradients (gnnqc_ig_cml) against the CPU oracle (torch autograd through
the literal model restatement).  Tolerance: attributions within 1e-4 absolute + 1e-3 relative of
the float64 oracle (they are sums of ~50 gradient products of magnitude <= 1); scores 1e-4."""
import os

import numpy as np
import pytest
import torch

import gnnqc_oracle as O
from conftest import MODELS, oracle_named
from gnnqc import NativeModel, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def model():
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    yield m
    m.close()


def run_gpu(model, batch, m_steps):
    b = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
    af, aa, p = model.ig_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], m_steps=m_steps)
    return af.cpu().numpy(), aa.cpu().numpy(), p.cpu().numpy()


def run_oracle(model, batch, m_steps):
    feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
    w = O.to_torch_weights(oracle_named("model_cml"), torch.float64)
    af, aa, p = O.integrated_gradients_cml(torch.tensor(feat).double(), batch["anom"].double(), torch.tensor(idx), torch.tensor(samp), w, m_steps=m_steps)
    return af.numpy(), aa.numpy(), p.numpy()


def close(got, want, atol, rtol):
    err = np.abs(got.astype(np.float64) - want)
    bad = err > atol + rtol * np.abs(want)
    assert not bad.any(), "max abs err %.3e (|want| there %.3e, max |want| %.3e), %d bad of %d" % (
        err.max(), np.abs(want).flat[err.argmax()], np.abs(want).max(), bad.sum(), bad.size)


@pytest.mark.parametrize("n,ragged,m_steps", [(3, False, 4), (70, True, 3)])
def test_ig_matches_oracle(model, n, ragged, m_steps):
    batch = synth.make_cml_batch(n, seed=500 + n, ragged=ragged, n_nodes=9)
    gf, ga, gp = run_gpu(model, batch, m_steps)
    of, oa, op = run_oracle(model, batch, m_steps)
    close(gp, op, 1e-4, 1e-4)
    close(ga, oa, 1e-4, 1e-3)
    close(gf, of, 1e-4, 1e-3)


def test_ig_completeness_and_dead_steps(model):
    """Low-amplitude inputs: sum of attributions ~= f(x) - f(0); steps past 162 get exactly 0."""
    batch = synth.make_cml_batch(64, seed=9, n_nodes=6)
    batch["features"] = batch["features"] * 0.05
    batch["anom"] = batch["anom"] * 0.05
    gf, ga, gp = run_gpu(model, batch, 32)
    zero = {k: (torch.zeros_like(v) if k in ("features", "anom") else v) for k, v in batch.items()}
    # f(0) through the same fp32 kernels as the attributions (the scoring path runs its LSTM layers on
    # the tensor cores with split-bf16 products, 1e-5-level differences that matter for |f(x)-f(0)| < 1e-2)
    _, _, f0 = run_gpu(model, zero, 2)
    T, N = 181, 6
    tot = gf.reshape(64, T, N, 2).sum(axis=(1, 2, 3)) + ga.sum(axis=(1, 2))
    want = gp - f0
    smooth = np.abs(want) < 1e-2   # windows with an injected jump need far more than 32 steps to converge
    assert smooth.sum() > 32
    assert np.all(np.abs(tot - want)[smooth] <= 2e-2 * np.abs(want)[smooth] + 1e-6), np.abs(tot - want)[smooth].max()
    assert np.all(gf.reshape(64, T, N, 2)[:, 162:] == 0) and np.all(ga[:, 162:] == 0)
    # scores returned by the IG call are the plain scores
    b = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
    p = model.score_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"]).cpu().numpy()
    np.testing.assert_allclose(gp, p, rtol=1e-4, atol=5e-5)


def test_explainer_writes_the_six_files_per_sample(model, tmp_path):
    from gnnqc.config import to_config
    from libs.create_model import load_model
    from xai.libs.integrated_gradients import IntegratedGradientsExplainer, split_work

    clf = load_model(os.path.join(MODELS, "model_cml"), device=0)
    xai = to_config({"output_dir": str(tmp_path), "project": "unit-ig",
                     "integrated_gradients": {"m_steps": 4, "baseline": "zero", "scale_gradients": True, "negative_values": "keep",
                                              "threshold": 0.956, "dataset": "test", "which_samples": ["TN", "TP", "FP", "FN"]}})
    pre = to_config({"ds_type": "cml", "timestep_before": 120, "timestep_after": 60})
    ex = IntegratedGradientsExplainer(clf, xai, pre)
    batch = synth.make_cml_batch(5, seed=2, ragged=True, n_nodes=7)
    meta = ([("cml%d" % i) for i in range(5)], ["20190101_000%d00" % i for i in range(5)], [0, 1, 0, 0, 1])
    res = ex.get_gradients([batch], meta=[meta])[0]
    assert len(res["files"]) == 30
    no = batch["node_off"].numpy()
    gf, ga, gp = run_gpu(model, batch, 4)
    for i in range(5):
        d = [p for p in res["files"] if p.name.startswith("gradients_features_unwrapped_") and ("cml%d_" % i) in p.name]
        arr = np.load(d[0])
        n = no[i + 1] - no[i]
        assert arr.shape == (n, 181, 2)
        rows = gf[no[i] * 181: no[i + 1] * 181]
        for k in range(n):                       # the reference's a[i::n_neighbors]
            np.testing.assert_array_equal(arr[k], rows[k::n])
    assert split_work(list(range(7)), 1, 3) == [1, 4]


@pytest.mark.gpu
def test_ig_large_batch_equals_its_chunks(model):
    """602 tile pairs: more than the 592 group slots of the tensor-core backward kernels, so they run a
    second round; attributions must not depend on the dealing (one call == 8,192-window calls)."""
    from test_gpu_scoring import _slice_cml
    T, B = 181, 77000
    batch = synth.make_cml_batch(B, seed=6, n_nodes=4, device="cuda")
    af, aa, p = model.ig_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"], m_steps=1)
    for w0 in range(0, B, 8192):
        w1 = min(B, w0 + 8192)
        s = _slice_cml(batch, w0, w1, T)
        bf, ba, bp = model.ig_cml(s["features"], s["anom"], s["edges"], s["node_off"], s["edge_off"], m_steps=1)
        n0, n1 = int(batch["node_off"][w0]), int(batch["node_off"][w1])
        assert torch.equal(bf, af[n0 * T:n1 * T]) and torch.equal(ba, aa[w0:w1]) and torch.equal(bp, p[w0:w1])


def test_unscaled_gradients_with_zeros_in_the_input(model):
    """scale_gradients: False (reference :1006-1012 without :1185-1191): the trapezoid mean of the raw gradients, which is
    not zero where the input is exactly zero (rolling-median normalisation produces exact zeros)."""
    batch = synth.make_cml_batch(5, seed=77, n_nodes=6)
    batch["features"][::7] = 0.0
    batch["anom"][:, ::5] = 0.0
    b = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
    af, aa, p = model.ig_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], m_steps=4, scale_by_input=False)
    feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
    w = O.to_torch_weights(oracle_named("model_cml"), torch.float64)
    of, oa, op = O.integrated_gradients_cml(torch.tensor(feat).double(), batch["anom"].double(), torch.tensor(idx), torch.tensor(samp), w, m_steps=4,
                                            scale_gradients=False)
    live = np.zeros(batch["anom"].shape[1], bool)
    live[:162] = True                                   # dead steps get 0 from the kernel; their true raw gradient is 0 as well
    close(p.cpu().numpy(), op.numpy(), 1e-4, 1e-4)
    close(aa.cpu().numpy(), oa.numpy(), 1e-5, 1e-3)
    close(af.cpu().numpy(), of.numpy(), 1e-5, 1e-3)
    zero_rows = (batch["features"] == 0).all(dim=1).numpy()
    assert np.abs(af.cpu().numpy()[zero_rows]).max() > 0
