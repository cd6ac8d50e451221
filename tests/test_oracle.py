"""Oracle self-consistency (it is 'parity unpinned' w.r.t. TensorFlow - see its header) and
agreement with the committed golden vectors."""
import os

import numpy as np
import torch

import gnnqc_oracle as O
from conftest import GOLDEN, MODELS
from gnnqc import NativeModel, synth


def _named(name):
    return O.load_named(os.path.join(MODELS, name))


def _cml_inputs(g):
    batch = {k: torch.as_tensor(g[k]) for k in ("features", "anom", "edges", "node_off", "edge_off")}
    batch["T"] = int(g["anom"].shape[1])
    feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
    return torch.tensor(feat), batch["anom"], torch.tensor(idx), torch.tensor(samp)


def test_cml_gcn_golden_fp32_and_fp64():
    g = np.load(os.path.join(GOLDEN, "cml_gcn_scores.npz"))
    named = _named("model_cml")
    feat, anom, idx, samp = _cml_inputs(g)
    p64 = O.gcn_forward_cml(feat.double(), anom.double(), idx, samp, O.to_torch_weights(named, torch.float64)).reshape(-1)
    np.testing.assert_allclose(p64.numpy(), g["scores"], rtol=1e-9, atol=1e-12)
    p32 = O.gcn_forward_cml(feat, anom, idx, samp, O.to_torch_weights(named, torch.float32)).reshape(-1)
    np.testing.assert_allclose(p32.numpy(), g["scores"], rtol=1e-4, atol=1e-5)


def test_baseline_golden_and_numpy_cross_check():
    for fixture, model, first in (("cml_baseline_scores.npz", "model_cml_baseline", "dense1"),
                                  ("soilnet_baseline_scores.npz", "model_soilnet_baseline", "dense1")):
        g = np.load(os.path.join(GOLDEN, fixture))
        named = _named(model)
        p = O.baseline_forward(torch.tensor(g["x"]).double(), O.to_torch_weights(named, torch.float64)).reshape(-1)
        np.testing.assert_allclose(p.numpy(), g["scores"], rtol=1e-9, atol=1e-12)
        # second, independently written implementation (NumPy loops)
        np.testing.assert_allclose(O.np_lstm_stack_head(g["x"], named, "", first), g["scores"], rtol=1e-9, atol=1e-12)


def test_node_coefficient_collapse_equals_literal_graph_conv():
    """pooled = sum_j coef_j z_j (what the CUDA front-end computes) == GeneralConv + node mean."""
    g = np.load(os.path.join(GOLDEN, "cml_gcn_scores.npz"))
    named = _named("model_cml")
    w = O.to_torch_weights(named, torch.float64)
    feat, anom, idx, samp = _cml_inputs(g)
    T = anom.shape[1]
    lit = O.timeseries_pooling(O.general_conv(feat.double(), idx, w), samp, anom.shape[0], T)
    no, eo, edges = g["node_off"], g["edge_off"], g["edges"]
    for b in range(anom.shape[0]):
        n = int(no[b + 1] - no[b])
        e = edges[eo[b]: eo[b + 1]]
        deg = np.bincount(e[:, 0], minlength=n).astype(np.float64)
        coef = np.zeros(n)
        np.add.at(coef, e[:, 1], 1.0 / deg[e[:, 0]])
        coef /= n
        x = feat[no[b] * T: no[b + 1] * T].double().reshape(T, n, 2)
        z = x @ w["gcn_layer/kernel"] + w["gcn_layer/bias"]
        z = (z - w["gcn_layer/batch_norm/moving_mean"]) / torch.sqrt(w["gcn_layer/batch_norm/moving_variance"] + O.BN_EPS)
        z = z * w["gcn_layer/batch_norm/gamma"] + w["gcn_layer/batch_norm/beta"]
        z = torch.clamp(z, min=0) + w["gcn_layer/activation/alpha"] * torch.clamp(z, max=0)
        pooled = (z * torch.tensor(coef)[None, :, None]).sum(dim=1)
        np.testing.assert_allclose(pooled.numpy(), lit[b].numpy(), rtol=1e-10, atol=1e-12)


def test_dead_steps_do_not_change_scores():
    """Steps after 27*floor(floor(floor(T/3)/3)/3) cannot reach the output (StackPlan)."""
    g = np.load(os.path.join(GOLDEN, "cml_baseline_scores.npz"))
    named = _named("model_cml_baseline")
    w = O.to_torch_weights(named, torch.float64)
    x = torch.tensor(g["x"]).double()
    keep = 27 * (((x.shape[1] // 3) // 3) // 3)
    assert keep == 162
    x2 = x.clone()
    x2[:, keep:] = 123.0
    a = O.baseline_forward(x, w)
    b = O.baseline_forward(x2, w)
    assert torch.equal(a, b)


def test_integrated_gradients_completeness():
    """sum of attributions -> f(x) - f(0) as m_steps grows.  The shipped model is strongly
    non-monotonic along the straight path for full-amplitude windows (the trapezoid rule then
    needs hundreds of steps), so the axiom is checked on low-amplitude inputs."""
    named = _named("model_cml")
    w = O.to_torch_weights(named, torch.float64)
    batch = synth.make_cml_batch(3, seed=7, n_nodes=5)
    feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
    feat, idx, samp = 0.05 * torch.tensor(feat).double(), torch.tensor(idx), torch.tensor(samp)
    anom = 0.05 * batch["anom"].double()
    f0 = O.gcn_forward_cml(torch.zeros_like(feat), torch.zeros_like(anom), idx, samp, w).reshape(-1)
    errs = []
    for m in (4, 32):
        af, aa, p = O.integrated_gradients_cml(feat, anom, idx, samp, w, m_steps=m)
        tot = torch.stack([af[samp == b].sum() + aa[b].sum() for b in range(3)])
        errs.append(((tot - (p - f0)).abs() / (p - f0).abs()).max().item())
    assert errs[1] < errs[0]
    assert errs[1] < 2e-2
