"""Regenerates everything under tests/golden/ .  Run in the build container, where the reference
is mounted at /root/reference (it does not exist on the GPU box):

    python tests/golden/make_fixtures.py

1. models/<name>/variables/variables.{index,data-*}: the four shipped checkpoints re-emitted by
   this repo's own tensor-bundle writer with the optimizer slots and Keras metric counters
   dropped (weights, model_info and normalisation strings are bit-identical; 0.75 MB instead of
   2.3 MB each).  These are data fixtures, not reference source.
2. *.npz golden vectors: seeded synthetic inputs plus the float64 oracle's outputs.  The oracle
   itself is "parity unpinned" (oracle/gnnqc_oracle.py header): these vectors freeze its
   behaviour, they are not TensorFlow outputs.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
sys.path.insert(0, os.path.join(ROOT, "gnn-xai-timeseries-qualitycontrol_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from gnnqc import NativeModel, synth  # noqa: E402
import gnnqc_oracle as O  # noqa: E402

REF = os.environ.get("GNNQC_REFERENCE", "/root/reference")
OUT = os.path.join(ROOT, "tests", "golden")
MODELS = ["model_cml", "model_cml_baseline", "model_soilnet", "model_soilnet_baseline"]


def strip_models():
    for name in MODELS:
        m = NativeModel.load(os.path.join(REF, name), device=None)
        m.save(os.path.join(OUT, "models", name))
        back = NativeModel.load(os.path.join(OUT, "models", name), device=None)
        assert np.array_equal(back.get_params(), m.get_params())
        print(name, "->", m.n_params, "floats")


def golden_cml():
    named = O.load_named(os.path.join(REF, "model_cml"))
    w = O.to_torch_weights(named, torch.float64)
    b = synth.make_cml_batch(12, seed=44, ragged=True)
    feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(b), 2)
    p = O.gcn_forward_cml(torch.tensor(feat).double(), b["anom"].double(), torch.tensor(idx), torch.tensor(samp), w)
    np.savez_compressed(
        os.path.join(OUT, "cml_gcn_scores.npz"),
        features=b["features"].numpy(), anom=b["anom"].numpy(), edges=b["edges"].numpy(),
        node_off=b["node_off"].numpy(), edge_off=b["edge_off"].numpy(), scores=p.reshape(-1).numpy())
    print("cml gcn", p.reshape(-1).numpy())

    nb = O.load_named(os.path.join(REF, "model_cml_baseline"))
    x = synth.make_seq_batch(24, 2, 181, seed=45)
    pb = O.baseline_forward(x.double(), O.to_torch_weights(nb, torch.float64))
    np.savez_compressed(os.path.join(OUT, "cml_baseline_scores.npz"), x=x.numpy(), scores=pb.reshape(-1).numpy())
    print("cml baseline", pb.reshape(-1).numpy())

    ns = O.load_named(os.path.join(REF, "model_soilnet_baseline"))
    xs = synth.make_seq_batch(10, 3, 337, seed=46)
    ps = O.baseline_forward(xs.double(), O.to_torch_weights(ns, torch.float64))
    np.savez_compressed(os.path.join(OUT, "soilnet_baseline_scores.npz"), x=xs.numpy(), scores=ps.reshape(-1).numpy())
    print("soilnet baseline", ps.reshape(-1).numpy())


if __name__ == "__main__":
    strip_models()
    golden_cml()
