"""A/B helper (not the bench contract): per-kernel CUDA-event times of one CML GCN scoring pass at full size, and the
maximum score error of 1,024 windows against the float64 oracle.  Run once per library build:
    GNNQC_LIB=.../libgnnqc_p3.so python tools/ab_scoring.py [windows]"""
import os
import sys

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path[:0] = [os.path.join(ROOT, "gnn-xai-timeseries-qualitycontrol_b200"), os.path.join(ROOT, "oracle")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

import gnnqc_oracle as O  # noqa: E402
from gnnqc import NativeModel, synth  # noqa: E402

MODELS = os.path.join(ROOT, "tests", "golden", "models")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
which = sys.argv[2] if len(sys.argv) > 2 else "cml"
tag = os.path.basename(os.environ.get("GNNQC_LIB", "libgnnqc.so"))
if which == "cml":
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
    b = synth.make_cml_batch(n, seed=44, device="cuda")
    run = lambda: m.score_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], n_nodes_total=n * 23)  # noqa: E731
else:
    m = NativeModel.load(os.path.join(MODELS, "model_soilnet_baseline"), device=0)
    x = synth.make_seq_batch(n, 3, 337, seed=44, device="cuda")
    run = lambda: m.score_seq(x)  # noqa: E731
for _ in range(3):
    s = run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    s = run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
m.profile(True)
for _ in range(5):
    run()
prof = m.profile_read()
m.profile(False)
k = 1024
if which == "cml":
    cpu = {kk: (v.cpu() if isinstance(v, torch.Tensor) else v) for kk, v in synth.make_cml_batch(k, seed=44, device="cuda").items()}
    got = m.score_cml(*(cpu[kk].cuda() for kk in ("features", "anom", "edges", "node_off", "edge_off"))).cpu().numpy()
    feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(cpu), 2)
    w = O.load_weights(os.path.join(MODELS, "model_cml"), torch.float64)
    with torch.no_grad():
        want = O.gcn_forward_cml(torch.tensor(feat).double(), cpu["anom"].double(), torch.tensor(idx), torch.tensor(samp), w).reshape(-1).numpy()
else:
    xs = synth.make_seq_batch(k, 3, 337, seed=17)
    got = m.score_seq(xs.cuda()).cpu().numpy()
    with torch.no_grad():
        want = O.baseline_forward(xs.double(), O.load_weights(os.path.join(MODELS, "model_soilnet_baseline"), torch.float64)).reshape(-1).numpy()
err = np.abs(got - want)
print("%s %s n=%d: %.3f ms/pass = %.2f M units/s | max abs err %.2e max rel %.2e | %s" % (
    tag, which, n, ms, n / ms / 1e3, err.max(), (err / np.maximum(want, 1e-12)).max(),
    " ".join("%s=%.3f" % (kk, v[0] / 5) for kk, v in sorted(prof.items()))), flush=True)
