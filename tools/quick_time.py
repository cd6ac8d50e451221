"""Scratch timing helper (not the bench contract): fp32 peak probe + per-kernel-stack timing."""
import ctypes as C
import os
import sys
import time

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(ROOT, "gnn-xai-timeseries-qualitycontrol_b200"))
import torch  # noqa: E402

from gnnqc import NativeModel, synth  # noqa: E402
from gnnqc._lib import lib  # noqa: E402

for mode in (0, 1):
    v = C.c_double()
    rc = lib.gnnqc_probe_fp32_peak(0, mode, C.byref(v))
    print("fp32 peak mode", mode, rc, "%.1f TFLOP/s" % v.value, flush=True)

MODELS = os.path.join(ROOT, "tests", "golden", "models")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
m = NativeModel.load(os.path.join(MODELS, "model_cml_baseline"), device=0)
x = synth.make_seq_batch(n, 2, 181, seed=1, device="cuda")
for it in range(3):
    torch.cuda.synchronize(); t = time.time()
    s = m.score_seq(x)
    torch.cuda.synchronize(); dt = time.time() - t
    print("baseline n=%d: %.2f ms -> %.0f windows/s (%.2f TFLOP/s useful)" % (n, dt * 1e3, n / dt, n / dt * 6.38e6 / 1e12), flush=True)
g = NativeModel.load(os.path.join(MODELS, "model_cml"), device=0)
b = synth.make_cml_batch(n, seed=2, device="cuda")
for it in range(3):
    torch.cuda.synchronize(); t = time.time()
    s = g.score_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], n_nodes_total=n * 23)
    torch.cuda.synchronize(); dt = time.time() - t
    print("gcn n=%d: %.2f ms -> %.0f windows/s" % (n, dt * 1e3, n / dt), flush=True)
print(s[:8])
