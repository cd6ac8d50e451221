"""Scratch: IG throughput + per-kernel breakdown."""
import os, sys, time
ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(ROOT, "gnn-xai-timeseries-qualitycontrol_b200"))
import torch
from gnnqc import NativeModel, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
m_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
m = NativeModel.load(os.path.join(ROOT, "tests", "golden", "models", "model_cml"), device=0)
b = synth.make_cml_batch(n, seed=2, device="cuda")
for it in range(2):
    torch.cuda.synchronize(); t = time.time()
    out = m.ig_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], m_steps=m_steps, n_nodes_total=n * 23)
    torch.cuda.synchronize(); dt = time.time() - t
    print("ig n=%d m=%d: %.1f ms -> %.0f attributions/s" % (n, m_steps, dt * 1e3, n / dt), flush=True)
m.profile(True)
m.ig_cml(b["features"], b["anom"], b["edges"], b["node_off"], b["edge_off"], m_steps=m_steps, n_nodes_total=n * 23)
prof = m.profile_read()
tot = sum(v[0] for v in prof.values())
for k, (ms, c) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    print("  %-16s %9.2f ms %5.1f%%  (%d launches)" % (k, ms, 100 * ms / tot, c))
