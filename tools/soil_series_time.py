"""Throughput of gnnqc_score_soilnet_series (windows built on the device from a resident SoilNet network):
210 sensors, 256 windows per step = 53,760 sensor-sequences of 337 samples, CUDA events."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gnn-xai-timeseries-qualitycontrol_b200"))
from gnnqc import NativeModel  # noqa: E402
from gnnqc.series import SoilSeries, soilnet_adjacency  # noqa: E402

rng = np.random.default_rng(0)
S, Tt, before, after, B = 210, 4000, 240, 96, 256
xy = rng.uniform(0, 300, size=(S // 3, 2))
place = np.arange(S) // 3
depth = np.tile([5.0, 20.0, 50.0], S // 3)
dist = np.linalg.norm(xy[place][:, None] - xy[place][None], axis=-1)
adj = soilnet_adjacency(dist, np.abs(depth[:, None] - depth[None]), 40.0, 45.0)
raw = np.stack([rng.uniform(5, 55, (S, Tt)), rng.uniform(-10, 35, (S, Tt)), rng.uniform(2900, 3500, (S, Tt))], -1).astype(np.float32)
raw[rng.random((S, Tt)) < 1e-4] = np.nan
target = (rng.random((S, Tt)) < 0.05).astype(np.float32)
m = NativeModel.load(os.path.join(ROOT, "tests", "golden", "models", "model_soilnet"), device=0)
series = SoilSeries(raw, target, adj, device=0)
t_h = torch.from_numpy(rng.integers(before, Tt - after, size=B).astype(np.int32)).pin_memory()
t_d = t_h.cuda()
for _ in range(3):
    scores, valid = series.score(m, t_d, before, after)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    scores, valid = series.score(m, t_d, before, after)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
import time  # noqa: E402
t0 = time.perf_counter()
for _ in range(5):
    s_h = series.score(m, t_h.cuda(non_blocking=True), before, after)[0].cpu()
dt = (time.perf_counter() - t0) / 5
print("mean degree %.1f, valid %.3f" % (adj.sum(1).mean(), valid.float().mean().item()))
print("device-resident: %.3f ms per %d sequences = %.2f M sequences/s; end to end (window list H2D, scores D2H): %.2f M sequences/s"
      % (ms, B * S, B * S / ms / 1e3, B * S / dt / 1e6))
