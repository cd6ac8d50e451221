#!/usr/bin/env python
"""bench.py - windows scored per second by the CML GCN model (BASELINE.json configs[0]) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
                    [--workload cml_gcn|cml_baseline|soilnet_gcn|ig|train]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the scoring hot path over one batch of synthetic CML-shaped windows
(T=181, 23 graph nodes per window).  Prints ONE JSON line.
  value     windows/s, whole job, inputs resident in HBM, CUDA-event timed, max over ranks
  e2e       windows/s through the Python API from pinned HOST buffers (H2D + D2H inside the region)
  roofline  the largest single kernel (the graph front-end, HBM bound) against the measured copy
            bandwidth; nested: the tcgen05 LSTM layers against the measured bf16 peak ("tensor"), the
            MUFU pipe that actually bounds them ("xu"), and whole-step algorithmic HBM traffic ("hbm")
  cpu_baseline  the CPU oracle (torch fp32, all host cores) on a bounded sample - "port", since
            the reference's TensorFlow path is not installable here
--impl reference times that CPU oracle alone (the reference's own CPU implementation is unavailable).

Other workloads (not the driver's default line): cml_baseline (configs[1]), soilnet_gcn (configs[2],
210-sensor windows, unit = sensor-sequences/s), ig (configs[3], 50 interpolation steps, unit =
attributions/s), train (configs[4], data-parallel training step, unit = windows/s).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "gnn-xai-timeseries-qualitycontrol_b200"))

MODELS = os.path.join(ROOT, "tests", "golden", "models")
T_CML, N_NODES = 181, 23
METRIC = "cml_windows_scored_per_sec"
METRICS = {"cml_gcn": ("cml_windows_scored_per_sec", "windows/s"), "cml_baseline": ("cml_baseline_windows_scored_per_sec", "windows/s"),
           "soilnet_gcn": ("soilnet_sensor_sequences_scored_per_sec", "sequences/s"), "ig": ("cml_ig_attributions_per_sec", "attributions/s"),
           "train": ("cml_gcn_training_windows_per_sec", "windows/s"),
           "cml_series": ("cml_windows_scored_per_sec_from_resident_series", "windows/s")}

# Algorithmic FLOPs (2*MACs of every matmul) per window and per kernel.
LSTM_SHAPES = {"cml_gcn": [(181, 18, 16), (181, 16, 16), (60, 16, 32), (60, 32, 32), (20, 32, 64), (20, 64, 64), (6, 64, 128)],
               "cml_baseline": [(181, 2, 16), (181, 16, 16), (60, 16, 32), (60, 32, 32), (20, 32, 64), (20, 64, 64), (6, 64, 128)]}
# Steps that can reach the score (dead-step elimination): 162/162/54/54/18/18/6.
LIVE_STEPS = [162, 162, 54, 54, 18, 18, 6]
HEAD_FLOPS = 2 * (128 * 64 + 64 * 64 + 64)


def layer_flops(workload, live=False):
    out = []
    for i, (t, inp, hid) in enumerate(LSTM_SHAPES[workload]):
        steps = LIVE_STEPS[i] if live else t
        out.append(2 * steps * (inp + hid) * 4 * hid)
    return out


def bytes_per_window(workload):
    if workload == "cml_gcn":
        return 4 * 2 * T_CML * (N_NODES + 1) + 4 * N_NODES + 4      # 34,848 B
    return 4 * 2 * T_CML + 4


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index):
        self.samples = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "25"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.t = threading.Thread(target=self._read, daemon=True)
        self.t.start()

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            parts = [p.strip() for p in s.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


class CpuOracle:
    """The CPU oracle (torch fp32, all host threads) on a fixed sample of the workload, fed in batches
    of 128 windows - the reference's batch_size (preprocessing_config.yml:5)."""

    BATCH = 128

    def __init__(self, workload, sample_windows):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import torch

        import gnnqc_oracle as O
        from gnnqc import NativeModel, synth

        self.torch = torch
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.n = sample_windows
        name = "model_cml" if workload == "cml_gcn" else "model_cml_baseline"
        named = NativeModel.load(os.path.join(MODELS, name), device=None).named_params()
        w = O.to_torch_weights(named, torch.float32)
        self.calls = []
        for i, lo in enumerate(range(0, sample_windows, self.BATCH)):
            n = min(self.BATCH, sample_windows - lo)
            if workload == "cml_gcn":
                batch = synth.make_cml_batch(n, seed=44 + i, n_nodes=N_NODES)
                feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
                args = (torch.tensor(feat), batch["anom"], torch.tensor(idx), torch.tensor(samp), w)
                self.calls.append((O.gcn_forward_cml, args))
            else:
                self.calls.append((O.baseline_forward, (synth.make_seq_batch(n, 2, T_CML, seed=44 + i), w)))

    def step(self):
        with self.torch.no_grad():
            t0 = time.perf_counter()
            for fn, args in self.calls:
                fn(*args)
            return time.perf_counter() - t0


def cpu_oracle_throughput(workload, sample_windows, repeats=1):
    o = CpuOracle(workload, sample_windows)
    o.step()  # warm-up (thread pool, allocator)
    best = min(o.step() for _ in range(repeats))
    return sample_windows / best, o.cores, best


def run_reference(args, rank, world):
    """--impl reference: the reference's own implementation (TensorFlow/Keras/spektral) cannot be
    installed in this image (no network, not in the wheelhouse), so - as the task's tier rules
    say - this arm times the CPU oracle port on the host cores."""
    if rank != 0:
        return
    sample = 2048 if args.workload == "cml_gcn" else 16384
    o = CpuOracle(args.workload, sample)
    cores = o.cores
    times = []
    for i in range(args.warmup + args.steps):
        dt = o.step()
        if i >= args.warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    value = sample / (ms * 1e-3)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "windows/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "windows_per_step": sample, "batch_size": 128, "seq_len": T_CML, "nodes_per_window": N_NODES,
                   "note": "CPU oracle port (torch fp32); reference TF/Keras CPU path unavailable on this box"},
        "cpu_baseline": {"value": value, "unit": "windows/s", "cores": cores, "kind": "port",
                         "sample": "%d windows per step in batches of 128" % sample},
        "e2e": {"value": value, "unit": "windows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gnnqc", choices=["gnnqc", "reference"])
    ap.add_argument("--workload", default="cml_gcn", choices=sorted(METRICS))
    ap.add_argument("--windows", type=int, default=0, help="units per GPU per step (0 = workload default)")
    ap.add_argument("--m-steps", type=int, default=50, help="IG interpolation steps")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "gnnqc" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if args.workload not in ("cml_gcn", "cml_baseline"):
            args.workload = "cml_gcn"
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    from gnnqc import NativeModel, dp, synth
    from gnnqc._lib import lib
    from gnnqc.host_pipeline import CmlHostScorer, SeqHostScorer

    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    wl = args.workload
    metric, unit = METRICS[wl]
    model_name = {"cml_gcn": "model_cml", "ig": "model_cml", "train": "model_cml", "cml_series": "model_cml", "cml_baseline": "model_cml_baseline",
                  "soilnet_gcn": "model_soilnet"}[wl]
    model = NativeModel.load(os.path.join(MODELS, model_name), device=local_rank)
    B = args.windows or {"cml_gcn": 65536, "cml_baseline": 65536, "soilnet_gcn": 256, "ig": 65536, "train": 4096, "cml_series": 65536}[wl]
    T = T_CML
    cfg_extra = {}
    e2e_path = None

    # Synthetic inputs, generated on the device (weak scaling: every rank has its own B units per step).
    if wl in ("cml_gcn", "ig", "train"):
        batch = synth.make_cml_batch(B, seed=44 + rank, n_nodes=N_NODES, device=str(dev))
        n_nodes_total = B * N_NODES
        keys = ("features", "anom", "edges", "node_off", "edge_off")
        in_bytes = sum(batch[k].numel() * batch[k].element_size() for k in keys)
        units = B
        if wl == "cml_gcn":
            step = lambda: model.score_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"], n_nodes_total)  # noqa: E731
        elif wl == "ig":
            cfg_extra["m_steps"] = args.m_steps
            step = lambda: model.ig_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"], args.m_steps, n_nodes_total)[2]  # noqa: E731
        else:
            labels = (torch.rand(B, device=dev) < 1.0 / 6.0).long()      # 16.7 % positives
            cfg_extra.update({"class_weights": [1, 5], "optimizer": "adam", "learning_rate": 1e-4, "global_batch": B * world,
                              "collectives_per_step": "all-reduce of 6 f64 moments + all-reduce of the %d-float reduction vector" % model.train_vector_len()})

            def step():
                mom = model.train_moments_cml(batch["features"])
                dp.all_reduce_sum(mom)
                rvec = model.train_grads_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"], labels, mom,
                                             (1.0, 5.0), global_batch=B * world, n_nodes_total=n_nodes_total)
                dp.all_reduce_sum(rvec)
                model.train_apply(rvec, mom, 1e-4)
                return rvec[-1:]
    elif wl == "cml_series":
        # On-device window builder: a resident raw network, windows named by (sensor, time index).
        from gnnqc.series import CmlSeries
        S, per = 2048, max(1, B // 2048)
        net = synth.make_cml_network(S, 4096, seed=42 + rank)
        series = CmlSeries(net["raw"], net["rmed"], net["nbr_off"], net["nbr_idx"], net["edge_off"], net["edges"], device=local_rank)
        B = S * per
        tgt_h = torch.arange(S, dtype=torch.int32).repeat_interleave(per).pin_memory()
        tau_h = (2000 + torch.arange(per, dtype=torch.int32)).repeat(S).pin_memory()
        tgt_d, tau_d = tgt_h.to(dev), tau_h.to(dev)
        step = lambda: model.score_cml_series(series, tgt_d, tau_d, 120, 60)  # noqa: E731
        in_bytes, units = B * 8, B
        cfg_extra.update({"sensors": S, "consecutive_windows_per_sensor": per, "resident_series_bytes": int(net["raw"].nbytes * 2),
                          "mean_neighbourhood": float(np.diff(net["nbr_off"]).mean())})
    elif wl == "cml_baseline":
        x = synth.make_seq_batch(B, 2, T_CML, seed=44 + rank, device=str(dev))
        step = lambda: model.score_seq(x)  # noqa: E731
        in_bytes, units = x.numel() * 4, B
    else:  # soilnet_gcn: B windows of 210 sensors, T = 337
        T = 337
        batch = synth.make_soilnet_batch(B, seed=44 + rank, n_nodes=210, device=str(dev))
        n_nodes_total = B * 210
        keys = ("features", "edges", "node_off", "edge_off")
        in_bytes = sum(batch[k].numel() * batch[k].element_size() for k in keys)
        units = n_nodes_total
        cfg_extra.update({"windows_per_gpu_per_step": B, "sensors_per_window": 210})
        step = lambda: model.score_soilnet(batch["features"], batch["edges"], batch["node_off"], batch["edge_off"], T, n_nodes_total)  # noqa: E731

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing -------------------------------------------------------------
    for _ in range(args.warmup):
        result = step()
    barrier()
    launches0 = model.launch_count()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        result = step()
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = model.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms_total], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = world * units / (ms_step * 1e-3)

    # ---- per-kernel timing (events around every launch; not for the training loop, whose state moves) ----
    prof = {}
    prof_steps = args.steps if wl != "ig" else 1
    model.profile(True)
    for _ in range(prof_steps):
        step()
    prof = model.profile_read()
    model.profile(False)

    # ---- end to end from pinned host buffers ----------------------------------------------------
    h2d = d2h = 0
    if wl == "cml_gcn":
        host = {k: batch[k].cpu().pin_memory() for k in keys}
        class _Native:   # the C-ABI call with host buffers; keeps the byte counts like the Python scorers do
            last_h2d_bytes = last_d2h_bytes = 0

            def score(self):
                out, moved = model.score_cml_host(host["features"], host["anom"], host["edges"], host["node_off"], host["edge_off"], chunk_windows=8192)
                self.last_h2d_bytes, self.last_d2h_bytes = moved, out.numel() * 4
                return out
        scorer = _Native()
        e2e_step = scorer.score
        e2e_path = "gnnqc_score_cml_host (pinned host -> chunked H2D of the live timesteps on a copy stream -> scoring kernels -> D2H)"
    elif wl == "cml_series":
        scorer = None
        slab_raw = np.ascontiguousarray(net["raw"][:, 2000 + 60: 2000 + 60 + per])
        slab_med = np.ascontiguousarray(net["rmed"][:, 2000 + 60: 2000 + 60 + per])

        def e2e_step():
            # the samples that are new for this step's windows arrive from the host, then the window list
            series.update(2000 + 60, slab_raw, slab_med)
            return model.score_cml_series(series, tgt_h.to(dev, non_blocking=True), tau_h.to(dev, non_blocking=True), 120, 60).cpu()
        h2d, d2h = int(slab_raw.nbytes * 2 + B * 8), B * 4
        e2e_path = "gnnqc_series_update (new samples) + window list H2D -> gnnqc_score_cml_series -> D2H"
    elif wl == "cml_baseline":
        host_x = x.cpu().pin_memory()
        scorer = SeqHostScorer(model, chunk_seqs=16384)
        e2e_step = lambda: scorer.score(host_x)  # noqa: E731
        e2e_path = "SeqHostScorer"
    else:
        host = {k: batch[k].cpu().pin_memory() for k in keys}
        scorer = None
        if wl == "train":
            host_labels = labels.cpu().pin_memory()
        pinned_out = {}

        def to_host(name, t):   # device -> pinned host buffer (allocated once), asynchronous
            if name not in pinned_out:
                pinned_out[name] = torch.empty(t.shape, dtype=t.dtype).pin_memory()
            pinned_out[name].copy_(t, non_blocking=True)
            return pinned_out[name]

        def e2e_step():
            d = {k: host[k].to(dev, non_blocking=True) for k in keys}
            if wl == "soilnet_gcn":
                return model.score_soilnet(d["features"], d["edges"], d["node_off"], d["edge_off"], T, n_nodes_total).cpu()
            if wl == "ig":
                af, aa, p = model.ig_cml(d["features"], d["anom"], d["edges"], d["node_off"], d["edge_off"], args.m_steps, n_nodes_total)
                res = to_host("af", af), to_host("aa", aa), to_host("p", p)
                torch.cuda.synchronize()
                return res
            lab = host_labels.to(dev, non_blocking=True)
            mom = model.train_moments_cml(d["features"])
            dp.all_reduce_sum(mom)
            rvec = model.train_grads_cml(d["features"], d["anom"], d["edges"], d["node_off"], d["edge_off"], lab, mom, (1.0, 5.0),
                                         global_batch=B * world, n_nodes_total=n_nodes_total)
            dp.all_reduce_sum(rvec)
            model.train_apply(rvec, mom, 1e-4)
            return rvec[-1:].cpu()
        h2d = in_bytes + (B * 8 if wl == "train" else 0)
        d2h = {"soilnet_gcn": n_nodes_total * 4, "ig": batch["features"].numel() * 4 + (batch["anom"].numel() * 4 if wl == "ig" else 0) + B * 4, "train": 4}[wl]
        e2e_path = "pinned host -> H2D -> %s -> D2H of the result" % {"soilnet_gcn": "gnnqc_score_soilnet", "ig": "gnnqc_ig_cml", "train": "gnnqc_train_*"}[wl]
    for _ in range(2):
        out = e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(e2e_steps):
        out = e2e_step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * units * e2e_steps / float(t.item())
    if scorer is not None:
        h2d, d2h = scorer.last_h2d_bytes, scorer.last_d2h_bytes
        if not torch.equal(out.cuda(dev), result):
            raise SystemExit("e2e scores differ from the device-resident scores")

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline ------------------------------------------------------------------------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak, hbm_src = 6650.0, "fallback"
    if os.path.exists(peaks_path):
        hbm_peak, hbm_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    import ctypes as C
    fp32 = C.c_double()
    lib.gnnqc_probe_fp32_peak(local_rank, 1, C.byref(fp32))
    shapes = "cml_baseline" if wl == "cml_baseline" else "cml_gcn"
    if wl == "soilnet_gcn":
        fl_alg = [2 * t_ * (i_ + h_) * 4 * h_ for t_, i_, h_ in [(337, 19, 16), (337, 16, 16), (112, 16, 32), (112, 32, 32), (37, 32, 64), (37, 64, 64), (12, 64, 128)]]
        fl_live = [2 * t_ * (i_ + h_) * 4 * h_ for t_, i_, h_ in [(324, 19, 16), (324, 16, 16), (108, 16, 32), (108, 32, 32), (36, 32, 64), (36, 64, 64), (12, 64, 128)]]
    else:
        fl_alg, fl_live = layer_flops(shapes), layer_flops(shapes, live=True)
    passes = {"ig": args.m_steps + 1}.get(wl, 1)
    kernels = []
    for name, (ms, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
        k = {"name": name, "ms_per_step": ms / prof_steps, "launches_per_step": cnt / prof_steps}
        if "_l" in name and name[-1].isdigit() and ("lstm" in name or "fwd" in name or "bwd" in name):
            layers = [0, 1] if name.endswith("_l1_l2") else [int(name[-1]) - 1]     # lstm_l1_l2: the fused first pair
            k["tflops_executed"] = sum(fl_live[i] for i in layers) * units * passes / (ms / prof_steps * 1e-3) / 1e12
            k["tflops_algorithmic"] = sum(fl_alg[i] for i in layers) * units * passes / (ms / prof_steps * 1e-3) / 1e12
        kernels.append(k)
    timed = [k for k in kernels if "tflops_executed" in k]
    total_kernel_ms = sum(k["ms_per_step"] for k in kernels)
    peaks = json.load(open(peaks_path)) if os.path.exists(peaks_path) else {}
    # a kernel timed inside a long step is held against the sustained figure (B200_PROFILING.md)
    tensor_peak = peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 2250.0))
    tensor_src = "measured sustained cuBLAS bf16 (MEASURED_PEAKS.json)" if peaks else "nominal dense bf16 (fallback)"
    sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
    n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
    # The LSTM layers run their gate contraction on tcgen05 with split-bf16 operands: three MMA passes
    # (hi*hi + hi*lo + lo*hi) per algorithmic product, so the tensor pipe executes 3x the fp32-equivalent
    # FLOPs; ig adds the backward contraction (same size), train still uses the fp32 CUDA-core kernels.
    lstm_ms = sum(k["ms_per_step"] for k in timed)
    lstm_flops = sum(k["tflops_executed"] * k["ms_per_step"] for k in timed) * 1e9   # FLOP per step
    on_tensor = wl != "train"
    tensor = {"kernel": "lstm layers (%d launches per step)" % round(sum(k["launches_per_step"] for k in timed)),
              "achieved": (3 if on_tensor else 0) * lstm_flops / (lstm_ms * 1e-3) / 1e12 if lstm_ms else None, "peak": tensor_peak,
              "unit": "TFLOP/s", "peak_source": tensor_src, "fp32_equivalent_tflops": lstm_flops / (lstm_ms * 1e-3) / 1e12 if lstm_ms else None,
              "fp32_cuda_core_peak_tflops": fp32.value, "ms_per_step": lstm_ms, "share_of_kernel_time": lstm_ms / total_kernel_ms}
    tensor["frac"] = tensor["achieved"] / tensor_peak if tensor["achieved"] else None
    # what actually bounds those kernels: the cell update's transcendentals (5 ex2 + 2 rcp per unit and step
    # forward, 1 ex2 + 1 rcp backward) on the MUFU pipe, 16 lanes per clock and SM
    units_steps = sum(LIVE_STEPS[i] * h for i, (_, _, h) in enumerate(LSTM_SHAPES["cml_gcn"]))
    if wl == "soilnet_gcn":
        units_steps = sum(t_ * h_ for t_, h_ in [(324, 16), (324, 16), (108, 32), (108, 32), (36, 64), (36, 64), (12, 128)])
    mufu_per_unit = {"ig": 7 + 2, "train": 10 + 2}.get(wl, 7)
    mufu_rate = units_steps * mufu_per_unit * units * passes / (lstm_ms * 1e-3) if lstm_ms else 0.0
    mufu_peak = 16.0 * n_sm * sm_mhz * 1e6
    xu = {"achieved": mufu_rate / 1e12, "peak": mufu_peak / 1e12, "unit": "T transcendental op/s", "frac": mufu_rate / mufu_peak,
          "peak_source": "16 lanes/clk/SM x %d SMs x %.0f MHz (median SM clock sampled in this run)" % (n_sm, sm_mhz)}
    top = kernels[0]
    bpu = {"cml_gcn": bytes_per_window("cml_gcn"), "cml_series": 16 + 8 + 4, "cml_baseline": bytes_per_window("cml_baseline"), "soilnet_gcn": 4 * 3 * 337 + 4,
           "ig": 2 * bytes_per_window("cml_gcn"), "train": bytes_per_window("cml_gcn") + 4}[wl]
    step_hbm = {"achieved": bpu * units / (ms_step * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": bpu * units / (ms_step * 1e-3) / 1e9 / hbm_peak, "peak_source": hbm_src,
                "note": "whole step, algorithmic bytes (inputs + outputs) per unit: %d" % bpu}
    # Algorithmic bytes of the graph front-end kernels per window: live steps only (the 19 dead steps are never read).
    live0 = LIVE_STEPS[0]
    fe_bytes = {"cml_frontend": 8 * live0 * (N_NODES + 1) + 4 * N_NODES + 8 + 4 * 18 * live0,
                "ig_frontend_bwd": 3 * 8 * live0 * (N_NODES + 1) + 4 * 18 * live0 + 4 * N_NODES}
    fe_roof = None
    fe_k = next((k for k in kernels if k["name"] in fe_bytes), None) if wl in ("cml_gcn", "ig") else None
    if fe_k is not None:
        per_launch = fe_bytes[fe_k["name"]] * units
        ach = per_launch * fe_k["launches_per_step"] / (fe_k["ms_per_step"] * 1e-3) / 1e9
        fe_roof = {"bound": "hbm", "kernel": fe_k["name"], "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                   "traffic": None, "peak_source": hbm_src,
                   "note": "%.0f%% of kernel time; algorithmic bytes per launch %d (live steps of the features, coefficients, "
                           "output tile)" % (100 * fe_k["ms_per_step"] / total_kernel_ms, per_launch)}
    if fe_roof is not None and top["name"] == fe_roof["kernel"]:
        roofline = dict(fe_roof)
    else:
        roofline = {"bound": "tensor", "kernel": tensor["kernel"], "achieved": tensor["achieved"], "peak": tensor_peak, "unit": "TFLOP/s",
                    "frac": tensor["frac"], "traffic": None, "peak_source": tensor_src,
                    "note": "split-bf16 (3 MMA passes) tcgen05 LSTM layers, %.0f%% of kernel time (largest single kernel: %s); the binding "
                            "unit is the MUFU pipe, see xu" % (100 * lstm_ms / total_kernel_ms, top["name"])}
    if fe_roof is not None:
        roofline["frontend"] = fe_roof
    roofline["tensor"] = tensor
    roofline["xu"] = xu
    roofline["hbm"] = step_hbm

    cpu = None
    if not args.no_cpu_baseline and wl in ("cml_gcn", "cml_baseline"):
        sample = 1024 if wl == "cml_gcn" else 16384
        wps, cores, secs = cpu_oracle_throughput(wl, sample, repeats=2)
        cpu = {"value": wps, "unit": "windows/s", "cores": cores, "kind": "port",
               "sample": "%d windows in batches of 128, best of 2 after warm-up (%.1f s each); reference TF CPU path unavailable on this box" % (sample, secs)}

    config = {"workload": wl, "model": model_name, "units_per_gpu_per_step": units, "seq_len": T,
              "nodes_per_window": {"cml_baseline": 1, "soilnet_gcn": 210}.get(wl, N_NODES), "input_bytes_per_step": in_bytes,
              "l2_policy": "inputs plus inter-layer tiles per step (%.0f MB in, GBs of tiles) exceed the 126 MB L2" % (in_bytes / 1e6),
              "parallelism": "windows sharded, no collective" if wl != "train" else "data parallel, NCCL all-reduce of moments + gradients"}
    config.update(cfg_extra)
    line = {
        "metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config,
        "e2e": {"value": e2e_value, "unit": unit, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "path": e2e_path},
        "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "kernels": kernels, "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
