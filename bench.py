#!/usr/bin/env python
"""bench.py - the hot path of GNN-XAI-TimeSeries-QualityControl on N B200s (BASELINE.json's metric:
windows scored/s and IG attributions/s; % of roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Prints ONE JSON line.  Without --workload the line is the CML GCN scoring workload (BASELINE.json config 1
at throughput batch size) and every other workload of the metric is measured in the same run and nested
under "workloads": cml_baseline (config 2), soilnet_gcn (config 3), cml_series (config 1 with the windows
built on the device from a resident network), ig (config 4, 50 interpolation steps), train_b128 and
train_b4096 (config 5, data parallel with NCCL all-reduce, per-GPU batch 128 = the reference's and 4,096).

One "step" = `passes_per_step` passes of the workload over one batch of synthetic input; the number of passes
is chosen after warm-up so that the K timed steps last at least 0.5 s (clock samples, sustained power).
  value     units/s, whole job, inputs resident in HBM, CUDA events around the K steps, max over ranks
  e2e       the same metric through the host-buffer call: pinned HOST inputs, H2D + kernels + D2H timed
  roofline  the LSTM stack (the dominant kernel family): ALGORITHMIC FLOPs (2*MACs of every
            matmul at the full sequence length) / summed kernel time (CUDA events per launch) against the
            measured bf16 peak; `frac_executed` = what the tensor pipe really executes (3 split passes over
            the live steps); nested: the single largest launch, the graph front-end against HBM on SURVEY
            8(d)'s input bytes, the MUFU pipe, and the whole step against HBM
  parity    max abs / rel error of the first units of the timed batch against the float64 CPU oracle
  cpu_baseline  the CPU oracle (torch fp32, all host cores) on a bounded sample: kind "port", because the
            reference's TensorFlow path is not installable here (no network; not in the wheelhouse)
--impl reference times that CPU oracle alone, on the same config; it never loads libgnnqc.so.
"""
import argparse
import ctypes
import importlib.util
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "gnn-xai-timeseries-qualitycontrol_b200")
MODELS = os.path.join(ROOT, "tests", "golden", "models")
T_CML, T_SOIL, N_NODES, N_SOIL = 181, 337, 23, 210
MIN_TIMED_SECONDS = float(os.environ.get("GNNQC_BENCH_MIN_SECONDS", "0.5"))   # profiling runs under ncu set it to 0

# One full wave of the LSTM kernels on a B200: 148 SMs x 4 window groups x 128 windows.  The recurrences are serial chains
# per group, so a pass over fewer windows (65,536 = 512 of the 592 groups) takes the same time; throughput batches are
# sized to the wave.
WAVE = 148 * 4 * 128   # 75,776
WORKLOADS = {
    # name: (metric, unit, model, units per GPU per pass)
    "cml_gcn": ("cml_windows_scored_per_sec", "windows/s", "model_cml", WAVE),
    "cml_baseline": ("cml_baseline_windows_scored_per_sec", "windows/s", "model_cml_baseline", WAVE),
    "soilnet_gcn": ("soilnet_sensor_sequences_scored_per_sec", "sequences/s", "model_soilnet", 360 * N_SOIL),
    "cml_series": ("cml_windows_scored_per_sec_from_resident_series", "windows/s", "model_cml", WAVE),
    "ig": ("cml_ig_attributions_per_sec", "attributions/s", "model_cml", WAVE),
    "ig_series": ("cml_ig_attributions_per_sec_from_resident_series", "attributions/s", "model_cml", 72000),
    "train": ("cml_gcn_training_windows_per_sec", "windows/s", "model_cml", 4096),
}
NESTED = ["cml_baseline", "soilnet_gcn", "cml_series", "ig", "ig_series", "train_b128", "train_b4096"]

# ---- algorithmic work per unit ----------------------------------------------------
LSTM = {  # (T, in, H) per layer at the full sequence length
    "cml_gcn": [(181, 18, 16), (181, 16, 16), (60, 16, 32), (60, 32, 32), (20, 32, 64), (20, 64, 64), (6, 64, 128)],
    "cml_baseline": [(181, 2, 16), (181, 16, 16), (60, 16, 32), (60, 32, 32), (20, 32, 64), (20, 64, 64), (6, 64, 128)],
    "soilnet_gcn": [(337, 19, 16), (337, 16, 16), (112, 16, 32), (112, 32, 32), (37, 32, 64), (37, 64, 64), (12, 64, 128)],
}
LIVE = {181: [162, 162, 54, 54, 18, 18, 6], 337: [324, 324, 108, 108, 36, 36, 12]}   # steps that can reach the score
HEAD_FLOPS = 2 * (128 * 64 + 64 * 64 + 64)


def lstm_flops(kind, live=False):
    T = LSTM[kind][0][0]
    return [2 * (LIVE[T][i] if live else t) * (inp + h) * 4 * h for i, (t, inp, h) in enumerate(LSTM[kind])]


def shapes_of(wl):
    return {"cml_baseline": "cml_baseline", "soilnet_gcn": "soilnet_gcn"}.get(base_name(wl), "cml_gcn")


def base_name(wl):
    return "train" if wl.startswith("train") else wl


def alg_flops_per_unit(wl, m_steps=50):
    """LSTM stack + head forward; IG = (m+1) x (forward + input gradient); training = 3 x forward."""
    fwd = sum(lstm_flops(shapes_of(wl))) + HEAD_FLOPS
    return {"ig": (m_steps + 1) * 2 * fwd, "ig_series": (m_steps + 1) * 2 * fwd, "train": 3 * fwd}.get(base_name(wl), fwd)


def alg_bytes_per_unit(wl):
    """compulsory HBM traffic with materialised windows."""
    cml = 4 * 2 * T_CML * (N_NODES + 1) + 4 * N_NODES + 4                      # 34,848 B
    return {"cml_gcn": cml, "cml_baseline": 4 * 2 * T_CML + 4, "soilnet_gcn": 4 * 3 * T_SOIL + 4, "cml_series": 16 + 8 + 4,
            "ig": 2 * cml, "ig_series": 16 + 8 + 4 + 4 * 2 * T_CML * (32 + 1), "train": cml + 4}[base_name(wl)]


def make_config(wl, units, world=1, m_steps=50):
    """The workload description; identical for the GPU arm and the reference arm."""
    b = base_name(wl)
    cfg = {"workload": wl, "model": WORKLOADS[b][2], "units_per_gpu_per_pass": units, "seq_len": T_SOIL if b == "soilnet_gcn" else T_CML,
           "nodes_per_window": {"cml_baseline": 1, "soilnet_gcn": N_SOIL}.get(b, N_NODES),
           "l2_policy": "no flush: inputs plus inter-layer tiles of one pass exceed the 126 MB L2 many times over",
           "parallelism": "data parallel: NCCL all-reduce of 6 f64 moments + the flat gradient vector per step" if b == "train"
           else "windows sharded over ranks, no collective"}
    if b in ("ig", "ig_series"):
        cfg["m_steps"] = m_steps
    if b == "train":
        cfg.update({"class_weights": [1, 5], "optimizer": "adam", "learning_rate": 1e-4, "positives": "1/6"})
    if b == "soilnet_gcn":
        cfg["windows_per_gpu_per_pass"] = units // N_SOIL
    return cfg


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm": p["hbm_gbs"], "bf16": p.get("bf16_tflops", 2250.0), "bf16_sustained": p.get("bf16_tflops_sustained", p.get("bf16_tflops", 2250.0)),
                "sm_max_mhz": p.get("sm_max_mhz", 1965.0), "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm": 6650.0, "bf16": 2250.0, "bf16_sustained": 2250.0, "sm_max_mhz": 1965.0, "source": "fallback (B200_PROFILING.md)"}


def load_module(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index):
        self.samples, self.proc, self.gpu = [], None, gpu_index

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
        time.sleep(0.1)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, power, reasons = [], None, [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            parts = [p.strip() for p in s.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
                power.append(float(parts[2]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm),
                "power_w_max": max(power) if power else None}


# ---- the CPU oracle as the timed baseline / the reference arm -----------------------------------------------
class CpuOracle:
    """The CPU oracle (oracle/gnnqc_oracle.py, torch fp32, all host threads) on a bounded sample of a
    workload, fed in batches of 128 windows - the reference's batch_size (preprocessing_config.yml:5).
    Touches neither libgnnqc.so nor the gnnqc package: the synthetic inputs come from gnnqc/synth.py loaded
    as a plain file, the weights from the oracle's own checkpoint reader."""

    BATCH = 128

    def __init__(self, wl, sample_units, m_steps=50):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import torch

        import gnnqc_oracle as O
        synth = load_module("gnnqc_synth_standalone", os.path.join(PKG, "gnnqc", "synth.py"))
        self.torch, self.O = torch, O
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        b = base_name(wl)
        self.no_grad = b not in ("ig", "ig_series", "train")
        w = O.load_weights(os.path.join(MODELS, WORKLOADS[b][2]), torch.float32)
        named = O.load_named(os.path.join(MODELS, WORKLOADS[b][2]))
        self.calls, self.units = [], 0
        if b == "soilnet_gcn":
            n_win = max(1, sample_units // N_SOIL)
            for i in range(n_win):
                batch = synth.make_soilnet_batch(1, seed=44 + i, n_nodes=N_SOIL)
                feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 3)
                self.calls.append((O.gcn_forward_soilnet, (torch.tensor(feat), torch.tensor(idx), torch.tensor(samp), 1, T_SOIL, w)))
                self.units += N_SOIL
            self.sample = "%d windows of %d sensors, one window per call" % (n_win, N_SOIL)
            return
        per_call = {"ig": 2, "ig_series": 2, "train": 128}.get(b, self.BATCH)
        for i, lo in enumerate(range(0, sample_units, per_call)):
            n = min(per_call, sample_units - lo)
            self.units += n
            if b == "cml_baseline":
                self.calls.append((O.baseline_forward, (synth.make_seq_batch(n, 2, T_CML, seed=44 + i), w)))
                continue
            batch = synth.make_cml_batch(n, seed=44 + i, n_nodes=N_NODES)
            feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(batch), 2)
            args = (torch.tensor(feat), batch["anom"], torch.tensor(idx), torch.tensor(samp))
            if b in ("ig", "ig_series"):
                self.calls.append((O.integrated_gradients_cml, args + (w, m_steps)))
            elif b == "train":
                labels = (torch.rand(n, generator=torch.Generator().manual_seed(i)) < 1.0 / 6.0).long()
                self.calls.append((O.train_grads_cml, args + (labels, named, (1.0, 5.0), torch.float32)))
            else:
                self.calls.append((O.gcn_forward_cml, args + (w,)))
        self.sample = "%d %s in batches of %d" % (self.units, "windows", per_call)
        if b in ("cml_series", "ig_series"):
            self.sample += " (materialised windows: the oracle has no resident-series path)"

    def step(self):
        t0 = time.perf_counter()
        if self.no_grad:
            with self.torch.no_grad():
                for fn, args in self.calls:
                    fn(*args)
        else:
            for fn, args in self.calls:
                fn(*args)
        return time.perf_counter() - t0


CPU_SAMPLE = {"cml_gcn": 1024, "cml_baseline": 8192, "soilnet_gcn": 2 * N_SOIL, "cml_series": 512, "ig": 4, "ig_series": 4, "train": 256}
REF_SAMPLE = {"cml_gcn": 2048, "cml_baseline": 16384, "soilnet_gcn": 4 * N_SOIL, "cml_series": 2048, "ig": 8, "ig_series": 8, "train": 512}


def cpu_baseline(wl, unit, m_steps):
    o = CpuOracle(wl, CPU_SAMPLE[base_name(wl)], m_steps)
    o.step()                                   # warm-up (thread pool, allocator)
    secs = min(o.step() for _ in range(1 if base_name(wl) in ("ig", "ig_series") else 2))
    return {"value": o.units / secs, "unit": unit, "cores": o.cores, "kind": "port",
            "sample": "%s, best pass after warm-up (%.1f s); reference TF CPU path unavailable on this box" % (o.sample, secs)}


def run_reference(args, rank):
    """--impl reference: the reference's own implementation (TensorFlow/Keras/spektral) cannot be installed in
    this image, so - as the tier rules say - this arm times the CPU oracle port on the host cores, on the GPU
    arm's config; each step is a bounded sample of the workload's pass (a full pass of 75,776 windows would take
    two minutes per step on these cores)."""
    if rank != 0:
        return
    wl = args.workload or "cml_gcn"
    b = base_name(wl)
    metric, unit, _, units = WORKLOADS[b]
    units = args.windows or units
    o = CpuOracle(wl, REF_SAMPLE[b], args.m_steps)
    times = []
    for i in range(args.warmup + args.steps):
        dt = o.step()
        if i >= args.warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    value = o.units / (ms * 1e-3)
    line = {
        "impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": make_config(wl, units, 1, args.m_steps),
        "sample_note": "each step scores a bounded sample (%s) of the config's %d units per pass; the rate per unit is what compares" % (o.sample, units),
        "cpu_baseline": {"value": value, "unit": unit, "cores": o.cores, "kind": "port",
                         "sample": o.sample + " per step; CPU oracle port (torch fp32) - reference TF/Keras CPU path unavailable on this box"},
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ---- NUMA placement of the host staging buffers ----------------------------------------------------------------
def numa_node_from_nvml(local_rank, info):
    """The NUMA node that holds most of the CPUs NVML names as close to this GPU (None if it cannot be told)."""
    import glob
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
    words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
    cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
    info["nvml_cpus"] = len(cpus)
    best, best_n = None, 0
    nodes = glob.glob("/sys/devices/system/node/node[0-9]*")
    info["numa_nodes_visible"] = len(nodes)
    for path in nodes:
        members = set()
        for part in open(os.path.join(path, "cpulist")).read().strip().split(","):
            if part:
                lo, _, hi = part.partition("-")
                members.update(range(int(lo), int(hi or lo) + 1))
        n = len(members & cpus)
        if n > best_n:
            best, best_n = int(os.path.basename(path)[4:]), n
    return best if len(nodes) > 1 else None


def bind_near_gpu(torch, local_rank):
    """Pinned staging buffers should live on the NUMA node the GPU hangs off.  Returns what was found / done."""
    info = {"numa_node": None, "cpus_allowed": len(os.sched_getaffinity(0))}
    try:
        props = torch.cuda.get_device_properties(local_rank)
        bus = "%04x:%02x:%02x.0" % (getattr(props, "pci_domain_id", 0), props.pci_bus_id, props.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read().strip())
        info["numa_node"] = node
        if node < 0:   # containers often hide the PCI device's node: ask NVML for the GPU's CPU set and look up the node of those CPUs
            node = numa_node_from_nvml(local_rank, info)
            if node is None or node < 0:
                return info
            info["numa_node"] = node
        # prefer this node for the pages of the pinned buffers allocated from here on (MPOL_PREFERRED = 1); the CPU
        # affinity is left alone (the CPU baseline and the parity oracle use every core)
        mask = ctypes.c_ulong(1 << node)
        rc = ctypes.CDLL(None, use_errno=True).syscall(238, 1, ctypes.byref(mask), ctypes.c_ulong(8 * ctypes.sizeof(mask)))
        info["mempolicy"] = "preferred node %d" % node if rc == 0 else "set_mempolicy failed (errno %d)" % ctypes.get_errno()
    except Exception as e:   # sysfs layout / permissions differ between boxes; placement is best effort
        info["error"] = str(e)[:80]
    return info


# ---- GPU workloads ---------------------------------------------------------------------------------------------
class Workload:
    """One workload on this rank: device-resident step, host-buffer step, parity check."""

    def __init__(self, wl, ctx, args):
        torch, synth = ctx.torch, ctx.synth
        from gnnqc import NativeModel, dp
        self.wl, self.ctx = wl, ctx
        b = self.base = base_name(wl)
        self.metric, self.unit, self.model_name, units = WORKLOADS[b]
        if wl == "train_b128":
            units = 128
        if args.windows and not ctx.nested:
            units = args.windows
        self.m_steps = args.m_steps
        dev, rank, world = ctx.dev, ctx.rank, ctx.world
        self.model = model = NativeModel.load(os.path.join(MODELS, self.model_name), device=ctx.local_rank)
        self.T = T_SOIL if b == "soilnet_gcn" else T_CML
        self.h2d = self.d2h = 0
        self.e2e_path = None
        self.extra = {}
        self.check_e2e = None
        B = units
        if b in ("cml_gcn", "ig", "train"):
            batch = self.batch = synth.make_cml_batch(B, seed=44 + rank, n_nodes=N_NODES, device=str(dev))
            ntot = B * N_NODES
            keys = ("features", "anom", "edges", "node_off", "edge_off")
            self.in_bytes = sum(batch[k].numel() * batch[k].element_size() for k in keys)
            if b == "cml_gcn":
                self.step = lambda: model.score_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"], ntot)
                host = {k: batch[k].cpu().pin_memory() for k in keys}
                out_host = torch.empty(B, dtype=torch.float32).pin_memory()

                def e2e():
                    out, moved = model.score_cml_host(host["features"], host["anom"], host["edges"], host["node_off"], host["edge_off"],
                                                      chunk_windows=args.chunk_windows, out=out_host)
                    self.h2d, self.d2h = moved, out.numel() * 4
                    return out
                self.e2e_step = e2e
                self.check_e2e = True
                self.e2e_path = "gnnqc_score_cml_host: pinned host -> chunked H2D of the live timesteps on a copy stream -> scoring kernels -> D2H"
            elif b == "ig":
                self.step = lambda: model.ig_cml(batch["features"], batch["anom"], batch["edges"], batch["node_off"], batch["edge_off"], self.m_steps, ntot)[2]
                host = {k: batch[k].cpu().pin_memory() for k in keys}
                pinned = {}

                def to_host(name, t):
                    if name not in pinned:
                        pinned[name] = torch.empty(t.shape, dtype=t.dtype).pin_memory()
                    pinned[name].copy_(t, non_blocking=True)
                    return pinned[name]

                def e2e():
                    d = {k: host[k].to(dev, non_blocking=True) for k in keys}
                    af, aa, p = model.ig_cml(d["features"], d["anom"], d["edges"], d["node_off"], d["edge_off"], self.m_steps, ntot)
                    res = to_host("af", af), to_host("aa", aa), to_host("p", p)
                    torch.cuda.synchronize()
                    return res[2]
                self.e2e_step = e2e
                self.h2d = self.in_bytes
                self.d2h = batch["features"].numel() * 4 + batch["anom"].numel() * 4 + B * 4
                self.e2e_path = "pinned host -> H2D -> gnnqc_ig_cml -> D2H of both attribution arrays and the scores"
            else:
                labels = self.labels = (torch.rand(B, device=dev, generator=torch.Generator(device=dev).manual_seed(7 + rank)) < 1.0 / 6.0).long()
                self.extra.update({"global_batch": B * world, "reduction_vector_floats": model.train_vector_len()})
                self.comm_ms = []

                def train(d, lab, timed=False):
                    mom = model.train_moments_cml(d["features"])
                    if timed:
                        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
                        ev[0].record()
                    dp.all_reduce_sum(mom)
                    if timed:
                        ev[1].record()
                    rvec = model.train_grads_cml(d["features"], d["anom"], d["edges"], d["node_off"], d["edge_off"], lab, mom, (1.0, 5.0),
                                                 global_batch=B * world, n_nodes_total=ntot)
                    if timed:
                        ev[2].record()
                    dp.all_reduce_sum(rvec)
                    if timed:
                        ev[3].record()
                    model.train_apply(rvec, mom, 1e-4)
                    if timed:
                        torch.cuda.synchronize()
                        self.comm_ms.append(ev[0].elapsed_time(ev[1]) + ev[2].elapsed_time(ev[3]))
                    return rvec[-1:]
                self.train = train
                self.step = lambda: train(batch, labels)
                host = {k: batch[k].cpu().pin_memory() for k in keys}
                host_labels = labels.cpu().pin_memory()

                copy_stream, pending = torch.cuda.Stream(dev), []

                def upload():   # as libs/fit_model.py's epoch loop: the next step's batch travels on a copy stream while this step runs
                    with torch.cuda.stream(copy_stream):
                        d = {k: host[k].to(dev, non_blocking=True) for k in keys}
                        lab = host_labels.to(dev, non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(copy_stream)
                    return d, lab, ev

                def e2e():
                    d, lab, ev = pending.pop() if pending else upload()
                    pending.append(upload())
                    cur = torch.cuda.current_stream(dev)
                    cur.wait_event(ev)
                    for v in list(d.values()) + [lab]:
                        v.record_stream(cur)
                    return train(d, lab).cpu()
                self.e2e_step = e2e
                self.h2d, self.d2h = self.in_bytes + B * 8, 4
                self.e2e_path = ("pinned host -> H2D of the NEXT step's batch and labels on a copy stream (one upload per step) | gnnqc_train_moments_cml / all-reduce / "
                                 "gnnqc_train_grads_cml / all-reduce / gnnqc_train_apply -> D2H of the loss")
        elif b == "cml_series":
            from gnnqc.series import CmlSeries
            np = ctx.np
            S = 2048
            per = max(1, B // S)
            net = self.net = synth.make_cml_network(S, 4096, seed=42 + rank)
            series = self.series = CmlSeries(net["raw"], net["rmed"], net["nbr_off"], net["nbr_idx"], net["edge_off"], net["edges"], device=ctx.local_rank)
            units = B = S * per
            tgt_h = torch.arange(S, dtype=torch.int32).repeat_interleave(per).pin_memory()
            tau_h = (2000 + torch.arange(per, dtype=torch.int32)).repeat(S).pin_memory()
            tgt_d, tau_d = tgt_h.to(dev), tau_h.to(dev)
            self.step = lambda: model.score_cml_series(series, tgt_d, tau_d, 120, 60)
            self.in_bytes = B * 8
            slab_raw = np.ascontiguousarray(net["raw"][:, 2060: 2060 + per])
            slab_med = np.ascontiguousarray(net["rmed"][:, 2060: 2060 + per])

            pin_raw, pin_med = torch.from_numpy(slab_raw).pin_memory(), torch.from_numpy(slab_med).pin_memory()

            def e2e():   # the samples that are new for this pass's windows arrive from the host, then the window list
                series.update(2060, pin_raw, pin_med, stream=torch.cuda.current_stream().cuda_stream)   # gnnqc_series_update_async
                return model.score_cml_series(series, tgt_h.to(dev, non_blocking=True), tau_h.to(dev, non_blocking=True), 120, 60).cpu()
            self.e2e_step = e2e
            self.h2d, self.d2h = int(slab_raw.nbytes * 2 + B * 8), B * 4
            self.e2e_path = "gnnqc_series_update_async (new samples, pinned host slabs) + window list H2D -> gnnqc_score_cml_series -> D2H"
            self.extra.update({"sensors": S, "consecutive_windows_per_sensor": per, "resident_series_bytes": int(net["raw"].nbytes * 2),
                               "mean_neighbourhood": float(np.diff(net["nbr_off"]).mean())})
        elif b == "ig_series":
            # BASELINE.json config 4: the IG sweep over a 4,000-link network that stays resident in HBM, sharded by target
            # link over the ranks (each rank holds its links plus their neighbour halo; no collective).  The time extent is
            # cut from one year (525,600 samples) to 16,384 samples per link so that the host can synthesise it in seconds:
            # the resident arrays scale linearly with it (full year: 16.8 GB raw + 16.8 GB medians + 8.4 GB validity
            # prefix, inside one B200's 180 GB) and a window only ever touches 181 consecutive samples.
            from gnnqc.series import CmlSeries, shard_network_by_link
            np = ctx.np
            S, Tt = 4000, 16384
            net = synth.make_cml_network(S, Tt, seed=42)
            sh = shard_network_by_link(net, rank, world)
            del net
            series = self.series = CmlSeries(sh["raw"], sh["rmed"], sh["nbr_off"], sh["nbr_idx"], sh["edge_off"], sh["edges"], device=ctx.local_rank)
            owned = torch.tensor(sh["owned_local"], dtype=torch.int32)
            per = max(1, B // owned.numel())
            units = B = owned.numel() * per
            tgt_h = owned.repeat_interleave(per).pin_memory()
            tau_h = (8000 + torch.arange(per, dtype=torch.int32)).repeat(owned.numel()).pin_memory()
            tgt_d, tau_d = tgt_h.to(dev), tau_h.to(dev)
            self.step = lambda: model.ig_cml_series(series, tgt_d, tau_d, 120, 60, self.m_steps)[2]
            self.in_bytes = B * 8
            pinned = {}

            def to_host(name, t):
                if name not in pinned:
                    pinned[name] = torch.empty(t.shape, dtype=t.dtype).pin_memory()
                pinned[name].copy_(t, non_blocking=True)
                return pinned[name]

            def e2e():   # window list up, both attribution arrays, the node lists and the scores down
                an, aa, p, sens, cnt = model.ig_cml_series(series, tgt_h.to(dev, non_blocking=True), tau_h.to(dev, non_blocking=True), 120, 60, self.m_steps)
                res = to_host("an", an), to_host("aa", aa), to_host("p", p), to_host("sens", sens), to_host("cnt", cnt)
                torch.cuda.synchronize()
                return res[2]
            self.e2e_step = e2e
            self.h2d = B * 8
            self.d2h = B * (T_CML * 33 * 8 + 4 + 32 * 4 + 4)
            self.e2e_path = "window list H2D -> gnnqc_ig_cml_series on the resident network -> D2H of attributions [B,T,32,2] + [B,T,2], node lists, scores"
            self.extra.update({"links_total": S, "links_owned_by_rank0": int(owned.numel()), "links_resident_on_rank0_incl_halo": int(sh["global_ids"].size),
                               "samples_per_link": Tt, "consecutive_windows_per_link": per,
                               "resident_bytes": int(sh["raw"].nbytes * 2 + sh["raw"].shape[0] * (Tt + 1) * 4),
                               "mean_neighbourhood": float(np.diff(sh["nbr_off"])[sh["owned_local"]].mean()),
                               "note": "time extent cut from 525,600 to 16,384 samples per link (host synthesis time); throughput does not depend on it"})
        elif b == "cml_baseline":
            from gnnqc.host_pipeline import SeqHostScorer
            x = self.x = synth.make_seq_batch(B, 2, T_CML, seed=44 + rank, device=str(dev))
            self.step = lambda: model.score_seq(x)
            self.in_bytes = x.numel() * 4
            host_x = x.cpu().pin_memory()
            scorer = SeqHostScorer(model)   # chunks of one wave of the LSTM kernels

            def e2e():
                out = scorer.score(host_x)
                self.h2d, self.d2h = scorer.last_h2d_bytes, scorer.last_d2h_bytes
                return out
            self.e2e_step = e2e
            self.check_e2e = True
            self.e2e_path = "SeqHostScorer: pinned host -> chunked H2D on a copy stream -> gnnqc_score_seq -> D2H"
        else:   # soilnet_gcn: whole windows of 210 sensors, T = 337
            nw = B // N_SOIL
            batch = self.batch = synth.make_soilnet_batch(nw, seed=44 + rank, n_nodes=N_SOIL, device=str(dev))
            units = ntot = nw * N_SOIL
            keys = ("features", "edges", "node_off", "edge_off")
            self.in_bytes = sum(batch[k].numel() * batch[k].element_size() for k in keys)
            self.step = lambda: model.score_soilnet(batch["features"], batch["edges"], batch["node_off"], batch["edge_off"], T_SOIL, ntot)
            host = {k: batch[k].cpu().pin_memory() for k in keys}

            def e2e():
                d = {k: host[k].to(dev, non_blocking=True) for k in keys}
                return model.score_soilnet(d["features"], d["edges"], d["node_off"], d["edge_off"], T_SOIL, ntot).cpu()
            self.e2e_step = e2e
            self.check_e2e = True
            self.h2d, self.d2h = self.in_bytes, ntot * 4
            self.e2e_path = "pinned host -> H2D -> gnnqc_score_soilnet -> D2H"
        self.units = units

    # ---- parity of the timed batch against the float64 oracle (rank 0, outside every timed region) ----
    def parity(self, result):
        torch, synth, np = self.ctx.torch, self.ctx.synth, self.ctx.np
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import gnnqc_oracle as O
        b = self.base
        w = O.load_weights(os.path.join(MODELS, self.model_name), torch.float64)
        torch.set_num_threads(os.cpu_count() or 1)
        t0 = time.perf_counter()
        if b == "ig_series":
            return None      # bit-identical to gnnqc_ig_cml on the materialised windows (tests/test_gpu_series.py), which the "ig" workload checks
        if b in ("cml_gcn", "ig", "train"):
            n = {"cml_gcn": 1024 if self.ctx.world == 1 else 256, "ig": 1, "train": min(128, self.units)}[b]   # rank 0 checks while the other ranks wait
            eo, no = self.batch["edge_off"][: n + 1].cpu(), self.batch["node_off"][: n + 1].cpu()
            cpu = {"features": self.batch["features"][: n * T_CML * N_NODES].cpu(), "anom": self.batch["anom"][:n].cpu(),
                   "edges": self.batch["edges"][: int(eo[-1])].cpu(), "node_off": no, "edge_off": eo, "T": T_CML}
            feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(cpu), 2)
            a = (torch.tensor(feat).double(), cpu["anom"].double(), torch.tensor(idx), torch.tensor(samp))
            if b == "cml_gcn":
                with torch.no_grad():
                    want = O.gcn_forward_cml(*a, w).reshape(-1).numpy()
                got = result[:n].cpu().numpy()
                what = "scores of the first %d windows of the timed batch" % n
            elif b == "ig":
                d = {k: v.to(self.ctx.dev) for k, v in cpu.items() if k != "T"}
                af, aa, p = self.model.ig_cml(d["features"], d["anom"], d["edges"], d["node_off"], d["edge_off"], self.m_steps, n * N_NODES)
                of, oa, op = O.integrated_gradients_cml(*a, w, m_steps=self.m_steps)
                got = np.concatenate([af.cpu().numpy().ravel(), aa.cpu().numpy().ravel(), p.cpu().numpy().ravel()])
                want = np.concatenate([of.numpy().ravel(), oa.numpy().ravel(), op.numpy().ravel()])
                what = "all attributions + score of the first window of the timed batch (%d values)" % got.size
            else:
                from gnnqc import NativeModel
                fresh = NativeModel.load(os.path.join(MODELS, self.model_name), device=self.ctx.local_rank)   # the timed model has moved on
                d = {k: v.to(self.ctx.dev) for k, v in cpu.items() if k != "T"}
                lab = self.labels[:n]
                mom = fresh.train_moments_cml(d["features"])
                rvec = fresh.train_grads_cml(d["features"], d["anom"], d["edges"], d["node_off"], d["edge_off"], lab, mom, (1.0, 5.0))
                oloss, ograds, _, _ = O.train_grads_cml(a[0], a[1], a[2], a[3], lab.cpu(), O.load_named(os.path.join(MODELS, self.model_name)))
                flat = rvec[: fresh.n_params].cpu().numpy()
                names = [s for s in fresh.segments() if s[0] in ograds and not s[0].startswith("gcn_layer")]
                got = np.concatenate([flat[o: o + int(np.prod(s))] for _, o, s, _ in names] + [rvec[-1:].cpu().numpy()])
                want = np.concatenate([ograds[p].numpy().ravel() for p, _, _, _ in names] + [np.array([oloss.item()])])
                fresh.close()
                what = "loss + LSTM/dense gradients of the first %d windows, shipped weights" % n
        elif b == "cml_baseline":
            n = 1024
            with torch.no_grad():
                want = O.baseline_forward(self.x[:n].cpu().double(), w).reshape(-1).numpy()
            got = result[:n].cpu().numpy()
            what = "scores of the first %d windows of the timed batch" % n
        elif b == "soilnet_gcn":
            nw = 5
            no, eo = self.batch["node_off"][: nw + 1].cpu(), self.batch["edge_off"][: nw + 1].cpu()
            cpu = {"features": self.batch["features"][: int(no[-1]) * T_SOIL].cpu(), "edges": self.batch["edges"][: int(eo[-1])].cpu(),
                   "node_off": no, "edge_off": eo, "T": T_SOIL}
            feat, idx, samp = O.reference_batch_layout(synth.cml_batch_to_windows(cpu), 3)
            with torch.no_grad():
                want = O.gcn_forward_soilnet(torch.tensor(feat).double(), torch.tensor(idx), torch.tensor(samp), nw, T_SOIL, w).reshape(-1).numpy()
            got = result[: int(no[-1])].cpu().numpy()
            what = "scores of the first %d windows (%d sensor-sequences) of the timed batch" % (nw, got.size)
        else:
            return None
        err = np.abs(got.astype(np.float64) - want)
        scale = np.abs(want).max() if b in ("ig", "train") else 1.0
        return {"max_abs_err": float(err.max()), "max_rel_err": float((err / np.maximum(np.abs(want), 1e-6 * max(scale, 1e-30))).max()),
                "max_abs_err_over_max_abs": float(err.max() / max(scale, 1e-30)), "oracle": "float64 CPU restatement (oracle/gnnqc_oracle.py), weights via oracle/ckpt_reader.py",
                "checked": what, "seconds": round(time.perf_counter() - t0, 1),
                "tolerance": "1e-4 abs/rel on scores (north_star); attributions/gradients: 1e-4 abs + 1e-3 rel of the float64 value"}

    def close(self):
        if hasattr(self, "series"):
            self.series.close()
        self.model.close()
        for k in ("batch", "x", "net", "labels"):
            if hasattr(self, k):
                delattr(self, k)
        self.step = self.e2e_step = None
        self.ctx.torch.cuda.empty_cache()


class Ctx:
    pass


_T0 = time.time()


def log(ctx, msg):
    """Progress on stderr (GNNQC_BENCH_VERBOSE=1): which rank is where, for diagnosing multi-rank runs."""
    if os.environ.get("GNNQC_BENCH_VERBOSE"):
        sys.stderr.write("[bench %6.1fs rank %d] %s\n" % (time.time() - _T0, ctx.rank, msg))
        sys.stderr.flush()


def measure(wl, ctx, args, steps, warmup, with_cpu=True):
    """Runs one workload; returns the line (rank 0) or None."""
    torch, dist, np = ctx.torch, ctx.dist, ctx.np
    log(ctx, "%s: building inputs" % wl)
    W = Workload(wl, ctx, args)
    log(ctx, "%s: warm-up" % wl)
    model, dev, world, rank = W.model, ctx.dev, ctx.world, ctx.rank
    units = W.units

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- warm-up, then size the step so that the timed region lasts >= MIN_TIMED_SECONDS -------------
    for _ in range(warmup):
        result = W.step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    W.step()
    e1.record()
    torch.cuda.synchronize()
    pass_s = max_over_ranks(e0.elapsed_time(e1) * 1e-3)
    passes = max(1, int(math.ceil(MIN_TIMED_SECONDS / (steps * pass_s))))
    barrier()
    log(ctx, "%s: timed region, %d passes per step" % (wl, passes))
    launches0 = model.launch_count()
    sampler = ClockSampler(ctx.local_rank)
    if rank == 0:
        sampler.start()
    e0.record()
    for _ in range(steps):
        for _ in range(passes):
            result = W.step()
    e1.record()
    barrier()
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / steps
    launches = model.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    value = world * units * passes / (ms_step * 1e-3)

    log(ctx, "%s: per-kernel profile" % wl)
    # ---- per-kernel times (CUDA events around every launch, on the launching stream) ---------------------
    prof_passes = 1 if W.base in ("ig", "ig_series") else max(1, min(passes * steps, 10))
    model.profile(True)
    for _ in range(prof_passes):
        W.step()
    prof = model.profile_read()
    model.profile(False)
    comm = None
    if W.base == "train":
        for _ in range(3):
            W.train(W.batch, W.labels, timed=True)
        comm = sum(W.comm_ms) / len(W.comm_ms)

    # ---- end to end from pinned host buffers ------------------------------------------------------------------
    log(ctx, "%s: end to end" % wl)
    for _ in range(2):
        out = W.e2e_step()
    barrier()
    e2e_passes = max(3, min(steps, 10))
    t0 = time.perf_counter()
    for _ in range(e2e_passes):
        out = W.e2e_step()
    torch.cuda.synchronize()
    dt_local = time.perf_counter() - t0
    dt = max_over_ranks(dt_local)
    e2e_value = world * units * e2e_passes / dt
    if W.check_e2e and not torch.equal(out.to(dev), result):
        raise SystemExit("%s: e2e scores differ from the device-resident scores" % wl)
    per_rank = torch.tensor([W.h2d * e2e_passes / dt_local / 1e9, float(ctx.placement.get("numa_node") if ctx.placement.get("numa_node") is not None else -9)],
                            device=dev, dtype=torch.float64)
    gathered = [per_rank.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(gathered, per_rank)
    log(ctx, "%s: parity / barrier" % wl)
    # at N > 1 only the primary workload is checked (the nested ones are checked by the N = 1 run; the other ranks wait here)
    parity = W.parity(result) if rank == 0 and not args.no_parity and (world == 1 or not ctx.nested) else None
    if world > 1:
        dist.barrier()
    log(ctx, "%s: measured" % wl)
    if rank != 0:
        W.close()
        return None

    # ---- roofline -----------------------------------------------------------------------------------------------
    pk = ctx.peaks
    shapes = shapes_of(wl)
    fl_alg, fl_live = lstm_flops(shapes), lstm_flops(shapes, live=True)
    n_pass = (W.m_steps + 1) if W.base in ("ig", "ig_series") else 1
    kernels, lstm = [], []
    for name, (ms, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
        k = {"name": name, "ms_per_pass": ms / prof_passes, "launches_per_pass": cnt / prof_passes}
        is_layer = "_l" in name and name[-1].isdigit() and any(s in name for s in ("lstm", "fwd", "bwd"))
        if is_layer:
            layers = [0, 1] if name.endswith("_l1_l2") else [int(name[-1]) - 1]
            sec = k["ms_per_pass"] * 1e-3
            k["tflops_algorithmic"] = sum(fl_alg[i] for i in layers) * units * n_pass / sec / 1e12
            k["tflops_executed_3pass"] = 3 * sum(fl_live[i] for i in layers) * units * n_pass / sec / 1e12
            k["frac_of_bf16_peak"] = k["tflops_algorithmic"] / pk["bf16"]
            lstm.append(k)
        kernels.append(k)
    total_ms = sum(k["ms_per_pass"] for k in kernels)
    lstm_ms = sum(k["ms_per_pass"] for k in lstm)
    sm_mhz = (clocks or {}).get("sm_mhz") or pk["sm_max_mhz"]
    n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
    fp32 = ctypes.c_double()
    ctx.lib.gnnqc_probe_fp32_peak(ctx.local_rank, 1, ctypes.byref(fp32))
    traffic = ctx.traffic.get(wl)
    if W.base in ("ig", "ig_series", "train"):
        # forward + backward (+ weight gradients): the unit's whole algorithmic FLOPs against the whole pass
        sec = ms_step / passes * 1e-3
        ach = alg_flops_per_unit(wl, W.m_steps) * units / sec / 1e12
        roofline = {"bound": "tensor", "kernel": "whole %s pass (%d launches)" % (W.base, round(sum(k["launches_per_pass"] for k in kernels))),
                    "achieved": ach, "peak": pk["bf16_sustained"], "unit": "TFLOP/s", "frac": ach / pk["bf16_sustained"],
                    "traffic": traffic, "peak_source": "sustained cuBLAS bf16, " + pk["source"],
                    "algorithmic_flops_per_unit": alg_flops_per_unit(wl, W.m_steps), "fp32_cuda_core_peak_tflops": fp32.value,
                    "frac_of_fp32_cuda_core_peak": ach / fp32.value if fp32.value else None,
                    "note": "SURVEY 8(d) FLOPs per unit x units / pass time; the contractions run on tcgen05 with split-bf16 operands "
                            "(3 MMA passes per product), the weight gradients included"}
        if W.base in ("ig", "ig_series"):
            saved = 140.5e3 * (W.m_steps + 1) * 2   # saved gates (12 B per cell, 11,712 cells per window) written + read per attribution
            roofline["saved_gate_stream"] = {"bound": "hbm", "achieved": saved * units / sec / 1e9, "peak": pk["hbm"], "unit": "GB/s",
                                             "frac": saved * units / sec / 1e9 / pk["hbm"]}
    else:
        sec = lstm_ms * 1e-3
        ach = sum(fl_alg) * units / sec / 1e12 if lstm_ms else None
        exe = 3 * sum(fl_live) * units / sec / 1e12 if lstm_ms else None
        roofline = {"bound": "tensor", "kernel": "LSTM stack: %s (%d launches per pass)" % (" + ".join(k["name"] for k in sorted(lstm, key=lambda k: k["name"])),
                                                                                         round(sum(k["launches_per_pass"] for k in lstm))),
                    "achieved": ach, "peak": pk["bf16"], "unit": "TFLOP/s", "frac": ach / pk["bf16"] if ach else None,
                    "frac_executed": exe / pk["bf16"] if exe else None, "achieved_executed": exe, "traffic": traffic,
                    "peak_source": "burst cuBLAS bf16 (kernels timed one by one with CUDA events), " + pk["source"],
                    "algorithmic_flops_per_unit": sum(fl_alg), "ms_per_pass": lstm_ms, "share_of_kernel_time": lstm_ms / total_ms if total_ms else None,
                    "frac_of_sustained_peak": ach / pk["bf16_sustained"] if ach else None,
                    "note": "algorithmic = 2*MACs of the 7 LSTM layers at the full sequence length (SURVEY 8(d)); executed = 3 split-bf16 MMA "
                            "passes over the steps that can reach the score; with 3 passes per product the algorithmic ceiling is 1/3"}
        if lstm:
            top = max(lstm, key=lambda k: k["ms_per_pass"])
            roofline["dominant_launch"] = {"kernel": top["name"], "ms": top["ms_per_pass"], "achieved": top["tflops_algorithmic"],
                                           "frac": top["frac_of_bf16_peak"], "frac_executed": top["tflops_executed_3pass"] / pk["bf16"]}
        # MUFU pipe: 5 ex2 + 2 rcp per unit and step on 16 lanes per clock and SM - the unit that bounds the cell update
        live = LIVE[LSTM[shapes][0][0]]
        unit_steps = sum(live[i] * h for i, (_, _, h) in enumerate(LSTM[shapes]))
        rate = unit_steps * 7 * units / sec if lstm_ms else 0.0
        peak = 16.0 * n_sm * sm_mhz * 1e6
        roofline["xu"] = {"achieved": rate / 1e12, "peak": peak / 1e12, "unit": "T transcendental op/s", "frac": rate / peak,
                          "peak_source": "16 lanes/clk/SM x %d SMs x %.0f MHz (median SM clock sampled in this run)" % (n_sm, sm_mhz)}
        fe = next((k for k in kernels if "frontend" in k["name"]), None)
        if fe is not None and W.base in ("cml_gcn", "soilnet_gcn"):
            bpu = alg_bytes_per_unit(wl)
            a = bpu * units / (fe["ms_per_pass"] * 1e-3) / 1e9
            roofline["frontend"] = {"bound": "hbm", "kernel": fe["name"], "achieved": a, "peak": pk["hbm"], "unit": "GB/s", "frac": a / pk["hbm"],
                                    "traffic": ctx.traffic.get(wl + ":frontend"), "share_of_kernel_time": fe["ms_per_pass"] / total_ms,
                                    "note": "SURVEY 8(d) input bytes per unit (%d) x units / launch time" % bpu}
    bpu = alg_bytes_per_unit(wl)
    hb = bpu * units * passes / (ms_step * 1e-3) / 1e9
    roofline["hbm"] = {"achieved": hb, "peak": pk["hbm"], "unit": "GB/s", "frac": hb / pk["hbm"], "peak_source": pk["source"],
                       "note": "whole pass, algorithmic bytes per unit: %d" % bpu}

    config = make_config(wl, units, world, W.m_steps)
    config.update(W.extra)
    line = {
        "metric": W.metric, "value": value, "unit": W.unit, "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16x3/f32acc", "data": "synthetic",
        "config": config, "passes_per_step": passes, "ms_per_pass": ms_step / passes, "input_bytes_per_pass": W.in_bytes,
        "e2e": {"value": e2e_value, "unit": W.unit, "h2d_bytes_per_step": int(W.h2d), "d2h_bytes_per_step": int(W.d2h), "passes": e2e_passes,
                "path": W.e2e_path, "h2d_gbs_per_rank": [round(float(g[0]), 2) for g in gathered],
                "gpu_numa_node_per_rank": [int(g[1]) if g[1] > -9 else None for g in gathered], "host_placement": ctx.placement},
        "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "parity": parity, "kernels": kernels,
        "dtype_note": "fp32 inputs/state/accumulators; LSTM gate contractions on tcgen05 with operands split into two 16-bit halves (hi*hi + hi*lo + lo*hi)",
    }
    if comm is not None:
        line["collectives"] = {"ms_per_step": comm, "share_of_step": comm / (ms_step / passes),
                               "what": "the two all-reduces (6 f64 moments, %d-float reduction vector) incl. waiting for the slowest rank" % model.train_vector_len()}
    if with_cpu and not args.no_cpu_baseline and world == 1:      # the CPU baseline is timed at N = 1 only
        line["cpu_baseline"] = cpu_baseline(wl, W.unit, W.m_steps)
    W.close()
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gnnqc", choices=["gnnqc", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS) + ["train_b128", "train_b4096"])
    ap.add_argument("--skip", default="", help="comma-separated nested workloads to leave out of the default run")
    ap.add_argument("--windows", type=int, default=0, help="units per GPU per pass (0 = workload default)")
    ap.add_argument("--m-steps", type=int, default=50, help="IG interpolation steps")
    ap.add_argument("--chunk-windows", type=int, default=8192, help="chunk size of the host-buffer scoring call")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--only-primary", action="store_true", help="skip the nested workloads")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if rank == 0:
        # torchrun exports OMP_NUM_THREADS=1 to its workers; rank 0 also runs the CPU oracle (parity check, CPU baseline,
        # the reference arm), which is meant to use every host core - set before torch is imported
        for var in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
            os.environ[var] = str(os.cpu_count() or 1)
    if args.impl == "reference":
        run_reference(args, rank)
        return
    args.warmup = max(args.warmup, 3)

    sys.path.insert(0, PKG)
    import numpy as np
    import torch
    import torch.distributed as dist

    from gnnqc import synth
    from gnnqc._lib import lib

    ctx = Ctx()
    ctx.torch, ctx.dist, ctx.np, ctx.synth, ctx.lib = torch, dist, np, synth, lib
    ctx.rank, ctx.world, ctx.local_rank = rank, world, local_rank
    torch.cuda.set_device(local_rank)
    ctx.placement = bind_near_gpu(torch, local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx.dev = torch.device("cuda", local_rank)
    ctx.peaks = load_peaks()
    ctx.nested = False
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")   # dram bytes per launch from the committed ncu --set full capture
    ctx.traffic = json.load(open(tpath)) if os.path.exists(tpath) else {}

    primary = args.workload or "cml_gcn"
    line = measure(primary, ctx, args, args.steps, args.warmup)
    if args.workload is None and not args.only_primary:
        ctx.nested = True
        nested = {}
        for wl in [w for w in NESTED if w not in args.skip.split(",")]:
            sub = measure(wl, ctx, args, max(2, min(args.steps, 5)), 3)
            if sub is not None:
                nested[wl] = sub
        if line is not None:
            line["workloads"] = nested
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
