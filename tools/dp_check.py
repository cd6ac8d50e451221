"""Run under torchrun with 2+ ranks: data-parallel training (NCCL all-reduce of the moments and the
reduction vector) must reproduce the single-GPU step."""
import os
import sys

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(ROOT, "gnn-xai-timeseries-qualitycontrol_b200"))
import numpy as np
import torch
import torch.distributed as dist

from gnnqc import NativeModel, synth
from libs import fit_model

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
path = os.path.join(ROOT, "tests", "golden", "models", "model_cml")
m = NativeModel.load(path, device=local)
batch = synth.make_cml_batch(128, seed=5, n_nodes=7)
labels = (torch.arange(128) % 5 == 0).long()
losses = [fit_model.train_step(m, batch, labels, {0: 1, 1: 5}, 1e-3) for _ in range(4)]
p_dp = m.get_params()
if rank == 0:
    dist_backup = dist.is_initialized
    ref = NativeModel.load(path, device=local)
    import gnnqc.dp as dpmod
    real_world = dpmod.world
    dpmod.world = lambda: (0, 1)
    real_ar = dpmod.all_reduce_sum
    dpmod.all_reduce_sum = lambda t: t
    ref_losses = [fit_model.train_step(ref, batch, labels, {0: 1, 1: 5}, 1e-3) for _ in range(4)]
    dpmod.world, dpmod.all_reduce_sum = real_world, real_ar
    p_ref = ref.get_params()
    print("losses dp ", losses)
    print("losses ref", ref_losses)
    err = np.abs(p_dp - p_ref).max()
    print("max |param_dp - param_single| = %.3e (max |param| %.3e)" % (err, np.abs(p_ref).max()))
    assert np.allclose(losses, ref_losses, rtol=1e-4), "loss mismatch"
    assert err < 5e-4, "parameter mismatch"
    print("DP_CHECK_OK world=%d" % world)
dist.barrier()
dist.destroy_process_group()
