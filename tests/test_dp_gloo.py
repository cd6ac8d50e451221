"""Host-side data-parallel logic over a world_size-2 gloo group on CPU: shards partition the batch,
the two all-reduces of the training step reproduce the single-process sums, and the epoch logic
(learning-rate schedule, class weights) matches the reference's rules."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gnnqc import dp, synth
from gnnqc.config import default_model_config
from libs import fit_model


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world_size, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        batch = synth.make_cml_batch(11, seed=1, ragged=True)
        labels = torch.arange(11) % 2
        sub, lab = dp.shard_cml_batch(batch, labels, rank, world_size)
        # stand-in for the device step: "moments" and a "reduction vector" that are plain sums over
        # the shard, so the all-reduced result must equal the full-batch sums
        f = sub["features"].double()
        moments = torch.stack([f[:, 0].sum(), f[:, 1].sum(), (f[:, 0] ** 2).sum(), (f[:, 1] ** 2).sum(), (f[:, 0] * f[:, 1]).sum(),
                               torch.tensor(float(f.shape[0]), dtype=torch.float64)])
        dp.all_reduce_sum(moments)
        rvec = torch.stack([sub["anom"].sum(), lab.float().sum(), torch.tensor(float(sub["anom"].shape[0]))])
        dp.all_reduce_sum(rvec)
        n = dp.global_count(sub["anom"].shape[0])
        out[rank] = (moments.numpy(), rvec.numpy(), n, int(sub["node_off"][-1]), sub["edges"].shape[0], dp.world())
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_all_reduce_matches_single_process():
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    batch = synth.make_cml_batch(11, seed=1, ragged=True)
    labels = torch.arange(11) % 2
    f = batch["features"].double()
    want_m = np.array([f[:, 0].sum(), f[:, 1].sum(), (f[:, 0] ** 2).sum(), (f[:, 1] ** 2).sum(), (f[:, 0] * f[:, 1]).sum(), f.shape[0]])
    for rank in (0, 1):
        m, r, n, nodes, edges, w = out[rank]
        np.testing.assert_allclose(m, want_m, rtol=1e-12)
        np.testing.assert_allclose(r, [batch["anom"].sum().item(), labels.sum().item(), 11], rtol=1e-5)
        assert n == 11 and w == (rank, 2)
    assert out[0][3] + out[1][3] == int(batch["node_off"][-1])
    assert out[0][4] + out[1][4] == batch["edges"].shape[0]


def test_shards_partition_every_batch_size():
    for n in (1, 2, 7, 64, 129):
        for ws in (1, 2, 4, 8):
            r = [dp.shard_range(n, k, ws) for k in range(ws)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(h - l for l, h in r) - min(h - l for l, h in r) <= 1


def test_shard_contents_are_the_original_windows():
    batch = synth.make_cml_batch(9, seed=4, ragged=True)
    labels = torch.arange(9)
    T = 181
    seen = []
    for rank in range(4):
        sub, lab = dp.shard_cml_batch(batch, labels, rank, 4)
        lo, hi = dp.shard_range(9, rank, 4)
        assert torch.equal(lab, labels[lo:hi]) and torch.equal(sub["anom"], batch["anom"][lo:hi])
        n0 = int(batch["node_off"][lo])
        assert torch.equal(sub["features"], batch["features"][n0 * T: int(batch["node_off"][hi]) * T])
        assert int(sub["node_off"][0]) == 0 and int(sub["edge_off"][0]) == 0
        seen.append(sub["anom"].shape[0])
    assert sum(seen) == 9


def test_schedule_and_class_weights_follow_the_reference_rules():
    mc = default_model_config()
    lr = mc.learning_rate
    lrs = []
    for epoch in range(8):
        lr = fit_model.learning_rate_for_epoch(mc, epoch, lr)
        lrs.append(lr)
    np.testing.assert_allclose(lrs[:5], [1e-4] * 5)
    np.testing.assert_allclose(lrs[5:], [1e-4 * 0.95, 1e-4 * 0.95 ** 2, 1e-4 * 0.95 ** 3])
    assert fit_model.calculate_weights(mc, []) == {0: 1, 1: 5}
    mc.weight_classes.calculate = True
    w = fit_model.calculate_weights(mc, [(None, torch.tensor([0, 0, 0, 1])), (None, torch.tensor([0, 1, 0, 0]))])
    assert w == {0: 8 / 6, 1: 2 * 8 / 2}
    mc.weight_classes.use = False
    assert fit_model.calculate_weights(mc, []) is None


class _FakeNative:
    """Stands in for NativeModel on the CPU: 'gradients' are plain sums over the shard, so the data-parallel result
    is known, and the device entry points refuse empty shards exactly like libgnnqc does (GNNQC_ERR_ARG)."""
    device = None
    soilnet = False

    def __init__(self, has_gcn):
        self.has_gcn = has_gcn
        self.applied = None

    def train_vector_len(self):
        return 4

    def train_moments_cml(self, features):
        assert features.shape[0] > 0, "gnnqc_train_moments_cml rejects n_rows <= 0"
        return torch.tensor([features.sum().item(), 0, 0, 0, 0, float(features.shape[0])], dtype=torch.float64)

    def train_grads_cml(self, features, anom, edges, node_off, edge_off, labels, moments, cw, global_batch=None, n_nodes_total=None):
        assert anom.shape[0] > 0, "gnnqc_train_grads_cml rejects B <= 0"
        return torch.tensor([anom.sum().item(), float(anom.shape[0]), labels.sum().item(), anom.shape[0] / global_batch], dtype=torch.float32)

    def train_grads_seq(self, x, labels, cw, global_batch=None):
        assert x.shape[0] > 0, "gnnqc_train_grads_seq rejects B <= 0"
        return torch.tensor([x.sum().item(), float(x.shape[0]), labels.sum().item(), x.shape[0] / global_batch], dtype=torch.float32)

    def train_apply(self, rvec, moments, lr):
        self.applied = (rvec.clone(), None if moments is None else moments.clone())


def _small_batch_worker(rank, world_size, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        batch = synth.make_cml_batch(1, seed=3)            # one window for two ranks: rank 1's shard is empty
        gcn, base = _FakeNative(True), _FakeNative(False)
        loss_g = fit_model.train_step(gcn, batch, torch.tensor([1]), {0: 1, 1: 5}, 1e-4)
        loss_b = fit_model.train_step(base, batch["anom"], torch.tensor([1]), None, 1e-4)
        out[rank] = (gcn.applied[0].numpy(), gcn.applied[1].numpy(), base.applied[0].numpy(), loss_g, loss_b)
    finally:
        dist.destroy_process_group()


def test_a_batch_smaller_than_the_world_does_not_hang():
    port = _free_port()
    out = mp.Manager().dict()
    mp.spawn(_small_batch_worker, args=(2, port, out), nprocs=2, join=True)
    batch = synth.make_cml_batch(1, seed=3)
    for rank in (0, 1):                                     # both ranks applied the same, full-batch reduction
        g, m, b, loss_g, loss_b = out[rank]
        np.testing.assert_allclose(g, [batch["anom"].sum().item(), 1, 1, 1.0], rtol=1e-6)
        np.testing.assert_allclose(m[[0, 5]], [batch["features"].sum().item(), batch["features"].shape[0]], rtol=1e-6)
        np.testing.assert_allclose(b, [batch["anom"].sum().item(), 1, 1, 1.0], rtol=1e-6)
        assert loss_g == loss_b == 1.0


def test_epoch_iterator_yields_every_batch_once_and_matches_train_step():
    """libs/fit_model._Uploads (shard -> upload one batch ahead -> step) on the CPU stand-in: every batch comes out once,
    in order, and step_on_shard on what it yields applies what train_step applies."""
    batches = [(synth.make_cml_batch(n, seed=10 + n), torch.arange(n) % 2) for n in (3, 1, 2)]
    a, b = _FakeNative(True), _FakeNative(True)
    seen = []
    for (inputs, labels), (kind, sub, lab, gb) in zip(batches, fit_model._Uploads(a, batches)):
        assert kind == "cml" and gb == inputs["anom"].shape[0] and torch.equal(lab, labels.to(torch.int64))
        loss = fit_model.step_on_shard(a, kind, sub, lab, gb, {0: 1, 1: 5}, 1e-4)
        ref = fit_model.train_step(b, inputs, labels, {0: 1, 1: 5}, 1e-4)
        assert float(loss.item()) == ref and torch.equal(a.applied[0], b.applied[0]) and torch.equal(a.applied[1], b.applied[1])
        seen.append(gb)
    assert seen == [3, 1, 2]
    assert list(fit_model._Uploads(a, [])) == []
