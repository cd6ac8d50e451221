"""gnnqc/layout.py on CPU tensors: the reference's disjoint-graph batch (tiled COO + sample_indices,
libs/preprocessing_functions.py:638-666) <-> the offsets layout of include/gnnqc.h, without any host read-back."""
import torch

from gnnqc import layout, synth


def test_disjoint_round_trip_with_trailing_empty_windows():
    for make, T in ((lambda: synth.make_cml_batch(7, seed=2, ragged=True), 181), (lambda: synth.make_soilnet_batch(3, seed=2, n_nodes=11, ragged=True, min_nodes=4), 337)):
        b = make()
        idx, samp = layout.offsets_to_disjoint(b["edges"], b["node_off"], b["edge_off"], T)
        B = b["node_off"].numel() - 1
        # shuffle the tiled edge list: the conversion must not depend on its order
        perm = torch.randperm(idx.shape[0], generator=torch.Generator().manual_seed(0))
        for n_windows in (B, B + 5):            # the reference partitions into batch_size windows; short batches leave empty ones
            edges, node_off, edge_off, n_nodes = layout.disjoint_to_offsets(idx[perm], samp, n_windows, T, validate=True)
            assert n_nodes == int(b["node_off"][-1])
            assert torch.equal(node_off[: B + 1], b["node_off"]) and (node_off[B:] == node_off[B]).all()
            assert torch.equal(edge_off[: B + 1], b["edge_off"]) and (edge_off[B:] == edge_off[B]).all()
            for w in range(B):
                got = edges[edge_off[w]: edge_off[w + 1]]
                want = b["edges"][b["edge_off"][w]: b["edge_off"][w + 1]]
                assert sorted(map(tuple, got.tolist())) == sorted(map(tuple, want.tolist()))
            assert torch.equal(layout.node_offsets(samp, n_windows, T), node_off)


def test_link_shards_keep_every_target_neighbourhood():
    """shard_network_by_link: the ranks' target ranges partition the links, every target's neighbourhood and adjacency
    survive the renumbering, halo links carry their series but no neighbourhood."""
    import numpy as np

    from gnnqc.series import shard_network_by_link
    net = synth.make_cml_network(60, 64, seed=5, mean_neighbours=6)
    seen = []
    for world in (1, 3, 8):
        for rank in range(world):
            sh = shard_network_by_link(net, rank, world)
            g = sh["global_ids"]
            assert np.array_equal(sh["raw"], net["raw"][g], equal_nan=True) and np.array_equal(sh["rmed"], net["rmed"][g])
            for s_glob, s_loc in zip(sh["owned_global"], sh["owned_local"]):
                assert g[s_loc] == s_glob
                a, b = net["nbr_off"][s_glob], net["nbr_off"][s_glob + 1]
                la, lb = sh["nbr_off"][s_loc], sh["nbr_off"][s_loc + 1]
                assert np.array_equal(g[sh["nbr_idx"][la:lb]], net["nbr_idx"][a:b])          # same sensors, same order
                ea, eb = net["edge_off"][s_glob], net["edge_off"][s_glob + 1]
                assert np.array_equal(sh["edges"][sh["edge_off"][s_loc]: sh["edge_off"][s_loc + 1]], net["edges"][ea:eb])
            halo = np.setdiff1d(np.arange(g.size), sh["owned_local"])
            assert (np.diff(sh["nbr_off"])[halo] == 0).all()
            if world == 8:
                seen.append(sh["owned_global"])
    assert np.array_equal(np.concatenate(seen), np.arange(60))


def test_constructors_refuse_hyper_parameters_the_kernels_do_not_implement():
    """ADVICE: dense.alpha, activation, pool_size, regularizers and dropout are hard-wired in the native code (and pinned to
    keras_metadata.pb); a config that differs must not silently build another model."""
    import pytest

    from gnnqc.config import default_model_config, default_preproc_config
    from libs.create_model import BaselineClassifier, GCNClassifier
    pc = default_preproc_config("cml")
    for section, key, value in (("dense", "alpha", 0.1), ("sequence_layer", "activation", "relu"), ("sequence_layer", "pool_size", 2),
                                ("graph_convolution", "dropout_rate", 0.2), ("graph_convolution", "regularizer", 1e-4),
                                ("sequence_layer", "regularizer", 1e-4), ("dense", "regularizer", 1e-4)):
        mc = default_model_config()
        mc[section][key] = value
        with pytest.raises(NotImplementedError, match=key):
            GCNClassifier(mc, pc, device=None)
    for key, value in (("alpha", 0.2), ("activation", "relu"), ("pool_size", 4)):
        mc = default_model_config()
        mc.baseline_model[key] = value
        with pytest.raises(NotImplementedError, match=key):
            BaselineClassifier(mc, pc, device=None)
    GCNClassifier(default_model_config(), pc, device=None)      # the shipped configuration itself builds (host-only handle)
