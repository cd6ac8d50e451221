"""Pins the oracle (and the constants the CUDA path hard-codes) to artefacts the REFERENCE itself
holds: the Keras-serialised layer configs in model_*/keras_metadata.pb and the original checkpoint
bytes.  tests/golden/make_reference_pins.py extracted them in the build container with the oracle's
independent reader; here they are compared with (a) oracle/gnnqc_oracle.py, (b) csrc/model_spec.h and
the library's own model description, (c) the product's checkpoint reader, (d) SURVEY.md section 8(c)'s
checkpoint-key table.  TensorFlow cannot run here, so this is as far as "pinned" goes: the semantics of
the Keras / spektral ops themselves remain a restatement (oracle header)."""
import hashlib
import json
import os
import re

import numpy as np
import pytest

import ckpt_reader as R
import gnnqc_oracle as O
from conftest import GOLDEN, MODELS, PKG, REFERENCE
from gnnqc import NativeModel

NAMES = ["model_cml", "model_cml_baseline", "model_soilnet", "model_soilnet_baseline"]
CFG = json.load(open(os.path.join(GOLDEN, "keras_layer_configs.json")))
KEYS = json.load(open(os.path.join(GOLDEN, "checkpoint_key_map.json")))


# ---- (c) two readers, one set of bytes -----------------------------------------------------------------
@pytest.mark.parametrize("name", NAMES)
def test_independent_reader_equals_native_reader(name):
    named, info = R.load_named(os.path.join(MODELS, name))
    m = NativeModel.load(os.path.join(MODELS, name), device=None)
    native = m.named_params()
    assert set(named) == set(native)
    for k in named:
        assert named[k].shape == native[k].shape and np.array_equal(named[k], native[k]), k
    assert list(info["model_info"]) == list(m.model_info)
    assert info["normalization"] == m.normalization
    if m.has_gcn:
        assert info["ds_type"] == m.ds_type


@pytest.mark.parametrize("name", NAMES)
def test_fixture_checkpoints_hold_the_reference_bytes(name):
    """sha256 per tensor of the re-emitted fixture == sha256 taken from the original checkpoint."""
    named, _ = R.load_named(os.path.join(MODELS, name))
    want = {row["path"]: row for row in KEYS[name].values()}
    assert set(named) == set(want)
    for path, arr in named.items():
        assert list(arr.shape) == want[path]["shape"]
        assert hashlib.sha256(arr.tobytes()).hexdigest() == want[path]["sha256"], path


@pytest.mark.skipif(not os.path.isdir(REFERENCE), reason="reference not mounted")
@pytest.mark.parametrize("name", NAMES)
def test_committed_pins_match_the_mounted_reference(name):
    rows = {k.split("/")[1]: (p, f, s) for k, p, f, s in R.key_map(os.path.join(REFERENCE, name))}
    assert {n: (r["path"], r["full_name"], r["shape"]) for n, r in KEYS[name].items()} == rows
    named, _ = R.load_named(os.path.join(REFERENCE, name))
    native = NativeModel.load(os.path.join(REFERENCE, name), device=None).named_params()
    assert all(np.array_equal(named[k], native[k]) for k in named) and set(named) == set(native)
    raw = R.keras_layer_configs(os.path.join(REFERENCE, name, "keras_metadata.pb"))
    for path, entry in CFG[name].items():
        assert raw[path]["class_name"] == entry["class_name"]


# ---- (d) SURVEY 8(c) weight map --------------------------------------------------------------------------
def _survey_table(gcn):
    """checkpoint variable index -> (attribute path, Keras name with the layer counters rebased to 0)."""
    t = {}
    if gcn:
        t.update({0: ("gcn_layer/kernel", "general_conv/kernel"), 1: ("gcn_layer/bias", "general_conv/bias"),
                  2: ("gcn_layer/activation/alpha", "general_conv/p_re_lu/alpha"),
                  3: ("gcn_layer/batch_norm/gamma", "general_conv/batch_normalization/gamma"),
                  4: ("gcn_layer/batch_norm/beta", "general_conv/batch_normalization/beta"),
                  5: ("gcn_layer/batch_norm/moving_mean", "general_conv/batch_normalization/moving_mean"),
                  6: ("gcn_layer/batch_norm/moving_variance", "general_conv/batch_normalization/moving_variance")})
        pre, kpre, dense = "time_layer/", "time_layer/", ("dense", "dense2", "dense_out")
        first = {"time1": 19, "time2": 22, "time_layers/0": 7, "time_layers/1": 10, "time_layers/2": 13, "time_layers/3": 16, "time4": 25}
        dense0 = 28
    else:
        pre, kpre, dense = "", "", ("dense1", "dense2", "dense_out")
        first = {"time1": 12, "time2": 15, "time_layers/0": 0, "time_layers/1": 3, "time_layers/2": 6, "time_layers/3": 9, "time4": 18}
        dense0 = 21
    # Keras names layers in creation order (libs/create_model.py:61-79): time1, time2, time_layers[0..3], time4
    keras = dict(zip(O.LSTM_ORDER, ["lstm", "lstm_1", "lstm_2", "lstm_3", "lstm_4", "lstm_5", "lstm_6"]))
    for attr, n in first.items():
        for j, leaf in enumerate(("kernel", "recurrent_kernel", "bias")):
            t[n + j] = ("%s%s/cell/%s" % (pre, attr, leaf), "%s%s/%s" % (kpre, keras[attr], leaf))
    for i, (attr, kn) in enumerate(zip(dense, ("dense", "dense_1", "dense_2"))):
        t[dense0 + 2 * i] = (attr + "/kernel", kn + "/kernel")
        t[dense0 + 2 * i + 1] = (attr + "/bias", kn + "/bias")
    return t


def _rebased_names(rows):
    """Keras numbers layers per process (lstm_cell_100, dense_69, ...): rebase every counter to the
    smallest one seen in this model and drop the lstm_cell_N segment."""
    pat = re.compile(r"^(lstm|dense|general_conv|p_re_lu|batch_normalization)(?:_(\d+))?$")
    lo = {}
    for row in rows.values():
        for seg in row["full_name"].split("/")[1:]:
            m = pat.match(seg)
            if m:
                lo[m.group(1)] = min(lo.get(m.group(1), 1 << 30), int(m.group(2) or 0))
    out = {}
    for n, row in rows.items():
        segs = []
        for seg in row["full_name"].split("/")[1:]:
            m = pat.match(seg)
            if m:
                k = int(m.group(2) or 0) - lo[m.group(1)]
                seg = m.group(1) + ("_%d" % k if k else "")
            if not seg.startswith("lstm_cell"):
                segs.append(seg)
        out[int(n)] = "/".join(segs)
    return out


@pytest.mark.parametrize("name", NAMES)
def test_checkpoint_key_to_layer_map_is_the_surveyed_one(name):
    table = _survey_table("baseline" not in name)
    rows = KEYS[name]
    assert sorted(int(k) for k in rows) == sorted(table)
    names = _rebased_names(rows)
    for n, (attr, keras_name) in table.items():
        assert rows[str(n)]["path"] == attr, (n, rows[str(n)])
        assert names[n] == keras_name, (n, rows[str(n)]["full_name"], keras_name)


# ---- (a)/(b) layer hyper-parameters ------------------------------------------------------------------------
def _shape(entry):
    s = entry["build_input_shape"]
    return s["items"] if isinstance(s, dict) else s


@pytest.mark.parametrize("name", NAMES)
def test_layer_constants_and_order_match_keras_metadata(name):
    cfg = CFG[name]
    gcn = "baseline" not in name
    tl = "root.time_layer." if gcn else "root."
    soil = "soilnet" in name
    T, F = (337, 3) if soil else (181, 2)
    order = ["time1", "time2", "time_layers.0", "time_layers.1", "time_layers.2", "time_layers.3", "time4"]
    assert [o.replace(".", "/") for o in order] == O.LSTM_ORDER
    units = [16, 16, 32, 32, 64, 64, 128]
    pools = {1: "max_pooling", 3: "pooling_layers.0", 5: "pooling_layers.1"}
    t, c = T, ((16 + F if soil else 16 + 2) if gcn else F)
    m = NativeModel.load(os.path.join(MODELS, name), device=None)
    assert m.seq_len == T and m.n_features == F
    assert int(m.model_info[0] + m.model_info[1]) // (int(m.model_info[3]) if soil else 1) + 1 == T  # create_model.py:12,155
    for i, (lname, u) in enumerate(zip(order, units)):
        e = cfg[tl + lname]
        k = e["config"]
        assert e["class_name"] == "LSTM" and k["units"] == u == m.lstm_hid[i]
        # model_cml_baseline was traced with a dynamic time axis ([None, None, 2]); T = 181 then comes from model_info
        assert _shape(e) in ([None, t, c], [None, None, c]) and m.lstm_in[i] == c
        assert k["activation"] == "tanh" and k["recurrent_activation"] == "sigmoid"
        assert k["use_bias"] is True and k["unit_forget_bias"] is True
        assert k["go_backwards"] is False and k["stateful"] is False and k["time_major"] is False
        assert k["dropout"] == 0.0 and k["recurrent_dropout"] in (0, 0.0)
        assert k["return_sequences"] is (i < 6)
        cell = cfg[tl + lname + ".cell"]
        assert cell["class_name"] == "LSTMCell" and cell["config"]["units"] == u
        c = u
        if i in pools:
            p = cfg[tl + pools[i]]
            assert p["class_name"] == "MaxPooling1D" and _shape(p) in ([None, t, c], [None, None, c])
            assert p["config"]["pool_size"]["items"] == [O.POOL] and p["config"]["strides"]["items"] == [O.POOL]
            assert p["config"]["padding"] == "valid" and p["config"]["data_format"] == "channels_last"
            t = t // O.POOL
    d1 = "root.dense" if gcn else "root.dense1"
    for path, u, act, inp in ((d1, 64, "linear", 128), ("root.dense2", 64, "linear", 64), ("root.dense_out", 1, "sigmoid", 64)):
        e = cfg[path]
        assert e["class_name"] == "Dense" and e["config"]["units"] == u and e["config"]["activation"] == act
        assert e["config"]["use_bias"] is True and _shape(e) == [None, inp]
    for path in ("root.leakyrelu4", "root.leakyrelu5"):
        assert cfg[path]["class_name"] == "LeakyReLU"
        assert np.float32(cfg[path]["config"]["alpha"]) == np.float32(O.LEAKY)
    if gcn:
        e = cfg["root.time_layer"]
        assert e["config"]["layer_type"] == "lstm" and e["config"]["n_stacks"] == 2 and e["config"]["filter_1_size"] == 16
        assert e["config"]["pool_size"] == O.POOL and e["config"]["activation"] == "tanh"
        g = cfg["root.gcn_layer"]
        assert g["class_name"] == "GeneralConv"
        assert g["config"]["aggregate"] == "mean" and g["config"]["channels"] == 16 and g["config"]["use_bias"] is True
        assert g["config"]["activation"]["class_name"] == "PReLU" and g["config"]["activation"]["config"]["shared_axes"] is None
        assert _shape(g)[0]["items"] == [None, F]
        bn = cfg["root.gcn_layer.batch_norm"]["config"]
        assert bn["epsilon"] == O.BN_EPS and bn["momentum"] == O.BN_MOMENTUM and bn["axis"] == [1]
        assert bn["center"] is True and bn["scale"] is True
        assert cfg["root.gcn_layer.dropout"]["config"]["rate"] == 0
        assert _shape(cfg["root.gcn_layer.activation"]) == [None, 16]


def test_cuda_side_constants_match_keras_metadata():
    """csrc/model_spec.h and csrc/train.cu hard-code the same numbers; read them out of the source."""
    spec = open(os.path.join(PKG, "csrc", "model_spec.h")).read()
    leaky = float(re.search(r"leaky_alpha\s*=\s*([0-9.e+-]+)f", spec).group(1))
    eps = float(re.search(r"bn_eps\s*=\s*([0-9.e+-]+)f", spec).group(1))
    cfg = CFG["model_cml"]
    assert np.float32(leaky) == np.float32(cfg["root.leakyrelu4"]["config"]["alpha"])
    assert eps == cfg["root.gcn_layer.batch_norm"]["config"]["epsilon"]
    assert int(re.search(r"kGcnChannels\s*=\s*(\d+)", spec).group(1)) == cfg["root.gcn_layer"]["config"]["channels"]
    train = open(os.path.join(PKG, "csrc", "train.cu")).read()
    call = re.search(r"adam_kernel<<<.*?>>>\((.*?)\);", train, re.S).group(1)
    assert re.search(r"ts->stats,\s*%sf\b" % cfg["root.gcn_layer.batch_norm"]["config"]["momentum"], call)   # BN momentum argument
