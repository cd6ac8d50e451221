"""TFRecord reader (gnnqc_tfrecord_read_cml) against files written by INDEPENDENT encoders: the
record framing by tensorboard's RecordWriter and the tf.train.SequenceExample payload by
google.protobuf's own serialiser (message types declared here from the public schema of
tensorflow/core/example/{example,feature}.proto).  Schema of a window: reference
libs/preprocessing_functions.py:220-283 (create_example)."""
import numpy as np
import pytest
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
from tensorboard.summary.writer.record_writer import RecordWriter

from gnnqc import GnnqcError
from gnnqc.tfrecord import read_cml_tfrecord, read_soilnet_tfrecord


def _example_classes():
    fd = descriptor_pb2.FileDescriptorProto(name="gnnqc_test_example.proto", package="gnnqc_test", syntax="proto3")

    def msg(name):
        m = fd.message_type.add()
        m.name = name
        return m

    def field(m, name, number, ftype, label=1, type_name=None, packed=None):
        f = m.field.add()
        f.name, f.number, f.type, f.label = name, number, ftype, label
        if type_name:
            f.type_name = ".gnnqc_test." + type_name
        if packed is not None:
            f.options.packed = packed
        return f

    T = descriptor_pb2.FieldDescriptorProto
    field(msg("BytesList"), "value", 1, T.TYPE_BYTES, 3)
    field(msg("FloatList"), "value", 1, T.TYPE_FLOAT, 3, packed=True)
    field(msg("Int64List"), "value", 1, T.TYPE_INT64, 3, packed=True)
    feat = msg("Feature")
    field(feat, "bytes_list", 1, T.TYPE_MESSAGE, type_name="BytesList")
    field(feat, "float_list", 2, T.TYPE_MESSAGE, type_name="FloatList")
    field(feat, "int64_list", 3, T.TYPE_MESSAGE, type_name="Int64List")
    # map<string, Feature> is a repeated nested entry {key = 1, value = 2}
    for owner, value_type in (("Features", "Feature"), ("FeatureLists", "FeatureList")):
        if owner == "FeatureLists":
            field(msg("FeatureList"), "feature", 1, T.TYPE_MESSAGE, 3, type_name="Feature")
        m = msg(owner)
        entry = m.nested_type.add()
        entry.name = "Entry"
        field(entry, "key", 1, T.TYPE_STRING)
        field(entry, "value", 2, T.TYPE_MESSAGE, type_name=value_type)
        field(m, "feature" if owner == "Features" else "feature_list", 1, T.TYPE_MESSAGE, 3, type_name=owner + ".Entry")
    se = msg("SequenceExample")
    field(se, "context", 1, T.TYPE_MESSAGE, type_name="Features")
    field(se, "feature_lists", 2, T.TYPE_MESSAGE, type_name="FeatureLists")
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName("gnnqc_test." + n))  # noqa: E731
    return get("SequenceExample"), get("Feature")


SequenceExample, Feature = _example_classes()


def _ctx(ex, key, floats=None, ints=None, strings=None):
    e = ex.context.feature.add()
    e.key = key
    if floats is not None:
        e.value.float_list.value.extend([float(v) for v in floats])
    if ints is not None:
        e.value.int64_list.value.extend([int(v) for v in ints])
    if strings is not None:
        e.value.bytes_list.value.extend([s.encode() for s in strings])


def _flist(ex, key, rows, kind):
    e = ex.feature_lists.feature_list.add()
    e.key = key
    for r in rows:
        f = e.value.feature.add()
        if kind == "float":
            f.float_list.value.extend([float(v) for v in r])
        else:
            f.int64_list.value.extend([int(v) for v in r])


def make_window(rng, n_nodes, T, target, flag):
    trsl = (60 + rng.normal(0, 1, size=(2, n_nodes, T))).astype(np.float32)
    med = (60 + rng.normal(0, 0.2, size=(2, n_nodes))).astype(np.float32)
    adj = rng.random((n_nodes, n_nodes)) < 0.5
    adj |= adj.T | np.eye(n_nodes, dtype=bool)
    nodes, nbrs = np.nonzero(adj)
    ids = ["cml_%02d" % i for i in range(n_nodes)]
    ex = SequenceExample()
    _ctx(ex, "anomaly_ID", strings=[ids[target]])
    _ctx(ex, "CML_ids", strings=ids)
    _ctx(ex, "dates", strings=["2019-01-01T00:%02d:00" % (t % 60) for t in range(T)])
    _ctx(ex, "anomaly_flag", ints=[flag])
    _ctx(ex, "node_numb", ints=[n_nodes])
    _ctx(ex, "link_numb", ints=[len(nodes)])
    for ch in (0, 1):
        p = "TRSL%d_" % (ch + 1)
        _ctx(ex, p + "rolling_median", floats=med[ch])
        _ctx(ex, p + "mean", floats=trsl[ch].mean(axis=1))
        _ctx(ex, p + "std", floats=trsl[ch].std(axis=1))
        _ctx(ex, p + "anomalous_cml", floats=trsl[ch][target])
        _flist(ex, "TRSL%d" % (ch + 1), trsl[ch].T, "float")          # T features of N floats
    _flist(ex, "nodes", [[v] for v in nodes], "int")
    _flist(ex, "neighbours", [[v] for v in nbrs], "int")
    return ex, trsl, med, np.stack([nodes, nbrs], axis=1)


def test_reader_against_independent_encoders(tmp_path):
    rng = np.random.default_rng(0)
    T = 181
    path = str(tmp_path / "cml_07_2019-01-01.tfrec")
    windows = []
    with open(path, "wb") as f:
        w = RecordWriter(f)
        for n_nodes, target, flag in ((5, 2, 0), (9, 0, 1), (3, 2, 0)):
            ex, trsl, med, edges = make_window(rng, n_nodes, T, target, flag)
            w.write(ex.SerializeToString())
            windows.append((n_nodes, target, flag, trsl, med, edges))
        w.flush()
    batch, labels, ids, dates = read_cml_tfrecord(path)
    assert labels.tolist() == [0, 1, 0] and ids == ["cml_02", "cml_00", "cml_02"]
    assert len(dates[0]) == T and dates[1][5] == "2019-01-01T00:05:00"
    assert batch["node_off"].tolist() == [0, 5, 14, 17] and batch["T"] == T
    no, eo = batch["node_off"].numpy(), batch["edge_off"].numpy()
    for i, (n, tgt, flag, trsl, med, edges) in enumerate(windows):
        want = np.stack([(trsl[c] - med[c][:, None]).T for c in (0, 1)], axis=-1)       # [T, N, 2]
        got = batch["features"].numpy()[no[i] * T: no[i + 1] * T].reshape(T, n, 2)
        np.testing.assert_array_equal(got, want)
        np.testing.assert_array_equal(batch["anom"].numpy()[i], want[:, tgt])
        np.testing.assert_array_equal(batch["edges"].numpy()[eo[i]: eo[i + 1]], edges)
    # another normalisation mode of parse_cml_tfrecord_fn
    b2, _, _, _ = read_cml_tfrecord(path, "standarization")
    n, tgt, flag, trsl, med, edges = windows[0]
    want = (trsl[0] - trsl[0].mean(axis=1, dtype=np.float32)[:, None]) / trsl[0].std(axis=1, dtype=np.float32)[:, None]
    np.testing.assert_allclose(b2["features"].numpy()[: T * n].reshape(T, n, 2)[:, :, 0], want.T, rtol=1e-5, atol=1e-5)


def test_corrupt_and_truncated_files_are_rejected(tmp_path):
    rng = np.random.default_rng(1)
    path = tmp_path / "a.tfrec"
    with open(path, "wb") as f:
        w = RecordWriter(f)
        w.write(make_window(rng, 4, 30, 1, 0)[0].SerializeToString())
        w.flush()
    raw = bytearray(path.read_bytes())
    bad = tmp_path / "bad.tfrec"
    flipped = bytearray(raw)
    flipped[200] ^= 1
    bad.write_bytes(bytes(flipped))
    with pytest.raises(GnnqcError, match="checksum"):
        read_cml_tfrecord(bad)
    bad.write_bytes(bytes(raw[:-7]))
    with pytest.raises(GnnqcError, match="truncated"):
        read_cml_tfrecord(bad)
    with pytest.raises(GnnqcError):
        read_cml_tfrecord(tmp_path / "missing.tfrec")


def make_soil_window(rng, n_nodes, T):
    """A SoilNet SequenceExample with the schema of create_example(ds_type='soilnet') (reference :284-341)."""
    raw = {"moisture": rng.uniform(5, 55, size=(n_nodes, T)), "temp": rng.uniform(-10, 35, size=(n_nodes, T)),
           "battv": rng.uniform(2900, 3500, size=(n_nodes, T))}
    raw = {k: v.astype(np.float32) for k, v in raw.items()}
    adj = rng.random((n_nodes, n_nodes)) < 0.4
    adj |= adj.T | np.eye(n_nodes, dtype=bool)
    nodes, nbrs = np.nonzero(adj)
    flags = (rng.random(n_nodes) < 0.3).astype(np.int64)
    sids = rng.permutation(1000)[:n_nodes]
    ex = SequenceExample()
    _ctx(ex, "node_numb", ints=[n_nodes])
    _ctx(ex, "link_numb", ints=[len(nodes)])
    _ctx(ex, "dates", strings=["2014-06-01T%02d:%02d:00" % ((15 * t // 60) % 24, (15 * t) % 60) for t in range(T)])
    for k, v in raw.items():
        _ctx(ex, k + "_mean", floats=v.mean(axis=1))
        _ctx(ex, k + "_std", floats=v.std(axis=1))
        _ctx(ex, k + "_rolling_median", floats=np.median(v, axis=1))
        _flist(ex, k, v.T, "float")                                   # T features of N floats
    _flist(ex, "nodes", [[v] for v in nodes], "int")
    _flist(ex, "neighbours", [[v] for v in nbrs], "int")
    _flist(ex, "anomaly_flag", [[v] for v in flags], "int")
    _flist(ex, "sensor_ids", [[v] for v in sids], "int")
    return ex, raw, np.stack([nodes, nbrs], axis=1), flags, sids


def test_soilnet_reader_against_independent_encoders(tmp_path):
    rng = np.random.default_rng(5)
    T = 49
    path = str(tmp_path / "2014-06-01.tfrec")
    windows = []
    with open(path, "wb") as f:
        w = RecordWriter(f)
        for n_nodes in (6, 11, 4):
            ex, raw, edges, flags, sids = make_soil_window(rng, n_nodes, T)
            w.write(ex.SerializeToString())
            windows.append((n_nodes, raw, edges, flags, sids))
        w.flush()
    batch, labels, sensor_ids, dates = read_soilnet_tfrecord(path)
    assert batch["node_off"].tolist() == [0, 6, 17, 21] and batch["T"] == T and len(dates) == 3 and len(dates[2]) == T
    no, eo = batch["node_off"].numpy(), batch["edge_off"].numpy()
    f32 = np.float32
    for i, (n, raw, edges, flags, sids) in enumerate(windows):
        want = np.stack([raw["moisture"].T / f32(60), (raw["temp"].T - f32(-20)) / f32(60), (raw["battv"].T - f32(2800)) / f32(800)], axis=-1)
        got = batch["features"].numpy()[no[i] * T: no[i + 1] * T].reshape(T, n, 3)
        np.testing.assert_array_equal(got, want)                      # scale_range in float32, like the reference's tf ops
        np.testing.assert_array_equal(batch["edges"].numpy()[eo[i]: eo[i + 1]], edges)
        np.testing.assert_array_equal(labels.numpy()[no[i]: no[i + 1]], flags)
        np.testing.assert_array_equal(sensor_ids.numpy()[no[i]: no[i + 1]], sids)
    b2, _, _, _ = read_soilnet_tfrecord(path, "standarization")
    n, raw, _, _, _ = windows[1]
    v = raw["temp"]
    want = (v - v.mean(axis=1, dtype=f32)[:, None]) / v.std(axis=1, dtype=f32)[:, None]
    np.testing.assert_allclose(b2["features"].numpy()[no[1] * T: no[2] * T].reshape(T, n, 3)[:, :, 1], want.T, rtol=1e-5, atol=1e-5)
    b3, _, _, _ = read_soilnet_tfrecord(path, "rolling_median")
    np.testing.assert_array_equal(b3["features"].numpy()[: T * 6].reshape(T, 6, 3)[:, :, 0],
                                  (windows[0][1]["moisture"] - np.median(windows[0][1]["moisture"], axis=1).astype(f32)[:, None]).T)
    with pytest.raises(GnnqcError):
        read_soilnet_tfrecord(path, "no_such_mode")
    with pytest.raises(GnnqcError, match="moisture"):
        cml = tmp_path / "cml.tfrec"
        with open(cml, "wb") as f:
            w = RecordWriter(f)
            w.write(make_window(rng, 4, 30, 1, 0)[0].SerializeToString())
            w.flush()
        read_soilnet_tfrecord(cml)
