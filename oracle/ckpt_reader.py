"""CPU ORACLE - TEST INFRASTRUCTURE ONLY.  Independent pure-Python reader of the reference's Keras
SavedModel artefacts (no TensorFlow, no libgnnqc, no protobuf package).

Why it exists: the oracle must not receive its weights from the product's checkpoint reader
(csrc/tf_bundle.cc), otherwise a wrong checkpoint-key -> layer mapping would cancel out in every
parity test.  This file parses the same bytes a second time, written separately:

  read_bundle(prefix)         variables.index (LevelDB table) + variables.data-00000-of-00001
                              -> {checkpoint key: ndarray | bytes}
  object_graph(bundle)        _CHECKPOINTABLE_OBJECT_GRAPH (TrackableObjectGraph proto)
                              -> [(checkpoint_key, attribute path from the root, full_name)]
  load_named(dir)             {"gcn_layer/kernel": ndarray, "time_layer/time1/cell/kernel": ...}
                              keyed by object-graph attribute path (what gnnqc_oracle.py expects),
                              plus model_info / ds_type / normalization
  keras_layer_configs(path)   keras_metadata.pb (SavedMetadata proto) -> {node_path: metadata JSON}

Formats as found in /root/reference/model_*/ (SURVEY.md appendix): LevelDB table with 48-byte
footer and magic 0xdb4775248b80fb57, BundleEntryProto{dtype=1, shape=2, shard_id=3, offset=4,
size=5, crc32c=6}; reference load path: tf.keras.models.load_model (notebooks/pipeline.ipynb:202).
"""
import json
import os
import struct

import numpy as np

TABLE_MAGIC = 0xDB4775248B80FB57
DTYPES = {1: np.float32, 2: np.float64, 3: np.int32, 9: np.int64, 10: np.bool_}
DT_STRING = 7


# ---- protobuf wire format -------------------------------------------------------------------------
def _varint(buf, pos):
    out = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7


def wire_fields(buf):
    """[(field number, wire type, value)] of one message; value is int or bytes."""
    out, pos, end = [], 0, len(buf)
    while pos < end:
        tag, pos = _varint(buf, pos)
        num, wt = tag >> 3, tag & 7
        if wt == 0:
            val, pos = _varint(buf, pos)
        elif wt == 1:
            val, pos = buf[pos:pos + 8], pos + 8
        elif wt == 2:
            n, pos = _varint(buf, pos)
            val, pos = bytes(buf[pos:pos + n]), pos + n
        elif wt == 5:
            val, pos = buf[pos:pos + 4], pos + 4
        else:
            raise ValueError("unsupported wire type %d" % wt)
        out.append((num, wt, val))
    return out


# ---- crc32c (Castagnoli), masked as LevelDB / TensorFlow store it -----------------------------------
def _crc_table():
    tab = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
        tab.append(c)
    return np.array(tab, dtype=np.uint32)


_CRC = _crc_table()


def crc32c(data):
    c = 0xFFFFFFFF
    tab = _CRC
    for b in bytes(data):
        c = int(tab[(c ^ b) & 0xFF]) ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc(data):
    c = crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


# ---- LevelDB table --------------------------------------------------------------------------------
def _block(buf, offset, size):
    body = buf[offset:offset + size]
    ctype = buf[offset + size]
    if ctype != 0:
        raise ValueError("compressed table block (type %d)" % ctype)
    (stored,) = struct.unpack_from("<I", buf, offset + size + 1)
    if stored != masked_crc(buf[offset:offset + size + 1]):
        raise ValueError("table block checksum mismatch")
    (n_restart,) = struct.unpack_from("<I", body, len(body) - 4)
    end = len(body) - 4 - 4 * n_restart
    pos, key, out = 0, b"", []
    while pos < end:
        shared, pos = _varint(body, pos)
        non_shared, pos = _varint(body, pos)
        vlen, pos = _varint(body, pos)
        key = key[:shared] + body[pos:pos + non_shared]
        pos += non_shared
        out.append((key, body[pos:pos + vlen]))
        pos += vlen
    return out


def read_table(path):
    buf = open(path, "rb").read()
    (magic,) = struct.unpack_from("<Q", buf, len(buf) - 8)
    if magic != TABLE_MAGIC:
        raise ValueError("%s: not a table file" % path)
    foot = buf[-48:]
    _, p = _varint(foot, 0)
    _, p = _varint(foot, p)
    ioff, p = _varint(foot, p)
    isize, p = _varint(foot, p)
    entries = []
    for _, handle in _block(buf, ioff, isize):
        off, q = _varint(handle, 0)
        size, q = _varint(handle, q)
        entries.extend(_block(buf, off, size))
    return entries


# ---- tensor bundle --------------------------------------------------------------------------------
def read_bundle(prefix, verify=True):
    """prefix = '<dir>/variables/variables'.  Returns {key: ndarray (numeric) | [bytes] (string)}."""
    data = open(prefix + ".data-00000-of-00001", "rb").read()
    out = {}
    for key, val in read_table(prefix + ".index"):
        if key == b"":
            continue
        f = {n: v for n, _, v in wire_fields(val)}
        dtype = f.get(1, 0)
        shape = [dict((n, v) for n, _, v in wire_fields(d)).get(1, 0) for n, _, d in wire_fields(f.get(2, b"")) if n == 2]
        off, size = f.get(4, 0), f.get(5, 0)
        raw = data[off:off + size]
        if verify:
            (stored,) = struct.unpack("<I", f[6])
            if dtype != DT_STRING and stored != masked_crc(raw):
                raise ValueError("tensor checksum mismatch for %s" % key.decode())
        if dtype == DT_STRING:
            n = int(np.prod(shape)) if shape else 1
            lens, pos = [], 0
            for _ in range(n):
                ln, pos = _varint(raw, pos)
                lens.append(ln)
            pos += 4  # crc32c of the lengths
            strs = []
            for ln in lens:
                strs.append(raw[pos:pos + ln])
                pos += ln
            out[key.decode()] = strs
        else:
            out[key.decode()] = np.frombuffer(raw, dtype=DTYPES[dtype]).reshape(shape).copy()
    return out


def object_graph(bundle):
    """Walks the TrackableObjectGraph breadth-first from node 0 and returns
    [(checkpoint_key, 'path/from/root/<attribute name>', full_name)] for every serialized tensor
    (first path found per node, i.e. the shortest; slot variables are not followed)."""
    nodes = []
    for num, _, val in wire_fields(bundle["_CHECKPOINTABLE_OBJECT_GRAPH"][0]):
        if num != 1:
            continue
        children, attrs = [], []
        for n, _, v in wire_fields(val):
            d = {a: b for a, _, b in wire_fields(v)} if isinstance(v, bytes) else {}
            if n == 1:
                children.append((d.get(1, 0), d.get(2, b"").decode()))
            elif n == 2:
                attrs.append((d.get(1, b"").decode(), d.get(2, b"").decode(), d.get(3, b"").decode()))
        nodes.append((children, attrs))
    paths = {0: ""}
    order = [0]
    for nid in order:
        for child, name in nodes[nid][0]:
            if child not in paths:
                paths[child] = (paths[nid] + "/" + name) if paths[nid] else name
                order.append(child)
    out = []
    for nid in order:
        for name, full_name, key in nodes[nid][1]:
            out.append((key, paths[nid], full_name))
    return out


def key_map(savedmodel_dir):
    """[(checkpoint key, canonical attribute path, Keras full_name, shape)] of the float32 model
    variables (`variables/N/...`), sorted by N - the table in SURVEY.md section 8(c)."""
    prefix = os.path.join(savedmodel_dir, "variables", "variables")
    bundle = read_bundle(prefix)
    full = {k: f for k, _, f in object_graph(bundle)}
    rows = []
    for key, path in _canonical_paths(bundle).items():
        if key.startswith("variables/"):
            rows.append((int(key.split("/")[1]), key, path, full.get(key, ""), list(bundle[key].shape)))
    return [r[1:] for r in sorted(rows)]


# object-graph attribute names the oracle uses (Keras layer attribute paths; the variable itself is the
# leaf node, so its path already ends in kernel / bias / gamma / ...)
def load_named(savedmodel_dir):
    """Returns (named weights keyed by attribute path, info dict).  Only paths that do not go through
    'optimizer', 'keras_api', 'layer-N' / 'layer_with_weights-N' / 'variables' aliases are kept - the
    same canonical names libs/create_model.py gives its layers (gcn_layer, time_layer.time1, dense ...)."""
    prefix = os.path.join(savedmodel_dir, "variables", "variables")
    if not os.path.exists(prefix + ".index"):
        prefix = savedmodel_dir
    bundle = read_bundle(prefix)
    named = {}
    for key, path in _canonical_paths(bundle).items():
        val = bundle[key]
        if isinstance(val, np.ndarray) and val.dtype == np.float32:
            named[path] = val
    info = {}
    for k, v in bundle.items():
        head = k.split("/")[0]
        if head == "model_info":
            info["model_info"] = v.astype(np.int32)
        elif head in ("model_type", "model_normalization", "normalization"):
            info["ds_type" if head == "model_type" else "normalization"] = v[0].decode()
    return named, info


_SKIP_EDGES = ("optimizer", "keras_api", "variables", "trainable_variables", "non_trainable_variables",
               "regularization_losses", "layers", "signatures", "layer_regularization_losses", "layer_metrics", "metrics")


def _canonical_paths(bundle):
    """checkpoint key -> attribute path using only named-attribute edges (no list aliases)."""
    nodes = []
    for num, _, val in wire_fields(bundle["_CHECKPOINTABLE_OBJECT_GRAPH"][0]):
        if num != 1:
            continue
        children, attrs = [], []
        for n, _, v in wire_fields(val):
            if n == 1:
                d = {a: b for a, _, b in wire_fields(v)}
                children.append((d.get(1, 0), d.get(2, b"").decode()))
            elif n == 2:
                d = {a: b for a, _, b in wire_fields(v)}
                attrs.append(d.get(3, b"").decode())
        nodes.append((children, attrs))
    paths, order = {0: ""}, [0]
    for nid in order:
        for child, name in nodes[nid][0]:
            if name in _SKIP_EDGES or name.startswith(("layer-", "layer_with_weights-")):
                continue
            if child not in paths:
                paths[child] = (paths[nid] + "/" + name) if paths[nid] else name
                order.append(child)
    out = {}
    for nid in order:
        for key in nodes[nid][1]:
            if key and key not in out:
                out[key] = paths[nid]
    return out


# ---- keras_metadata.pb ------------------------------------------------------------------------------
def keras_layer_configs(path):
    """SavedMetadata{ repeated SavedObject nodes = 1 }, SavedObject{ node_id=2, node_path=3,
    identifier=4, metadata=5 (JSON) } -> {node_path: parsed metadata}."""
    out = {}
    for num, _, val in wire_fields(open(path, "rb").read()):
        if num != 1:
            continue
        d = {n: v for n, _, v in wire_fields(val)}
        meta = d.get(5, b"")
        if meta:
            out[d.get(3, b"").decode()] = json.loads(meta.decode())
    return out
