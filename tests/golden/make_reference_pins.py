"""Extracts what the reference itself holds about the hot path into small committed fixtures.  Run in
the build container, where the reference is mounted (it does not exist on the GPU box):

    python tests/golden/make_reference_pins.py

Outputs (both read by tests/test_reference_pins.py):
  keras_layer_configs.json   per model: every layer's class, config and build_input_shape exactly as
                             Keras serialised them into model_*/keras_metadata.pb (initializer /
                             constraint entries dropped, they do not affect inference) - this is the
                             reference-held statement of BN epsilon/momentum, LeakyReLU alpha, pooling
                             3/3/valid, the LSTM activations, GeneralConv aggregate/activation and the
                             layer order + tensor shapes between layers.
  checkpoint_key_map.json    per model: checkpoint key -> (attribute path, Keras full_name, shape,
                             sha256 of the tensor bytes) read from the ORIGINAL checkpoints with the
                             independent reader (oracle/ckpt_reader.py), so the re-emitted fixtures under
                             tests/golden/models and both readers can be checked against the reference's
                             own bytes on any box.
"""
import hashlib
import json
import os
import sys

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ckpt_reader as R  # noqa: E402

REF = os.environ.get("GNNQC_REFERENCE", "/root/reference")
OUT = os.path.join(ROOT, "tests", "golden")
MODELS = ["model_cml", "model_cml_baseline", "model_soilnet", "model_soilnet_baseline"]
DROP = ("_initializer", "_constraint", "shared_object_id")


def _trim(cfg):
    if isinstance(cfg, dict):
        return {k: _trim(v) for k, v in cfg.items() if not k.endswith(DROP)}
    if isinstance(cfg, list):
        return [_trim(v) for v in cfg]
    return cfg


def layer_configs():
    out = {}
    for name in MODELS:
        raw = R.keras_layer_configs(os.path.join(REF, name, "keras_metadata.pb"))
        out[name] = {
            path: {"class_name": m.get("class_name"), "config": _trim(m.get("config", {})),
                   "build_input_shape": _trim(m.get("build_input_shape"))}
            for path, m in raw.items() if ".keras_api" not in path
        }
    return out


def key_maps():
    out = {}
    for name in MODELS:
        bundle = R.read_bundle(os.path.join(REF, name, "variables", "variables"))
        rows = {}
        for key, path, full_name, shape in R.key_map(os.path.join(REF, name)):
            rows[key.split("/")[1]] = {"path": path, "full_name": full_name, "shape": shape,
                                      "sha256": hashlib.sha256(bundle[key].tobytes()).hexdigest()}
        out[name] = rows
    return out


if __name__ == "__main__":
    with open(os.path.join(OUT, "keras_layer_configs.json"), "w") as f:
        json.dump(layer_configs(), f, indent=1, sort_keys=True)
    with open(os.path.join(OUT, "checkpoint_key_map.json"), "w") as f:
        json.dump(key_maps(), f, indent=1, sort_keys=True)
    print("written")
