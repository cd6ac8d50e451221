"""Minimal stand-in for the OmegaConf nodes the reference passes around (attribute access on
nested YAML).  Users keep their own model_config.yml / preprocessing_config.yml files
(reference: libs/config/*.yml); `load_config` reads them with PyYAML."""
import copy

import yaml


class Config(dict):
    def __getattr__(self, k):
        try:
            v = self[k]
        except KeyError:
            raise AttributeError(k)
        return v

    def __setattr__(self, k, v):
        self[k] = v


def to_config(obj):
    if isinstance(obj, dict):
        return Config({k: to_config(v) for k, v in obj.items()})
    if isinstance(obj, list):
        return [to_config(v) for v in obj]
    return obj


def load_config(path):
    with open(path) as f:
        return to_config(yaml.safe_load(f))


_MODEL_DEFAULTS = {
    "train": False, "train_baseline": False, "epochs": 10, "model_path": "model_out", "optimizer": "adam",
    "es_patience": 10, "learning_rate": 1e-4, "calculate_threshold": True,
    "learning_learn_scheduler": {"use": True, "after_epochs": 5, "rate": 0.95},
    "sequence_layer": {"algorithm": "lstm", "filter_1_size": 16, "n_stacks": 2, "pool_size": 3, "alpha": 0.3,
                       "activation": "tanh", "regularizer": None, "kernel_size": None, "dropout": None},
    "graph_convolution": {"layer": "GeneralConv", "activation": "prelu", "units": 16, "aggregation_type": "mean",
                          "regularizer": None, "dropout_rate": 0},
    "dense": {"alpha": 0.3, "units": 64, "regularizer": None},
    "pooling": {"aggregation_type": "mean"},
    "weight_classes": {"use": True, "calculate": False, "class_0": 1, "class_1": 5},
    "baseline_model": {"type": "lstm", "model_path": "model_out_baseline", "n_stacks": 2, "filter_1_size": 16, "pool_size": 3,
                       "alpha": 0.3, "dense_layer_units": 64, "activation": "tanh", "regularizer": None},
}


def default_model_config():
    """The hyper-parameters every shipped checkpoint was built with."""
    return to_config(copy.deepcopy(_MODEL_DEFAULTS))


def default_preproc_config(ds_type="cml"):
    if ds_type == "cml":
        d = {"ds_type": "cml", "random_state": 44, "timestep_before": 120, "timestep_after": 60, "batch_size": 128,
             "normalization": "rolling_median"}
    else:
        d = {"ds_type": "soilnet", "random_state": 44, "timestep_before": 4320, "timestep_after": 720, "batch_size": 32,
             "normalization": "scale_range"}
    return to_config(d)
