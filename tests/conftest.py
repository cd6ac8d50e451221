import os
import sys

import pytest

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
PKG = os.path.join(ROOT, "gnn-xai-timeseries-qualitycontrol_b200")
for p in (PKG, os.path.join(ROOT, "oracle"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")
MODELS = os.path.join(GOLDEN, "models")
REFERENCE = "/root/reference"  # present in the build container only


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def models_dir():
    return MODELS


def oracle_named(name):
    """Shipped weights for the ORACLE side of a parity test: parsed by oracle/ckpt_reader.py, not by the
    library under test."""
    import gnnqc_oracle

    return gnnqc_oracle.load_named(os.path.join(MODELS, name))


def model_path(name):
    return os.path.join(MODELS, name)
