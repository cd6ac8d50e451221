"""The C-ABI library loads without a GPU and exports every symbol include/gnnqc.h declares."""
import ctypes
import os
import re

from conftest import ROOT
from gnnqc import LIB_PATH


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "gnnqc.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gnnqc_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported():
    lib = ctypes.CDLL(LIB_PATH)
    names = declared_symbols()
    assert "gnnqc_model_load" in names and "gnnqc_score_cml" in names
    for n in names:
        assert hasattr(lib, n), "missing export: " + n


def test_abi_version():
    lib = ctypes.CDLL(LIB_PATH)
    text = open(os.path.join(ROOT, "include", "gnnqc.h")).read()
    want = int(re.search(r"#define GNNQC_ABI_VERSION (\d+)", text).group(1))
    assert lib.gnnqc_abi_version() == want
