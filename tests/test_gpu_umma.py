"""tcgen05 / TMEM building blocks (csrc/umma.cuh): a 128x64x48 bf16 GEMM through the tensor cores
must reproduce a host float loop (inputs are exact in bf16, accumulation is fp32)."""
import ctypes as C

import pytest

from gnnqc._lib import lib

pytestmark = pytest.mark.gpu


def test_umma_selftest_gemm():
    err = C.c_float(-1.0)
    rc = lib.gnnqc_selftest_umma(0, C.byref(err))
    assert rc == 0
    assert 0.0 <= err.value < 1e-4, err.value
