"""gnnqc_score_seq_noalloc (the entry point for callers that may not allocate on the scoring
path: fixed-budget real-time loops, processes whose device memory belongs to another pool).

Collected last on purpose (file name), so a `-x` run has already exercised everything else when
it gets here.  Compares the no-allocation variant with gnnqc_score_seq through the C ABI on two
calls in a row - the second with a different batch, so contents the first call left in the shared
built-in scratch array cannot leak into the result - and with the CPU oracle directly, and checks
that an oversized call is refused with a message that names the limit."""
import os

import numpy as np
import pytest
import torch

import gnnqc_oracle as O
from conftest import MODELS, oracle_named
from gnnqc import GnnqcError, NativeModel, synth

pytestmark = pytest.mark.gpu

ATOL = RTOL = 1e-4   # the tolerance of tests/test_gpu_scoring.py


def _close(got, want):
    got, want = got.cpu().numpy().astype(np.float64), want.cpu().numpy().astype(np.float64)
    assert (np.abs(got - want) <= ATOL + RTOL * np.abs(want)).all(), "max abs err %.3e" % np.abs(got - want).max()


@pytest.mark.parametrize("name,F,T,sizes", [("model_cml_baseline", 2, 181, (1500, 333)),
                                            ("model_soilnet_baseline", 3, 337, (700, 65))])
def test_noalloc_matches_score_seq_twice_in_a_row(name, F, T, sizes):
    m = NativeModel.load(os.path.join(MODELS, name), device=0)
    try:
        for call, n in enumerate(sizes):
            x = synth.make_seq_batch(n, F, T, seed=300 + call, device="cuda")
            want = m.score_seq(x)
            out = torch.full((n,), -1.0, device="cuda")
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
            m.score_seq_noalloc(x, out)
            torch.cuda.synchronize()
            if call > 0:   # the very first use makes the driver load the module that holds the array
                assert torch.cuda.mem_get_info()[0] >= free_before, "the no-allocation entry point allocated device memory"
            _close(out, want)
            _close(out, O.baseline_forward(x.cpu(), O.to_torch_weights(oracle_named(name), torch.float32)).reshape(-1))
    finally:
        m.close()


def test_noalloc_rejects_oversized_calls():
    m = NativeModel.load(os.path.join(MODELS, "model_cml_baseline"), device=0)
    try:
        x = torch.zeros((20000, 181, 2), device="cuda")
        out = torch.empty(20000, device="cuda")
        with pytest.raises(GnnqcError, match="at most"):
            m.score_seq_noalloc(x, out)
        m.score_seq_noalloc(x[:0], out[:0])   # an empty call is fine and does nothing
    finally:
        m.close()
