"""Checkpoint reader / writer: known answers computed from the reference's own SavedModel
files and round trips through this repo's writer."""
import os

import numpy as np
import pytest

from conftest import MODELS, REFERENCE
from gnnqc import GnnqcError, NativeModel

# name -> (float count, float64 sum over variables/*, model_info, normalization, ds_type)
KAT = {
    "model_cml": (188225, -248.704764227, [120, 60, 128, 1], "rolling_median", "cml"),
    "model_cml_baseline": (187073, -1431.707777736, [120, 60, 128], "rolling_median", "cml"),
    "model_soilnet": (188305, 240.831901614, [4320, 720, 128, 15], "scale_range", "soilnet"),
    "model_soilnet_baseline": (187137, 202.759443051, [4320, 720, 32, 15], "scale_range", "soilnet"),
}


def _check(m, name):
    n, total, info, norm, ds = KAT[name]
    p = m.get_params()
    assert p.size == n == m.n_params
    assert abs(p.astype(np.float64).sum() - total) < 1e-6
    assert list(m.model_info) == info
    assert m.normalization == norm and m.ds_type == ds
    assert m.has_gcn == ("baseline" not in name)
    assert m.seq_len == (181 if ds == "cml" else 337)
    assert m.lstm_hid == [16, 16, 32, 32, 64, 64, 128]


@pytest.mark.parametrize("name", sorted(KAT))
def test_stripped_fixture_known_answers(name):
    _check(NativeModel.load(os.path.join(MODELS, name), device=None), name)


@pytest.mark.skipif(not os.path.isdir(REFERENCE), reason="reference checkpoints not mounted")
@pytest.mark.parametrize("name", sorted(KAT))
def test_reference_checkpoint_known_answers(name):
    m = NativeModel.load(os.path.join(REFERENCE, name), device=None)
    _check(m, name)
    # the stripped fixture holds exactly the same numbers
    f = NativeModel.load(os.path.join(MODELS, name), device=None)
    assert np.array_equal(m.get_params(), f.get_params())


def test_spot_values_and_segment_names():
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=None)
    named = m.named_params()
    np.testing.assert_allclose(named["gcn_layer/kernel"][0, :3], [-0.018752, 0.058746, 0.467282], atol=1e-6)
    np.testing.assert_allclose(named["dense_out/bias"], [-0.049159], atol=1e-6)
    assert named["time_layer/time1/cell/kernel"].shape == (18, 64)
    assert named["time_layer/time4/cell/recurrent_kernel"].shape == (128, 512)
    # unit_forget_bias: the f-gate slice of a Keras LSTM bias starts at 1
    b = named["time_layer/time1/cell/bias"]
    assert 0.5 < b[16:32].mean() < 1.5
    segs = m.segments()
    assert sum(int(np.prod(s)) for _, _, s, _ in segs) == m.n_params
    assert sum(int(np.prod(s)) for _, _, s, tr in segs if tr) == m.n_trainable == 188193


def test_save_round_trip(tmp_path):
    for name in KAT:
        m = NativeModel.load(os.path.join(MODELS, name), device=None)
        m.save(tmp_path / name)
        back = NativeModel.load(tmp_path / name, device=None)
        assert np.array_equal(back.get_params(), m.get_params())
        assert list(back.model_info) == list(m.model_info)
        assert back.normalization == m.normalization and back.ds_type == m.ds_type


def test_create_set_get(tmp_path):
    rng = np.random.default_rng(0)
    ref = NativeModel.load(os.path.join(MODELS, "model_cml_baseline"), device=None)
    p = rng.standard_normal(ref.n_params).astype(np.float32)
    m = NativeModel.create(False, False, p, model_info=[120, 60, 128, 1], normalization="rolling_median", device=None)
    assert np.array_equal(m.get_params(), p)
    m.set_params(p * 2)
    assert np.array_equal(m.get_params(), p * 2)
    with pytest.raises(GnnqcError):
        m.set_params(p[:-1])


def test_corruption_is_detected(tmp_path):
    m = NativeModel.load(os.path.join(MODELS, "model_cml"), device=None)
    m.save(tmp_path / "c")
    data = tmp_path / "c" / "variables" / "variables.data-00000-of-00001"
    raw = bytearray(data.read_bytes())
    raw[5000] ^= 0x40
    data.write_bytes(bytes(raw))
    with pytest.raises(GnnqcError, match="checksum"):
        NativeModel.load(tmp_path / "c", device=None)
    with pytest.raises(GnnqcError):
        NativeModel.load(tmp_path / "does_not_exist", device=None)


def test_host_only_handle_refuses_device_calls():
    import ctypes as C

    from gnnqc._lib import ERR_STATE, lib
    m = NativeModel.load(os.path.join(MODELS, "model_cml_baseline"), device=None)
    rc = lib.gnnqc_score_seq(m._h, None, 0, 181, None, None)
    assert rc == ERR_STATE
    assert b"no CPU path" in lib.gnnqc_last_error()
