"""ctypes binding of libgnnqc.so (C ABI declared in include/gnnqc.h).

The shared object is the product: there is no Python or CPU fallback.  If it is missing this
module raises at import, and device entry points raise when no B200 is present.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GNNQC_LIB") or os.path.join(_HERE, "libgnnqc.so")   # GNNQC_LIB: an A/B build (build_native.py --variant=...)

OK, ERR_ARG, ERR_CKPT, ERR_CUDA, ERR_STATE = 0, -1, -2, -3, -5


class GnnqcError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("libgnnqc error %d: %s" % (code, message))
        self.code = code


class ModelDesc(C.Structure):
    _fields_ = [
        ("has_gcn", C.c_int32),
        ("soilnet", C.c_int32),
        ("n_features", C.c_int32),
        ("n_lstm", C.c_int32),
        ("lstm_in", C.c_int32 * 8),
        ("lstm_hid", C.c_int32 * 8),
        ("dense_units", C.c_int32),
        ("n_model_info", C.c_int32),
        ("model_info", C.c_int32 * 4),
        ("seq_len", C.c_int32),
        ("n_params", C.c_int64),
        ("n_trainable", C.c_int64),
        ("ds_type", C.c_char * 16),
        ("normalization", C.c_char * 32),
    ]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libgnnqc.so is not built (%s). Run `python __graft_entry__.py` or "
            "`python gnn-xai-timeseries-qualitycontrol_b200/build_native.py`; there is no fallback implementation." % LIB_PATH
        )
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, fp = C.c_void_p, C.c_int32, C.c_int64, C.POINTER(C.c_float)
    sigs = {
        "gnnqc_abi_version": (C.c_int, []),
        "gnnqc_last_error": (C.c_char_p, []),
        "gnnqc_model_load": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(vp)]),
        "gnnqc_model_create": (
            C.c_int,
            [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, fp, i64, C.POINTER(i32), C.c_int, C.c_char_p, C.c_int, C.POINTER(vp)],
        ),
        "gnnqc_model_describe": (C.c_int, [vp, C.POINTER(ModelDesc)]),
        "gnnqc_model_get_params": (C.c_int, [vp, fp, i64]),
        "gnnqc_model_set_params": (C.c_int, [vp, fp, i64]),
        "gnnqc_model_segment": (C.c_int, [vp, C.c_int, C.c_char * 128, C.POINTER(i64), i64 * 4, C.POINTER(i32), C.POINTER(i32)]),
        "gnnqc_model_save": (C.c_int, [vp, C.c_char_p]),
        "gnnqc_model_free": (None, [vp]),
        "gnnqc_score_seq": (C.c_int, [vp, vp, i64, i32, vp, vp]),
        "gnnqc_score_seq_noalloc": (C.c_int, [vp, vp, i64, i32, vp, vp]),
        "gnnqc_score_cml": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i32, i64, vp, vp]),
        "gnnqc_score_cml_host": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i32, i64, vp, vp]),
        "gnnqc_select_threshold": (C.c_int, [C.c_int, vp, vp, i64, vp, vp, vp, vp]),
        "gnnqc_roc_auc": (C.c_int, [C.c_int, vp, vp, i64, vp, vp]),
        "gnnqc_score_soilnet": (C.c_int, [vp, vp, vp, vp, vp, i64, i32, i64, i64, vp, vp]),
        "gnnqc_series_create": (C.c_int, [C.c_int, vp, vp, i64, i64, vp, vp, vp, vp, C.POINTER(vp)]),
        "gnnqc_series_update": (C.c_int, [vp, i64, i64, vp, vp]),
        "gnnqc_series_update_async": (C.c_int, [vp, i64, i64, vp, vp, vp]),
        "gnnqc_score_rows": (C.c_int, [vp, vp, vp, i64, i32, i64, vp, vp]),
        "gnnqc_series_free": (None, [vp]),
        "gnnqc_score_cml_series": (C.c_int, [vp, vp, vp, vp, i64, i32, i32, vp, vp, vp]),
        "gnnqc_soil_series_create": (C.c_int, [C.c_int, vp, vp, i64, i64, vp, C.POINTER(vp)]),
        "gnnqc_soil_series_free": (None, [vp]),
        "gnnqc_score_soilnet_series": (C.c_int, [vp, vp, vp, i64, i32, i32, vp, vp, vp, vp, vp]),
        "gnnqc_tfrecord_read_cml": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(vp)]),
        "gnnqc_cml_records_info": (C.c_int, [vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]),
        "gnnqc_cml_records_copy": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
        "gnnqc_cml_records_id": (C.c_char_p, [vp, i64]),
        "gnnqc_cml_records_dates": (C.c_char_p, [vp, i64]),
        "gnnqc_cml_records_free": (None, [vp]),
        "gnnqc_tfrecord_read_soilnet": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(vp)]),
        "gnnqc_soilnet_records_info": (C.c_int, [vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]),
        "gnnqc_soilnet_records_copy": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
        "gnnqc_soilnet_records_dates": (C.c_char_p, [vp, i64]),
        "gnnqc_soilnet_records_free": (None, [vp]),
        "gnnqc_ig_cml_series": (C.c_int, [vp, vp, vp, vp, i64, i32, i32, i32, vp, vp, vp, vp, vp, vp]),
        "gnnqc_ig_cml_raw": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i32, i64, i32, vp, vp, vp, vp]),
        "gnnqc_ig_cml": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i32, i64, i32, vp, vp, vp, vp]),
        "gnnqc_train_vector_len": (i64, [vp]),
        "gnnqc_train_moments_cml": (C.c_int, [vp, vp, i64, vp, vp]),
        "gnnqc_train_grads_cml": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, i64, i32, i64, vp, C.c_float, C.c_float, i64, vp, vp]),
        "gnnqc_train_moments_soilnet": (C.c_int, [vp, vp, i64, vp, vp]),
        "gnnqc_train_grads_soilnet": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i32, i64, i64, vp, C.c_float, C.c_float, i64, vp, vp]),
        "gnnqc_train_grads_seq": (C.c_int, [vp, vp, vp, i64, i32, C.c_float, C.c_float, i64, vp, vp]),
        "gnnqc_train_set_optimizer": (C.c_int, [vp, C.c_int]),
        "gnnqc_train_iterations": (i64, [vp]),
        "gnnqc_train_apply": (C.c_int, [vp, vp, vp, C.c_float, vp]),
        "gnnqc_score_scratch_bytes": (i64, [vp, i64, i32]),
        "gnnqc_launch_count": (i64, [vp]),
        "gnnqc_profile_enable": (C.c_int, [vp, C.c_int]),
        "gnnqc_profile_read": (C.c_int, [vp, C.c_int, vp, C.POINTER(C.c_double), C.POINTER(i64), C.POINTER(C.c_int)]),
        "gnnqc_selftest_umma": (C.c_int, [C.c_int, C.POINTER(C.c_float)]),
        "gnnqc_probe_fp32_peak": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_double)]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc):
    if rc != OK:
        raise GnnqcError(rc, lib.gnnqc_last_error().decode("utf-8", "replace"))
    return rc
