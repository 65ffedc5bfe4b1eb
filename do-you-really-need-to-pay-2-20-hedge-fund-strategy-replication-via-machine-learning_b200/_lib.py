"""ctypes binding of libhfrep.so (the C ABI declared in include/hfrep.h).

There is no CPU fallback: importing this module raises if the shared library has not been
built (``python -c "import __graft_entry__ as g; g.build()"`` or ``sh csrc/build.sh``), and
every entry point raises ``HfrepError`` with the library's message on a non-zero return code.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhfrep.so")

MAX_E, MAX_L, MAX_F = 24, 24, 32
STATIC_SCRATCH_BYTES = 4 << 20     # HFREP_STATIC_SCRATCH_BYTES
N_STATS = 11                       # HFREP_N_STATS
SCHED_STRIDE = 8                   # HFREP_SCHED_STRIDE
WPT_CLUSTER = 60                   # HFREP_WPT_CLUSTER: warps_per_task value of the six-CTA cluster mapping
STAT_NAMES = ("sharpe", "tracking_error", "correlation", "omega0", "omega10", "cvar5", "ceq2", "ceq5", "ceq10",
              "skewness", "kurtosis")


class HfrepError(RuntimeError):
    pass


class FitCfg(C.Structure):
    """hfrep_fit_cfg (include/hfrep.h); defaults = literals of Autoencoder_encapsulate.py:79-96 + Nadam()."""
    _fields_ = [("E", C.c_int32), ("L_max", C.c_int32), ("T", C.c_int32), ("n_fit", C.c_int32),
                ("batch", C.c_int32), ("max_epochs", C.c_int32), ("patience", C.c_int32),
                ("early_stop", C.c_int32), ("alpha", C.c_float), ("lr", C.c_float), ("beta_1", C.c_float),
                ("beta_2", C.c_float), ("epsilon", C.c_float), ("schedule_decay", C.c_float),
                ("decay_base", C.c_float), ("n_x", C.c_int32), ("warps_per_task", C.c_int32),
                ("dump_step", C.c_int32)]


class EvalCfg(C.Structure):
    """hfrep_eval_cfg (include/hfrep.h)."""
    _fields_ = [("E", C.c_int32), ("F", C.c_int32), ("T_train", C.c_int32), ("T_test", C.c_int32),
                ("window", C.c_int32), ("beta_rolling", C.c_int32), ("want_oos", C.c_int32),
                ("alpha", C.c_float), ("cost_param", C.c_double), ("phi", C.c_double)]


EXPORTS = ("hfrep_last_error", "hfrep_version", "hfrep_window_index", "hfrep_window_gather", "hfrep_window_gather_f64",
           "hfrep_rolling_sigma", "hfrep_fit_cfg_default", "hfrep_nadam_schedule", "hfrep_fit_batched", "hfrep_fit_windows_workspace", "hfrep_fit_windows", "hfrep_fit_windows_static", "hfrep_fit_wide_workspace", "hfrep_fit_wide", "hfrep_predict_wide",
           "hfrep_eval_cfg_default", "hfrep_predict_metrics", "hfrep_series_metrics", "hfrep_ex_post_return", "hfrep_forward",
           "hfrep_ffma_peak", "hfrep_generator_weight_count", "hfrep_generator_forward", "hfrep_tc_selftest", "hfrep_tc_mma_rate",
           "hfrep_generator_tc_weight_bytes", "hfrep_generator_tc_prepare", "hfrep_generator_forward_tc")


def _load():
    if not os.path.isfile(LIB_PATH):
        raise HfrepError(
            f"{LIB_PATH} is missing: the CUDA library has not been built. Build it with "
            f"`sh {os.path.join(_HERE, 'csrc', 'build.sh')}` (needs nvcc; targets sm_100a). "
            "There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    lib.hfrep_last_error.restype = C.c_char_p
    lib.hfrep_last_error.argtypes = []
    lib.hfrep_version.restype = C.c_int
    lib.hfrep_window_index.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, C.c_int]
    lib.hfrep_window_gather.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
    lib.hfrep_window_gather_f64.argtypes = lib.hfrep_window_gather.argtypes
    lib.hfrep_rolling_sigma.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int, C.c_int, vp, vp]
    lib.hfrep_fit_cfg_default.argtypes = [C.POINTER(FitCfg)]
    lib.hfrep_fit_cfg_default.restype = None
    lib.hfrep_nadam_schedule.argtypes = [C.POINTER(FitCfg), C.c_int, vp]
    lib.hfrep_fit_batched.argtypes = [C.POINTER(FitCfg), vp, C.c_int, vp, vp, vp, C.c_int, vp, i64, vp, vp, vp, vp, vp,
                                      vp, vp, vp, vp]
    lib.hfrep_fit_windows_workspace.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.hfrep_fit_windows.argtypes = [C.POINTER(FitCfg), vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, vp, vp, vp,
                                      C.c_int, vp, i64, vp, vp, vp, C.c_size_t, vp, vp, vp, vp, vp, vp]
    lib.hfrep_fit_windows_static.argtypes = [C.POINTER(FitCfg), vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, vp, vp, vp,
                                             C.c_int, vp, i64, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.hfrep_fit_wide_workspace.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.hfrep_fit_wide.argtypes = [C.POINTER(FitCfg), vp, C.c_int, C.c_int, vp, vp, vp, i64, vp, vp, C.c_size_t, vp, vp, vp,
                                   vp, vp, vp]
    lib.hfrep_predict_wide.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, vp, vp,
                                       vp, vp, vp]
    lib.hfrep_eval_cfg_default.argtypes = [C.POINTER(EvalCfg)]
    lib.hfrep_eval_cfg_default.restype = None
    lib.hfrep_predict_metrics.argtypes = [C.POINTER(EvalCfg), vp, vp, vp, vp, vp, C.c_int, vp, vp, vp, C.c_int] + [vp] * 11
    lib.hfrep_series_metrics.argtypes = [vp, vp, vp, C.c_int, C.c_int, vp, vp]
    lib.hfrep_ex_post_return.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, vp, vp]
    lib.hfrep_forward.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_float, vp, vp]
    lib.hfrep_ffma_peak.argtypes = [C.c_int, vp, C.POINTER(C.c_double), vp]
    lib.hfrep_generator_weight_count.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.hfrep_generator_tc_weight_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.hfrep_generator_tc_weight_bytes.restype = C.c_size_t
    lib.hfrep_generator_tc_prepare.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp]
    lib.hfrep_generator_forward_tc.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, vp, C.c_int,
                                               C.c_int, vp, vp]
    lib.hfrep_tc_selftest.argtypes = [vp, vp, C.c_int, vp, C.c_int, vp]
    lib.hfrep_tc_mma_rate.argtypes = [C.c_int, C.c_int, vp]
    lib.hfrep_generator_forward.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, vp, C.c_int,
                                            C.c_int, vp, vp]
    for name in EXPORTS:
        fn = getattr(lib, name)
        if name not in ("hfrep_last_error", "hfrep_fit_cfg_default", "hfrep_eval_cfg_default", "hfrep_fit_windows_workspace",
                        "hfrep_fit_wide_workspace", "hfrep_generator_weight_count"):
            fn.restype = C.c_int
    lib.hfrep_fit_windows_workspace.restype = C.c_size_t
    lib.hfrep_fit_wide_workspace.restype = C.c_size_t
    lib.hfrep_generator_weight_count.restype = C.c_size_t
    return lib


lib = _load()


def check(rc: int) -> int:
    if rc < 0:
        raise HfrepError(f"libhfrep error {rc}: {lib.hfrep_last_error().decode()}")
    return rc


def fit_cfg(**kw) -> FitCfg:
    c = FitCfg()
    lib.hfrep_fit_cfg_default(C.byref(c))
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def eval_cfg(**kw) -> EvalCfg:
    c = EvalCfg()
    lib.hfrep_eval_cfg_default(C.byref(c))
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def ptr(t):
    """Device/host pointer of a torch tensor or numpy array (None -> NULL)."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return C.c_void_p(t.data_ptr())
    return C.c_void_p(t.ctypes.data)


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
