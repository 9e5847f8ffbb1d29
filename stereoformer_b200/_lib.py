"""ctypes binding of libsfb200.so (the C ABI declared in include/stereoformer_b200.h).

There is no CPU fallback: if the library is missing or does not load, importing an operator
raises.  Build it with ``python -m stereoformer_b200.build`` (or ``__graft_entry__.build()``).
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# tools/ that use the tuning hooks (sweeps, timelines) set SF_TUNING_LIB=1 to load the -DSF_TUNING build instead
LIB_PATH = os.path.join(HERE, "libsfb200_tuning.so" if os.environ.get("SF_TUNING_LIB") == "1" else "libsfb200.so")

SF_ALGO_AUTO, SF_ALGO_TMA, SF_ALGO_GENERIC, SF_ALGO_TC, SF_ALGO_EXACT = 0, 1, 2, 3, 4
SF_WIN_FIXED, SF_WIN_VARIANCE, SF_WIN_OFFSETS, SF_WIN_ABSOLUTE, SF_WIN_HEAD, SF_WIN_VARIANCE_BOUNDS = 0, 1, 2, 3, 4, 5
SF_DIV_TRUE, SF_DIV_RECIP = 0, 1

_c_int, _c_float, _p = ctypes.c_int, ctypes.c_float, ctypes.c_void_p

# name -> argtypes; mirrors include/stereoformer_b200.h (tests/test_abi.py checks both against the .so)
SIGNATURES = {
    "sf_corr_volume_f32": [_p, _p, _p] + [_c_int] * 7 + [_p],
    "sf_shift_volume_f32": [_p, _p, _p] + [_c_int] * 6 + [_p],
    "sf_soft_argmin_f32": [_p, _p, _p] + [_c_int] * 5 + [_p],
    "sf_local_window_f32": [_p] * 10 + [_c_int] * 5 + [_c_float, _c_float] + [_c_int] * 3 + [_p],
    "sf_local_window_gather_f32": [_p] * 5 + [_c_int] * 5 + [_c_float, _c_float] + [_c_int] * 3 + [_p, _c_int, ctypes.c_longlong, _p],
    "sf_local_window_gather_signal_f32": [_p] * 5 + [_c_int] * 5 + [_c_float, _c_float] + [_c_int] * 3 +
                                         [_p, _p, _p, _c_int, ctypes.c_longlong, ctypes.c_longlong, _p, _p],
    "sf_gather_wait_f32": [_p, _p, _c_int, _c_int, _p, ctypes.c_longlong, _p, ctypes.c_longlong, _p],
    "sf_sample_volume_f32": [_p] * 5 + [_c_int] * 5 + [_p],
    "sf_disp_warp_f32": [_p] * 6 + [_c_int] * 6 + [_p, _p],
    "sf_guidance_f32": [_p] * 6 + [_c_int] * 6 + [_p, _p],
    "sf_disp_warp_band_f32": [_p] * 6 + [_c_int] * 10 + [_p, _p],
    "sf_guidance_band_f32": [_p] * 6 + [_c_int] * 10 + [_p, _p],
    "sf_corr1d_pyramid_f32": [_p, _p, _p] + [_c_int] * 7 + [_p],
    "sf_corr1d_lookup_f32": [_p, _p, _c_int, _c_int, _p, _p, _p, _p] + [_c_int] * 4 + [_p],
    "sf_corr1d_alt_workspace_bytes": [_c_int] * 5,
    "sf_corr1d_alt_f32": [_p] * 6 + [ctypes.c_size_t] + [_c_int] * 8 + [_p],
    "sf_dcn_forward_f32": [_p] * 6 + [_c_int] * 15 + [_p],
    "sf_corr_volume_backward_f32": [_p] * 5 + [_c_int] * 6 + [_p],
    "sf_corr_volume_backward_ex_f32": [_p] * 5 + [_c_int] * 7 + [_p],
    "sf_soft_argmin_backward_f32": [_p] * 3 + [_c_int] * 5 + [_p],
    "sf_local_window_backward_f32": [_p] * 9 + [_c_int] * 5 + [_c_float] + [_c_int] * 3 + [_p],
    "sf_sample_volume_backward_f32": [_p] * 5 + [_c_int] * 5 + [_p],
    "sf_disp_warp_backward_f32": [_p] * 5 + [_c_int] * 6 + [_p],
    "sf_gru_reset_backward_f32": [_p, ctypes.c_longlong] + [_p] * 5 + [_c_int] * 5 + [_p],
    "sf_gru_blend_backward_f32": [_p, ctypes.c_longlong] + [_p] * 6 + [_c_int] * 5 + [_p],
    "sf_guidance_backward_f32": [_p] * 7 + [_c_int] * 6 + [_p],
    "sf_upsample_convex_backward_f32": [_p, _p, _c_float, _p, _p, _p] + [_c_int] * 4 + [_p],
    "sf_dcn_backward_f32": [_p] * 10 + [_c_int] * 15 + [_p],
    "sf_dcn_backward_workspace_bytes": [_c_int] * 15,
    "sf_dcn_backward_ws_f32": [_p] * 10 + [_c_int] * 15 + [_p, ctypes.c_size_t, _c_int, _p],
    "sf_dcn_forward_f64": [_p] * 6 + [_c_int] * 15 + [_p],
    "sf_dcn_backward_f64": [_p] * 10 + [_c_int] * 15 + [_p],
    "sf_dcn_workspace_bytes": [_c_int] * 5,
    "sf_dcn_workspace_bytes_ex": [_c_int] * 9,
    "sf_dcn_forward_ws_f32": [_p] * 6 + [_c_int] * 15 + [_p, ctypes.c_size_t, _c_int, _p],
    "sf_dcn_static_scratch_bytes": [],
    "sf_dcn_forward_static_f32": [_p] * 6 + [_c_int] * 15 + [_c_int, _p],
    "sf_epipolar_attn_bf16": [_p] * 5 + [_c_int] * 6 + [_p],
    "sf_upsample_convex_f32": [_p, _p, _p] + [_c_int] * 4 + [_p],
    "sf_upsample_convex_scaled_f32": [_p, _p, _c_float, _p] + [_c_int] * 4 + [_p],
    "sf_upsample_simple_f32": [_p, _p] + [_c_int] * 4 + [_p],
    "sf_upsample_simple_backward_f32": [_p, _p] + [_c_int] * 4 + [_p],
    "sf_upsample_convex_gather_f32": [_p, _p, _c_float, _p] + [_c_int] * 6 + [_p, _c_int, ctypes.c_longlong,
                                                                              ctypes.c_longlong, _p],
    "sf_gru_reset_f32": [_p, ctypes.c_longlong, _p, _p, _p] + [_c_int] * 5 + [_p],
    "sf_gru_blend_f32": [_p, ctypes.c_longlong, _p, _p, _p] + [_c_int] * 5 + [_p],
}

RESTYPES = {"sf_dcn_workspace_bytes": ctypes.c_size_t, "sf_dcn_backward_workspace_bytes": ctypes.c_size_t, "sf_dcn_workspace_bytes_ex": ctypes.c_size_t, "sf_dcn_static_scratch_bytes": ctypes.c_size_t, "sf_corr1d_alt_workspace_bytes": ctypes.c_size_t}   # everything else returns an int status code

_lib = None


class SFError(RuntimeError):
    pass


def lib():
    """Load (once) and return the ctypes handle; raises if the CUDA library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SFError("stereoformer_b200: %s not found -- build it with `python -m stereoformer_b200.build`; "
                          "there is no CPU fallback" % LIB_PATH)
        h = ctypes.CDLL(LIB_PATH)
        h.sf_last_error.restype = ctypes.c_char_p
        h.sf_last_error.argtypes = []
        h.sf_abi_version.restype = _c_int
        h.sf_device_sm_count.restype = _c_int
        for name, args in SIGNATURES.items():
            fn = getattr(h, name)
            fn.argtypes = args
            fn.restype = RESTYPES.get(name, _c_int)
        _lib = h
    return _lib


def call(name, *args):
    h = lib()
    rc = getattr(h, name)(*args)
    if rc != 0:
        msg = h.sf_last_error().decode("utf-8", "replace")
        if rc == 1:
            raise ValueError("%s: %s" % (name, msg))
        raise SFError("%s failed (code %d): %s" % (name, rc, msg))
