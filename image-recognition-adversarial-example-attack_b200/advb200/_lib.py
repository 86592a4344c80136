"""ctypes binding of libadvb200.so (the C ABI declared in include/advb200.h).

There is no CPU fallback: if the shared library is missing or cannot be loaded the import of any
compute entry point raises, loudly.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int64, c_uint8, c_void_p
from pathlib import Path

_PKG_DIR = Path(__file__).resolve().parent.parent
LIB_PATH = Path(os.environ.get("ADVB200_LIB", _PKG_DIR / "lib" / "libadvb200.so"))

PREC_BF16 = 0
PREC_FP32 = 1
LOSS_CE_MEAN = 0
LOSS_CW_UNTGT = 1
LOSS_CW_TGT = 2
NUM_CONVS = 53

_lib = None

# name -> (restype, argtypes); must list EVERY symbol include/advb200.h declares
SIGNATURES = {
    "advb200_last_error": (c_char_p, []),
    "advb200_version": (c_int, []),
    "advb200_model_create": (c_int, [POINTER(c_void_p)] * 5 + [POINTER(c_float), c_void_p, c_void_p, c_int,
                                                                c_void_p, POINTER(c_void_p)]),
    "advb200_model_destroy": (c_int, [c_void_p]),
    "advb200_model_precision": (c_int, [c_void_p]),
    "advb200_workspace_bytes": (c_int64, [c_void_p, c_int]),
    "advb200_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int,
                                c_void_p, c_void_p, c_int64, c_void_p]),
    "advb200_input_grad": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float,
                                   c_float, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_int64, c_void_p]),
    "advb200_jpeg_scratch_bytes": (c_int64, [c_int, c_int, c_int]),
    "advb200_jpeg_roundtrip_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int64, c_void_p]),
    "advb200_jpeg_roundtrip_f32_noalloc": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "advb200_preprocess_scratch_bytes": (c_int64, [c_int, c_int, c_int, c_int]),
    "advb200_preprocess_u8": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int64, c_void_p]),
    "advb200_to_u8_hwc": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "advb200_vgg19_create": (c_int, [POINTER(c_void_p)] * 4 + [c_void_p, POINTER(c_void_p)]),
    "advb200_vgg19_destroy": (c_int, [c_void_p]),
    "advb200_vgg19_workspace_bytes": (c_int64, [c_void_p, c_int]),
    "advb200_vgg19_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "advb200_test_div9": (c_int, [c_void_p, c_void_p]),
    "advb200_backward": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int64, c_void_p]),
    "advb200_attack_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_float,
                                    c_float, c_float, c_void_p, c_void_p, c_int64, c_void_p]),
    "advb200_linf_step_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_float, c_int64, c_int, c_void_p]),
    "advb200_random_start_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "advb200_pgd_steps": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_float, c_float, c_float,
                                  c_int, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "advb200_cw_parts": (c_int, [c_int64]),
    "advb200_cw_init_f32": (c_int, [c_void_p] * 10 + [c_int, c_int64, c_void_p]),
    "advb200_cw_step_f32": (c_int, [c_void_p] * 13 + [c_int, c_int64, c_float, c_float, c_double, c_double,
                                                     c_double, c_void_p]),
    "advb200_cw_finalize_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p]),
    "advb200_smooth3x3_f32": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "advb200_median3x3_f32": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "advb200_quantize_f32": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "advb200_defend_input_f32": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p]),
    "advb200_eval_counters": (c_int, [c_void_p] * 5 + [c_float, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "advb200_test_conv": (c_int, [c_int] + [c_void_p] * 6 + [c_int] * 9 + [c_void_p]),
    "advb200_test_force_tile": (c_int, [c_int, c_int, c_int, c_int]),
    "advb200_linear_create": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, POINTER(c_void_p)]),
    "advb200_linear_destroy": (c_int, [c_void_p]),
    "advb200_linear_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p]),
    "advb200_fc_scratch_bytes": (c_int64, [c_int, c_int]),
    "advb200_fc_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "advb200_patch_im2col_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "advb200_vit_tokens_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "advb200_add_layernorm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_int64, c_int, c_int, c_int,
                                      c_void_p]),
    "advb200_gelu_bf16": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "advb200_window_attention_bf16": (c_int, [c_void_p, c_void_p, c_void_p] + [c_int] * 13 + [c_float, c_void_p]),
    "advb200_test_attention_simt": (c_int, [c_int]),
    "advb200_patch_merge_gather_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "advb200_ln_pool_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                    c_void_p]),
    "advb200_debug_trace": (c_int, [c_void_p, c_int64]),
    "advb200_debug_trace_index": (c_int64, [c_char_p, c_int64]),
    "advb200_launch_count": (c_int64, []),
    "advb200_profile_enable": (c_int, [c_int]),
    "advb200_profile_read": (c_int, [POINTER(ctypes.c_double), POINTER(c_int64)]),
    "advb200_profile_dump": (c_int64, [c_char_p, c_int64]),
    "advb200_tune_report": (c_int, [c_void_p, POINTER(c_int), POINTER(ctypes.c_double)]),
}


def load():
    """Load libadvb200.so once; raise RuntimeError when it is not there (no fallback path)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.is_file():
        raise RuntimeError(
            f"libadvb200.so not found at {LIB_PATH}: build it with ./build.sh (or __graft_entry__.build()). "
            "There is no CPU fallback for the attack hot path.")
    lib = ctypes.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc == 0:
        return
    msg = load().advb200_last_error()
    text = msg.decode("utf-8", "replace") if msg else "unknown error"
    if rc == -1:
        raise ValueError(f"{what}: {text}")
    raise RuntimeError(f"{what}: {text} (code {rc})")


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count() -> int:
    return int(load().advb200_launch_count())
