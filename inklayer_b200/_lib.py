"""ctypes binding of libinksam.so (include/inksam.h, include/inksam_ops.h).

There is no fallback: if the CUDA library is missing or does not load, importing the product path raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int64, c_longlong, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("INKSAM_LIB", os.path.join(_HERE, "libinksam.so"))  # override: A/B builds of the library


class InksamConfig(ctypes.Structure):
    """struct inksam_config (include/inksam.h)."""

    _fields_ = [
        ("embed_dim", c_int), ("depth", c_int), ("num_heads", c_int), ("num_global", c_int),
        ("global_attn_indexes", c_int * 8), ("window_size", c_int), ("img_size", c_int), ("patch_size", c_int),
        ("out_chans", c_int), ("pixel_mean", c_float * 3), ("pixel_std", c_float * 3),
    ]


# symbol -> (restype, argtypes); every symbol declared in include/*.h
SIGNATURES = {
    "inksam_create": (c_int, [POINTER(InksamConfig), c_int, POINTER(c_void_p)]),
    "inksam_destroy": (None, [c_void_p]),
    "inksam_load_tensor": (c_int, [c_void_p, c_char_p, c_void_p, POINTER(c_int64), c_int, c_void_p]),
    "inksam_finalize_weights": (c_int, [c_void_p, c_void_p]),
    "inksam_encode": (c_int, [c_void_p, c_void_p, c_int, c_int, POINTER(c_int), c_int64, c_int64, c_int64, c_void_p, c_void_p]),
    "inksam_decode_boxes": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                    c_void_p, c_void_p, c_float, c_void_p]),
    "inksam_decode_batch": (c_int, [c_void_p, c_void_p, c_int, c_void_p, POINTER(c_int), POINTER(c_int), POINTER(c_int),
                                    c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p]),
    "inksam_decode_batch_packed": (c_int, [c_void_p, c_void_p, c_int, c_void_p, POINTER(c_int), POINTER(c_int), POINTER(c_int),
                                           c_void_p, c_void_p, c_void_p, c_float, c_void_p]),
    "inksam_workspace_bytes": (c_size_t, [c_void_p, c_int, c_int]),
    "inksam_reserve": (c_int, [c_void_p, c_int, c_int]),
    "inksam_workspace_generation": (c_longlong, [c_void_p]),
    "inksam_launch_count": (c_longlong, [c_void_p]),
    "inksam_profile_enable": (c_int, [c_void_p, c_int]),
    "inksam_profile_read": (c_int, [c_void_p, POINTER(ctypes.c_double), POINTER(ctypes.c_double), POINTER(c_longlong)]),
    "inksam_last_error": (c_char_p, []),
    "inksam_morph_kernel_size": (c_int, [c_int, c_int]),
    "inksam_clean_masks": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, POINTER(c_longlong), c_void_p]),
    "inksam_content_iou": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, POINTER(c_longlong), c_void_p]),
    "inksam_op_gemm": (c_int, [c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                               c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_float, c_void_p, c_int,
                               c_int, c_int, c_int, c_void_p]),
    "inksam_op_gemm_qkv_windowed": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "inksam_op_attention": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "inksam_op_attention_trace": (c_int, [c_void_p]),
    "inksam_op_layernorm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_void_p]),
    "inksam_op_fold_ln_weights": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "inksam_op_row_stats_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "inksam_op_gemm_stats_partials": (c_int, [c_int]),
    "inksam_op_gemm_fold": (c_int, [c_int, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_float, c_int, c_void_p]),
    "inksam_op_patchify": (c_int, [c_void_p, c_int, c_int, POINTER(c_int), c_longlong, c_longlong, c_longlong,
                                   POINTER(c_float), POINTER(c_float), c_void_p, c_void_p]),
    "inksam_op_fill_pad_rows": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "inksam_op_postprocess": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_float, c_void_p]),
    "inksam_op_postprocess_packed": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_float, c_void_p]),
    "inksam_op_decoder_t2i_scratch_bytes": (c_size_t, [c_int]),
    "inksam_op_decoder_t2i": (c_int, [c_void_p, c_void_p, c_int, c_longlong, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "inksam_op_decoder_i2t": (c_int, [c_void_p, c_int, c_longlong, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "inksam_op_gemm_f16": (c_int, [c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_int,
                                   c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_float, c_void_p, c_int, c_int, c_int, c_void_p]),
    "inksam_op_gemm_rowadd_f16": (c_int, [c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int,
                                          c_void_p, c_void_p, c_int, c_void_p]),
    "inksam_resize_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "inksam_resize_u8": (c_int, [c_void_p, c_int, c_int, c_longlong, c_void_p, c_int, c_int, c_longlong, c_void_p, c_void_p]),
    "inksam_resize_static_max_bytes": (c_size_t, []),
    "inksam_resize_u8_static": (c_int, [c_void_p, c_int, c_int, c_longlong, c_void_p, c_int, c_int, c_longlong, c_void_p]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Loads libinksam.so once and types every entry point.  Raises if the library is absent: there is no CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C inklayer_b200/csrc`). inklayer_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    """Non-zero status -> RuntimeError carrying inksam_last_error()."""
    if rc != 0:
        raise RuntimeError("libinksam: " + load().inksam_last_error().decode(errors="replace"))
