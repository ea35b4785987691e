"""ctypes binding of libstts_b200.so (include/stts_b200.h).  There is NO CPU fallback: a
missing library or a non-sm_100 device raises."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libstts_b200.so")

_vp, _i, _i64 = C.c_void_p, C.c_int, C.c_int64
_fp = C.c_void_p  # device / host float pointers are passed as raw addresses

SIGNATURES = {
    "stts_create": (_i, [C.POINTER(_vp), _i]),
    "stts_destroy": (None, [_vp]),
    "stts_last_error": (C.c_char_p, [_vp]),
    "stts_version": (C.c_char_p, []),
    "stts_load_tensor": (_i, [_vp, C.c_char_p, C.c_char_p, _fp, C.POINTER(_i64), _i]),
    "stts_finalize_weights": (_i, [_vp]),
    "stts_set_precision": (_i, [_vp, _i]),
    "stts_set_pdl": (_i, [_vp, _i]),
    "stts_set_small_batch_shapes": (_i, [_vp, _i]),
    "stts_set_layout": (_i, [_vp, _i, C.POINTER(_i), C.POINTER(_i)]),
    "stts_workspace_bytes": (_i64, [_vp]),
    "stts_set_token_layout": (_i, [_vp, _i, C.POINTER(_i)]),
    "stts_text_encoder": (_i, [_vp, C.c_char_p, _fp, _fp, _fp, _vp]),
    "stts_prosody_encoder": (_i, [_vp, C.c_char_p, _fp, _fp, _fp, _fp, _vp]),
    "stts_text_style_encoder": (_i, [_vp, C.c_char_p, _fp, _fp, _vp]),
    "stts_pe_cross": (_i, [_vp, _fp, _fp, _fp, _fp, _fp, _vp]),
    "stts_durations": (_i, [_vp, _fp, _i, _fp, _vp]),
    "stts_frame_index": (_i, [_vp, _fp, _fp, _i, _fp, _fp, _i, _fp, _vp]),
    "stts_frame_index_static": (_i, [_vp, _fp, _fp, _i, _fp, _i, _fp, _vp]),
    "stts_source_sinegen": (_i, [_vp, _fp, C.POINTER(_i), _i, _fp, _fp, C.POINTER(C.c_float), _i, C.c_float, C.c_float,
                                 C.c_float, C.c_float, _fp, _fp, _vp]),
    "stts_generator_block_snake": (_i, [_vp, C.c_char_p, _i, _i, C.POINTER(_i), _i, C.POINTER(_i), _fp, _fp, _fp, _vp]),
    "stts_stft_conv_transform": (_i, [_vp, C.c_char_p, _i, _i, _i, _i, _i, C.POINTER(_i), _fp, _fp, _fp, _fp, _vp]),
    "stts_stft_conv_inverse": (_i, [_vp, C.c_char_p, _i, _i, _i, _i, C.POINTER(_i), C.POINTER(_i), _fp, _fp, _fp, _fp,
                                    _vp]),
    "stts_pe_towers": (_i, [_vp, _fp, _fp, _fp, _fp, _vp]),
    "stts_decoder": (_i, [_vp, _fp, _fp, _fp, _fp, _fp, _vp]),
    "stts_prior_flow": (_i, [_vp, _fp, _fp, _fp, _fp, _fp, _fp, _vp]),
    "stts_source_pcph": (_i, [_vp, _fp, _fp, _fp, _fp, _vp]),
    "stts_generator": (_i, [_vp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _vp]),
    "stts_generator_from_stft": (_i, [_vp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _vp]),
    "stts_speech_frames": (_i, [_vp] + [_fp] * 13 + [_vp]),
    "stts_check_source_status": (_i, [_vp, _vp]),
    "stts_saturation_count": (_i64, [_vp, _vp]),
    "stts_pcm16": (_i, [_vp, _fp, _i64, _vp, _vp]),
    "stts_test_conv": (_i, [_vp, _i, _fp, _i, _i, _i, _i, _fp, _fp, _i, _fp, _i, _vp]),
    "stts_debug_read": (_i64, [_vp, C.c_char_p, _fp, _i64, C.POINTER(_i), C.POINTER(_i), _vp]),
    "stts_layout_info": (_i, [_vp, _i, C.POINTER(_i), C.POINTER(_i)]),
    "stts_launch_count": (_i64, [_vp]),
    "stts_set_profiling": (_i, [_vp, _i]),
    "stts_profile_json": (C.c_char_p, [_vp]),
}

_lib = None


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: run `python -m stylish_tts_b200.build` (or __graft_entry__.build()). "
            "There is no CPU / PyTorch fallback for the synthesis path.")
    import torch  # noqa: F401  (makes sure the CUDA runtime that torch ships is the one we bind to)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
