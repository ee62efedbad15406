"""ctypes binding of libb200clip.so (include/b200clip.h).  Importing this module never computes anything;
`load()` fails loudly when the CUDA library has not been built — there is no CPU fallback."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "libb200clip.so"

OK, EINVAL, ECUDA, ESTATE, EDEVICE, EKERNEL = 0, -1, -2, -3, -4, -5
EPI_F32, EPI_BIAS_BF16, EPI_BIAS_ACT_BF16, EPI_BIAS_RES_F32 = 0, 1, 2, 3


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("image_size", C.c_int32), ("patch", C.c_int32), ("width", C.c_int32),
                ("layers", C.c_int32), ("heads", C.c_int32), ("mlp_dim", C.c_int32), ("embed_dim", C.c_int32),
                ("activation", C.c_int32), ("max_batch", C.c_int32), ("ln_eps", C.c_float),
                ("mean", C.c_float * 3), ("std", C.c_float * 3)]


class TextConfig(C.Structure):
    _fields_ = [("context_length", C.c_int32), ("vocab_size", C.c_int32), ("width", C.c_int32), ("layers", C.c_int32),
                ("heads", C.c_int32), ("mlp_dim", C.c_int32), ("embed_dim", C.c_int32), ("activation", C.c_int32),
                ("max_batch", C.c_int32), ("ln_eps", C.c_float)]


class B200ClipError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libb200clip error {code}: {message}")
        self.code = code
        self.message = message


_P, _I32, _I64, _F = C.c_void_p, C.c_int32, C.c_int64, C.c_float
SIGNATURES = {
    "b200clip_create": (C.c_int, [C.POINTER(Config), C.POINTER(_P)]),
    "b200clip_destroy": (None, [_P]),
    "b200clip_last_error": (C.c_char_p, []),
    "b200clip_sync": (C.c_int, [_P, _P]),
    "b200clip_load_weight": (C.c_int, [_P, C.c_char_p, _P, C.POINTER(_I64), _I32]),
    "b200clip_finalize_weights": (C.c_int, [_P]),
    "b200clip_clone": (C.c_int, [_P, C.POINTER(_P)]),
    "b200clip_preprocess": (C.c_int, [_P, _P, _P, _P, _I32, _I32, _I32, _I32, _P, _P]),
    "b200clip_encode": (C.c_int, [_P, _P, _I32, _P, _P]),
    "b200clip_embed_u8": (C.c_int, [_P, _P, _P, _P, _I32, _I32, _P, _P]),
    "b200clip_embed_host": (C.c_int, [_P, _P, _I64, _P, _P, _I32, _I32, _P, _P]),
    "b200clip_text_init": (C.c_int, [_P, _P]),
    "b200clip_encode_text": (C.c_int, [_P, _P, _I32, _P, _P]),
    "b200clip_jpeg_probe": (C.c_int, [_P, C.c_size_t, _P, _P]),
    "b200clip_decode_jpeg": (C.c_int, [_P, _P, _P, _I32, _P, _I64, _P, _P, _P, _P, _I32, _P]),
    "b200clip_gallery_pack": (C.c_int, [_P, _P, _I64, _I32, _I32, _P, _P]),
    "b200clip_set_labels": (C.c_int, [_P, _I32, _P, _I32, _I32, _P]),
    "b200clip_classify": (C.c_int, [_P, _I32, _P, _I32, _F, _I32, _I32, _P, _P, _P]),
    "b200clip_topk": (C.c_int, [_P, _P, _I32, _P, _I64, _I32, _I32, _I64, _P, _P, _P]),
    "b200clip_topk_keys": (C.c_int, [_P, _P, _I32, _P, _I64, _I32, _I32, _P, _P]),
    "b200clip_merge_topk_keys": (C.c_int, [_P, _P, _P, _I32, _I32, _I32, _P, _P, _P]),
    "b200clip_search_sharded": (C.c_int, [_P, _P, _I32, _P, C.c_int64, _I32, _I32, C.c_int64, _P, _I32, _P, _P, _P]),
    "b200clip_topk_static": (C.c_int, [_P, _I32, _P, _I64, _I32, _I32, _I64, _P, _P, _P]),
    "b200clip_topk_static_status": (C.c_int, [_P]),
    "b200clip_merge_topk": (C.c_int, [_P, _P, _P, _I32, _I32, _I32, _P, _P, _P]),
    "b200clip_gemm_bf16": (C.c_int, [_P, _P, _P, _I32, _I32, _I32, _P, _P, _I32, _I32, _P]),
    "b200clip_attention": (C.c_int, [_P, _P, _P, _I32, _I32, _I32, _I32, _I32, _P]),
    "b200clip_layernorm": (C.c_int, [_P, _P, _P, _P, _P, _I32, _I32, _F, _P]),
    "b200clip_launch_count": (_I64, [_P]),
    "b200clip_profile": (C.c_int, [_P, _I32]),
    "b200clip_profile_read": (C.c_int, [_P, _P, _P, _P]),
}

_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m lumen_b200.build` "
                               "(nvcc, sm_100a). lumen_b200 has no CPU or PyTorch fallback.")
        lib = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise B200ClipError(rc, load().b200clip_last_error().decode("utf-8", "replace"))
