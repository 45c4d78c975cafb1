"""ctypes binding of the C-ABI library (include/b200attn.h).

There is deliberately no fallback: if `libb200attn.so` is missing or a symbol is
absent, importing a compute entry point raises.  Nothing in the product path ever
routes to a CPU or torch implementation.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

# B200ATTN_LIB: load another build of the same ABI (the development variant libb200attn_dbg.so); never a fallback
LIB_PATH = Path(os.environ.get("B200ATTN_LIB") or Path(__file__).resolve().parent / "libb200attn.so")

BF16, F16, FP8_E4M3, F32 = 0, 1, 2, 3
SCALE_DIV, SCALE_RCP = 0, 1
EXTEND_AUTO, EXTEND_TCGEN05, EXTEND_MMA = 0, 1, 2
WINDOW_NEW_TOKENS_ONLY = -2  # b200attn_extend_masked: no cached-prefix key at all
ROPE_FP32, ROPE_NATIVE = 0, 1
LAYOUT_LAYER_FIRST, LAYOUT_PAGE_FIRST, LAYOUT_LAYER_TABLE = 0, 1, 2

ERR_NAMES = {0: "OK", -1: "INVALID", -2: "UNSUPPORTED", -3: "CUDA", -4: "WORKSPACE"}

_vp, _i, _i64, _f, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t

# name -> (restype, argtypes); must list every symbol include/b200attn.h declares
SIGNATURES = {
    "b200attn_version": (_i, []),
    "b200attn_last_error": (C.c_char_p, []),
    "b200attn_device_info": (_i, [C.POINTER(_i)] * 4),
    "b200attn_kv_write": (_i, [_vp, _vp, _i64, _i64, _i64, _i64, _vp, _i, _i, _i, _vp, _vp, _i64, _i, _i,
                               _f, _f, _vp, _vp, _i, _i, _i, _vp]),
    "b200attn_rope_kv_write": (_i, [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp,
                                    _vp, _vp, _i64, _i, _f, _f, _vp, _vp, _i, _i, _i, _vp]),
    "b200attn_decode_plan_bytes": (_sz, [_i, _i]),
    "b200attn_decode_plan": (_i, [_vp, _i, _i, _i, _i, _vp, _sz, _vp]),
    "b200attn_decode_plan_window": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "b200attn_decode_workspace_bytes": (_sz, [_i, _i, _i]),
    "b200attn_decode": (_i, [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _i, _i, _i, _i,
                             _i, _i, _f, _f, _f, _f, _vp, _i, _vp, _sz, _vp, _vp]),
    "b200attn_decode_partial": (_i, [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _i, _i, _i, _i,
                                     _i, _i, _f, _f, _f, _f, _vp, _i, _vp, _sz, _vp, _vp]),
    "b200attn_decode_multiq": (_i, [_vp, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _i, _i,
                                    _i, _i, _i, _i, _i, _f, _f, _f, _f, _vp, _i, _vp, _sz, _vp, _i64, _vp]),
    "b200attn_decode_paged": (_i, [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _i, _i, _i, _i,
                                   _i, _i, _f, _f, _f, _f, _vp, _i, _vp, _sz, _vp, _i, _i, _vp]),
    "b200attn_decode_multiq_paged": (_i, [_vp, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _i,
                                          _i, _i, _i, _i, _i, _i, _f, _f, _f, _f, _vp, _i, _vp, _sz, _vp, _i64, _i,
                                          _vp]),
    "b200attn_decode_static_ws_capacity": (_sz, []),
    "b200attn_decode_static_ws": (_i, [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _i, _i, _i, _i,
                                       _i, _i, _f, _f, _f, _f, _vp, _i, _vp, _i, _vp]),
    "b200attn_decode_merge": (_i, [_vp, _i64, _i, _i, _i, _i, _vp, _i, _vp, _sz, _vp, _vp]),
    "b200attn_merge_state": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "b200attn_extend": (_i, [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _vp, _vp, _vp,
                             _i, _i, _i, _i, _i, _i, _i, _f, _f, _f, _f, _i, _i, _vp, _vp]),
    "b200attn_extend_masked": (_i, [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _vp, _vp, _vp,
                                    _i, _i, _i, _i, _i, _i, _i, _f, _f, _f, _f, _i, _i, _i, _vp, _vp, _i, _vp, _vp]),
    "b200attn_extend_paged": (_i, [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _vp, _vp, _vp,
                                   _i, _i, _i, _i, _i, _i, _i, _f, _f, _f, _f, _i, _i, _i, _vp, _vp, _i, _vp, _i, _vp]),
    "b200attn_extend_split_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "b200attn_extend_split": (_i, [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _i, _vp, _vp, _vp,
                                   _i, _i, _i, _i, _i, _i, _i, _f, _f, _f, _i, _vp, _i, _i, _i, _vp, _sz, _vp]),
    "b200attn_check_loc": (_i, [_vp, _i, _i64, _i64, _vp, _vp]),
    "b200attn_flatten_kv_indices": (_i, [_vp, _i64, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp]),
    "b200attn_move_kv": (_i, [_vp, _vp, _i, _i64, _vp, _vp, _i, _i, _vp]),
    "b200attn_gather_kv": (_i, [_vp, _i64, _vp, _i64, _i64, _vp, _vp]),
    "b200attn_transfer_kv": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i64, _i64, _vp, _vp, _i64, _i64, _i, _i, _i, _vp]),
}

_lib = None


class B200AttnError(RuntimeError):
    """Raised when a C-ABI call returns a negative code (reference style: TORCH_CHECK -> RuntimeError)."""


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m sglang_custom_b200.build` "
            "(there is no CPU or torch fallback for the attention path)")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is absent: fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().b200attn_last_error().decode("utf-8", "replace")
        raise B200AttnError(f"{what} failed [{ERR_NAMES.get(rc, rc)}]: {msg}")


def version() -> int:
    return load().b200attn_version()


def device_info() -> dict:
    vals = [_i(0) for _ in range(4)]
    check(load().b200attn_device_info(*[C.byref(v) for v in vals]), "b200attn_device_info")
    return dict(sm_count=vals[0].value, cc_major=vals[1].value, cc_minor=vals[2].value,
                max_smem_optin=vals[3].value)
