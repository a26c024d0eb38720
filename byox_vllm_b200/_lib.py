"""ctypes binding of libb200attn.so — the C-ABI declared in include/b200attn.h.

No fallback: if the shared library is missing or a symbol is absent, import of the compute path
raises. The prototypes below are the single Python-side statement of the ABI; tests check that
every symbol include/b200attn.h declares is exported by the built library.
"""
from __future__ import annotations

import ctypes
from ctypes import c_char_p, c_float, c_int32, c_int64, c_size_t, c_void_p
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "libb200attn.so"

ABI_VERSION = 10
DTYPE_BF16 = 0
DTYPE_F16 = 1


def dtype_code(dtype) -> int:
    """C-ABI dtype code of a torch dtype (bf16 -> 0, fp16 -> 1); anything else raises."""
    import torch

    if dtype == torch.bfloat16:
        return DTYPE_BF16
    if dtype == torch.float16:
        return DTYPE_F16
    raise RuntimeError(f"byox_vllm_b200: 16-bit dtypes only (bf16 / fp16), got {dtype}")

E_BADARG = -1
E_UNSUPPORTED = -2
E_WORKSPACE = -3
E_DRIVER = -4

_i32, _i64, _vp, _f32, _sz = c_int32, c_int64, c_void_p, c_float, c_size_t

# name -> (restype, argtypes); mirrors include/b200attn.h one to one.
PROTOTYPES = {
    "b200attn_abi_version": (_i32, []),
    "b200attn_last_error": (c_char_p, []),
    "b200attn_sm_count": (_i32, []),
    "b200attn_store_kvcache": (
        _i32, [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i32, _i32, _i64, _i32, _vp]),
    "b200attn_decode_pick_splits": (_i32, [_i32, _i32, _i32, _i32, _i32]),
    "b200attn_decode_partition_host": (_i32, [_vp, _i32, _i32, _i32, _i32, _i32, _vp, _i32]),
    "b200attn_prefill_schedule_host": (_i32, [_vp, _vp, _i32, _i32, _i32, _vp, _i32]),
    "b200attn_debug_decode_trace": (None, [_vp]),
    "b200attn_debug_prefill_trace": (None, [_vp]),
    "b200attn_set_decode_overlap": (_i32, [_i32]),
    "b200attn_decode_workspace_bytes": (_sz, [_i32, _i32, _i32, _i32, _i32]),
    "b200attn_decode": (
        _i32, [_vp, _vp, _vp, _vp, _i64, _i32, _vp, _vp, _vp, _sz,
               _i32, _i32, _i32, _i32, _i32, _i64, _f32, _i32, _i32, _vp]),
    "b200attn_decode_static_scratch": (
        _i32, [_vp, _vp, _vp, _vp, _i64, _i32, _vp, _vp,
               _i32, _i32, _i32, _i32, _i32, _i64, _f32, _i32, _i32, _vp]),
    "b200attn_store_decode": (
        _i32, [_vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _i64, _i32, _vp, _vp, _vp, _sz,
               _i32, _i32, _i32, _i32, _i32, _i64, _f32, _i32, _i32, _vp]),
    "b200attn_cascade_pick_chunks": (_i32, [_i32, _i32, _i32, _i32, _i32, _i32]),
    "b200attn_cascade_plan": (_i32, [_vp, _i64, _i32, _vp, _i32, _i32, _vp, _vp]),
    "b200attn_store_decode_cascade": (
        _i32, [_vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _i64, _i32, _vp, _vp, _vp, _sz,
               _i32, _i32, _i32, _i32, _i32, _i64, _f32, _i32, _vp, _i32, _i32, _vp]),
    "b200attn_prefill_varlen": (
        _i32, [_vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _i64, _i32, _vp, _vp,
               _i32, _i32, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _i32, _i64, _f32, _i32, _vp]),
    "b200attn_launch_count": (_i64, []),
    "b200attn_qk_norm_rope_store": (
        _i32, [_vp, _i64, _vp, _vp, _vp, _f32, _vp, _i32, _vp, _vp, _i64, _vp, _vp, _vp,
               _i32, _i32, _i32, _i32, _i64, _i32, _vp]),
    "b200attn_prepare_decode": (
        _i32, [_vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _vp]),
    "b200attn_set_split_counters": (_i32, [_vp, _i64]),
    "b200attn_prepare_prefill": (
        _i32, [_vp, _i32, _i32, _vp, _i32, _vp, _i32, _vp, _vp, _vp]),
    "b200attn_qkv_decode": (
        _i32, [_vp, _i64, _vp, _vp, _vp, _f32, _vp, _i32, _vp, _vp, _vp, _vp, _i64, _i32, _vp,
               _vp, _vp, _sz, _i32, _i32, _i32, _i32, _i32, _i64, _f32, _i32, _i32, _vp]),
}

_lib = None


class B200AttnError(RuntimeError):
    """Non-zero return of a libb200attn entry point (message = b200attn_last_error())."""

    def __init__(self, func: str, code: int, msg: str):
        super().__init__(f"{func} failed with code {code}: {msg}")
        self.func, self.code, self.msg = func, code, msg


def load() -> ctypes.CDLL:
    """Load libb200attn.so once and attach prototypes. Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is not built. Run `python -m byox_vllm_b200.build` (needs nvcc); "
            "there is no CPU or PyTorch fallback for the attention path.")
    lib = ctypes.CDLL(str(LIB_PATH))
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing: loud by design
        fn.restype = res
        fn.argtypes = args
    got = lib.b200attn_abi_version()
    if got != ABI_VERSION:
        raise ImportError(f"libb200attn ABI {got} != expected {ABI_VERSION}; rebuild the library")
    _lib = lib
    return lib


def check(func: str, code: int) -> None:
    if code != 0:
        msg = load().b200attn_last_error()
        raise B200AttnError(func, code, msg.decode("utf-8", "replace") if msg else "")
