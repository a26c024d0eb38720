"""byox_vllm_b200 — B200-native (sm_100a) attention hot path for nanovllm (lsynpy/byox_vllm).

Scope: the store_kvcache scatter, paged decode attention and varlen causal
prefill behind the unchanged `Attention.forward(q, k, v)` + `get_context()` boundary of
reference nanovllm/layers/attention.py. Everything else of nanovllm is used as-is.

  byox_vllm_b200.attention   drop-in Attention / store_kvcache (+ functional paged_decode,
                             prefill_varlen) over the C-ABI of include/b200attn.h
  byox_vllm_b200.context     mirror of the reference's Context contract (standalone use)
  byox_vllm_b200.install     puts the drop-in under an importable nanovllm
  byox_vllm_b200.build       nvcc build of libb200attn.so
"""
from .context import Context, get_context, reset_context, set_context  # noqa: F401

__all__ = ["Context", "get_context", "set_context", "reset_context", "Attention", "store_kvcache",
           "paged_decode", "prefill_varlen"]


def __getattr__(name):  # lazy: importing the package must not require the built library
    if name in ("Attention", "store_kvcache", "paged_decode", "prefill_varlen",
                "decode_num_splits"):
        from . import attention
        return getattr(attention, name)
    raise AttributeError(name)
