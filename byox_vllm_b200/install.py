"""Put the B200 attention path under nanovllm without touching nanovllm's sources.

The reference binds the op by name: `from nanovllm.layers.attention import Attention`
(models/qwen3.py:7) and reads metadata through `nanovllm.utils.context.get_context`
(layers/attention.py:7,68). install() therefore
  1. registers byox_vllm_b200.attention as `nanovllm.layers.attention` in sys.modules BEFORE any
     nanovllm module is imported (importing one runs nanovllm/__init__.py and with it
     models/qwen3.py), so the reference's attention.py - and triton / flash_attn - are never
     imported (tests/test_host_logic.py::test_install_never_imports_flash_attn_or_triton);
  2. imports nanovllm.utils.context (pure Python, no CUDA deps) and points our module's
     `get_context` at it, so ModelRunner.set_context(...) feeds our forward();
  3. if nanovllm.models.qwen3 was already imported, rebinds its `Attention` name and swaps the
     class of existing instances' constructor target.
  4. (`fuse_qkv`, default on; B200ATTN_FUSE_QKV=0 or fuse_qkv=False keeps the
     frozen Attention.forward boundary) rebinds Qwen3Attention.forward (models/qwen3.py:70-86) to
     qkv_attention.qwen3_attention_forward: q/k RMSNorm + RoPE + KV scatter (+ the paged decode)
     become one library call on the raw qkv_proj output.
  5. (`fast_meta`, default on; B200ATTN_FAST_META=0 or fast_meta=False disables)
     rebinds ModelRunner._prepare_decode / _run_model (engine/model_runner.py:206-254) to
     fast_meta.prepare_decode / run_model: the decode metadata is expanded on the device into the
     CUDA-graph input buffers by one copy + one launch.
  6. (RPC half, `fast_rpc`, default on; B200ATTN_FAST_RPC=0 or fast_rpc=False
     disables; the transport runs only at tensor_parallel_size > 1, the `run` / `__init__`
     replacements at every size; EVERY rank must bind it - rank 0 checks and raises) rebinds ModelRunner._write_shm / _read_shm /
     run and the shared-memory block size (engine/model_runner.py:44,57-65,85-100) to
     fast_rpc: rank 0 ships a compact int32 record per step instead of pickled Sequence objects,
     in a block sized from the engine's limits; `run` loses the debug f-string that dereferences
     `logits` on ranks > 0 (the unmodified reference cannot take a step at TP > 1).
  7. switches the decode launches to overlap mode (attention.set_decode_overlap; B200ATTN_DECODE_
     OVERLAP=0 keeps it off): each decode launch is a programmatic dependent of the kernel before
     it and runs its prologue while that kernel drains. Valid here because the engine writes
     context_lens / block_tables once per step (model_runner.py:206-226), never right before an
     attention call.
For tensor parallel runs nanovllm spawns ranks with multiprocessing "spawn"
(engine/llm_engine.py:24-31): call install() at module top level of the launcher script so each
spawned rank re-applies it on import of __main__.
"""
from __future__ import annotations

import importlib
import os
import sys

_ORIG_QWEN3_FORWARD = None


def install(verbose: bool = False, fuse_qkv: bool | None = None,
            fast_meta: bool | None = None, fast_rpc: bool | None = None) -> None:
    from . import attention as ours

    if fuse_qkv is None:
        fuse_qkv = os.environ.get("B200ATTN_FUSE_QKV", "1") != "0"
    if fast_meta is None:
        fast_meta = os.environ.get("B200ATTN_FAST_META", "1") != "0"
    if fast_rpc is None:
        fast_rpc = os.environ.get("B200ATTN_FAST_RPC", "1") != "0"

    # Register first, import second: importing ANY nanovllm submodule runs nanovllm/__init__.py,
    # which pulls in llm -> llm_engine -> model_runner -> models.qwen3 -> `from
    # nanovllm.layers.attention import Attention`. With our module already in sys.modules under
    # that name the reference's layers/attention.py (and with it triton and flash_attn) is never
    # executed; our module has no nanovllm imports of its own.
    sys.modules["nanovllm.layers.attention"] = ours
    ref_ctx = importlib.import_module("nanovllm.utils.context")
    ours.get_context = ref_ctx.get_context
    layers = sys.modules.get("nanovllm.layers")
    if layers is not None:
        setattr(layers, "attention", ours)
    qwen3 = sys.modules.get("nanovllm.models.qwen3")
    if qwen3 is not None:
        qwen3.Attention = ours.Attention
    if fuse_qkv:
        global _ORIG_QWEN3_FORWARD
        from . import qkv_attention as fused

        qwen3 = importlib.import_module("nanovllm.models.qwen3")  # binds OUR Attention (step 2)
        cls = getattr(qwen3, "Qwen3Attention", None)
        if cls is None:
            fuse_qkv = False
        elif cls.forward is not fused.qwen3_attention_forward:
            _ORIG_QWEN3_FORWARD = cls.forward
            cls.forward = fused.qwen3_attention_forward
    if fast_meta:
        try:
            mr = importlib.import_module("nanovllm.engine.model_runner")
        except ImportError:
            mr = None
        if mr is not None and hasattr(mr, "ModelRunner"):
            from . import fast_meta as fm

            fm.bind(mr.ModelRunner)
        else:
            fast_meta = False
    if fast_rpc:
        try:
            mr = importlib.import_module("nanovllm.engine.model_runner")
        except ImportError:
            mr = None
        if mr is not None and hasattr(mr, "ModelRunner") and hasattr(mr, "SharedMemory"):
            from . import fast_rpc as rpc

            rpc.bind(mr)
        else:
            fast_rpc = False
    ours.set_decode_overlap(os.environ.get("B200ATTN_DECODE_OVERLAP", "1") != "0")
    if verbose:
        print("[byox_vllm_b200] nanovllm.layers.attention -> byox_vllm_b200.attention"
              + (" (+ fused qkv path)" if fuse_qkv else "")
              + (" (+ device-side decode metadata)" if fast_meta else "")
              + (" (+ compact tensor-parallel step records)" if fast_rpc else ""),
              file=sys.stderr)


def uninstall() -> None:
    from . import attention as ours
    from . import context as own

    ours.get_context = own.get_context
    ours.set_decode_overlap(False)
    global _ORIG_QWEN3_FORWARD
    qwen3 = sys.modules.get("nanovllm.models.qwen3")
    if qwen3 is not None and _ORIG_QWEN3_FORWARD is not None and hasattr(qwen3, "Qwen3Attention"):
        qwen3.Qwen3Attention.forward = _ORIG_QWEN3_FORWARD
        _ORIG_QWEN3_FORWARD = None
    mr = sys.modules.get("nanovllm.engine.model_runner")
    if mr is not None and hasattr(mr, "ModelRunner"):
        from . import fast_meta as fm

        fm.unbind(mr.ModelRunner)
        from . import fast_rpc as rpc

        rpc.unbind(mr)
    if sys.modules.get("nanovllm.layers.attention") is ours:
        del sys.modules["nanovllm.layers.attention"]
