"""Put the B200 attention path under nanovllm without touching nanovllm's sources.

The reference binds the op by name: `from nanovllm.layers.attention import Attention`
(models/qwen3.py:7) and reads metadata through `nanovllm.utils.context.get_context`
(layers/attention.py:7,68). install() therefore
  1. imports nanovllm.utils.context (pure Python, no CUDA deps) and points our module's
     `get_context` at it, so ModelRunner.set_context(...) feeds our forward();
  2. registers byox_vllm_b200.attention as `nanovllm.layers.attention` in sys.modules BEFORE
     nanovllm.models.qwen3 is imported, so neither triton nor flash_attn is ever imported;
  3. if nanovllm.models.qwen3 was already imported, rebinds its `Attention` name and swaps the
     class of existing instances' constructor target.
For tensor parallel runs nanovllm spawns ranks with multiprocessing "spawn"
(engine/llm_engine.py:24-31): call install() at module top level of the launcher script so each
spawned rank re-applies it on import of __main__.
"""
from __future__ import annotations

import importlib
import sys


def install(verbose: bool = False) -> None:
    from . import attention as ours

    ref_ctx = importlib.import_module("nanovllm.utils.context")
    ours.get_context = ref_ctx.get_context
    sys.modules["nanovllm.layers.attention"] = ours
    layers = sys.modules.get("nanovllm.layers")
    if layers is not None:
        setattr(layers, "attention", ours)
    qwen3 = sys.modules.get("nanovllm.models.qwen3")
    if qwen3 is not None:
        qwen3.Attention = ours.Attention
    if verbose:
        print("[byox_vllm_b200] nanovllm.layers.attention -> byox_vllm_b200.attention",
              file=sys.stderr)


def uninstall() -> None:
    from . import attention as ours
    from . import context as own

    ours.get_context = own.get_context
    if sys.modules.get("nanovllm.layers.attention") is ours:
        del sys.modules["nanovllm.layers.attention"]
