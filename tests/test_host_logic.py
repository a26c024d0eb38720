"""Host-side logic that needs no GPU: argument validation of the C-ABI (errors are returned
before any CUDA call), the split heuristic, the module surface and the install shim."""
import importlib
import os
import subprocess
import sys
import textwrap

import pytest
import torch

from byox_vllm_b200 import _lib


@pytest.fixture(scope="module")
def lib(built_lib):
    return _lib.load()


STREAM = os.environ.get("B200ATTN_DECODE_ENGINE", "rows").startswith("s")


def test_pick_splits_is_host_only_and_sane(lib):
    """Rows engine (default): num_splits = equal pieces per (sequence, KV head) row. Big batches
    never split; tiny batches with long tables split towards one CTA per SM; pieces never shorter
    than 512 tokens of the longest possible context. Stream engine (B200ATTN_DECODE_ENGINE=stream,
    run by test_stream_engine_host_logic below): num_splits = persistent CTAs, two per SM, fewer
    when even full tables could not give every CTA its minimum range (16 units), but never fewer
    than one per row while the rows fit; the cut is decided on the device
    (tests/test_decode_partition.py)."""
    sm = 148
    pick = lib.b200attn_decode_pick_splits
    assert pick(4, 8, 16, 256, 0) >= 1  # unknown SM count -> 148
    assert pick(0, 8, 16, 256, sm) == 1
    if STREAM:
        assert pick(512, 8, 4, 256, sm) == 2 * sm
        assert pick(128, 8, 16, 256, sm) == 2 * sm
        assert pick(256, 1, 16, 256, sm) == 2 * sm  # TP=8 rank of config 2
        assert pick(1, 8, 512, 256, sm) == 2 * sm   # one long sequence
        assert pick(1, 8, 1, 256, sm) == 8          # never fewer CTAs than rows
        assert pick(1, 8, 8, 256, sm) == 16         # multiples of Hkv beyond
        assert pick(1, 1, 1, 256, sm) == 1
        for b in (1, 2, 4, 8, 16, 64):
            for w in (1, 4, 16, 128):
                s = pick(b, 1, w, 256, sm)
                assert 1 <= s <= 2 * sm
                assert s <= max(b, -(-b * (4 * w + 1) // 16))
        return
    assert pick(512, 8, 4, 256, sm) == 1
    assert pick(128, 8, 16, 256, sm) == 1
    assert 16 <= pick(1, 8, 512, 256, sm) <= 128
    assert pick(1, 8, 1, 256, sm) == 1
    for b in (1, 2, 4, 8, 16, 64):
        for w in (1, 4, 16, 128):
            assert 1 <= pick(b, 1, w, 256, sm) <= max(1, w)
    # calibration points of tools/sweep_splits.py (profiles/r01_sweep_splits.log): about one CTA
    # per SM, no split once (sequence, KV head) rows reach ~0.7 x SMs, pieces >= 512 tokens
    assert pick(16, 8, 16, 256, sm) == 1
    assert pick(128, 1, 16, 256, sm) == 1
    assert pick(8, 8, 64, 256, sm) == 2
    assert pick(4, 8, 16, 256, sm) == 5
    assert pick(1, 8, 16, 256, sm) == 8
    assert pick(1, 8, 512, 256, sm) in (18, 19)


def test_workspace_bytes(lib):
    slot = 8 * (128 + 1) * 4  # one partial piece: [8 heads][128] fp32 O + [8] lse
    ws = lib.b200attn_decode_workspace_bytes
    pre = 4 * 16 * 5 * (128 + 1) * 4
    assert ws(4, 16, 128, 1, 0) == 0
    assert ws(4, 16, 128, 1, 5) == pre
    assert ws(4, 16, 64, 3, 0) == 0  # unsupported head_dim
    if STREAM:
        # two slots per work item that can hold a cut row (<= num_splits + num_q_heads items)
        assert ws(4, 16, 128, 3, 0) == (3 + 16) * 2 * slot
        assert ws(400, 16, 128, 296, 0) == (296 + 16) * 2 * slot
        assert ws(4, 16, 128, 3, 5) == (3 + 16) * 2 * slot + pre
    else:
        # one slot per (sequence, KV head, split), KV heads bounded by the q heads
        assert ws(4, 16, 128, 3, 0) == 4 * 16 * 3 * slot
        assert ws(4, 16, 128, 3, 5) == 4 * 16 * 3 * slot + pre


def test_stream_engine_host_logic():
    """The same host-logic and partition tests with the persistent stream engine selected (the
    engine is read once per process)."""
    if STREAM:
        pytest.skip("already the stream engine")
    env = dict(os.environ, B200ATTN_DECODE_ENGINE="stream")
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "not gpu", __file__,
                        "-k", "pick_splits or workspace_bytes"], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_argument_errors_have_codes_and_messages(lib):
    # all rejected before touching CUDA, so this runs without a GPU
    rc = lib.b200attn_store_kvcache(None, 1024, None, 1024, None, None, None, 4, 1024, 64, 0, None)
    assert rc == _lib.E_BADARG and b"null" in lib.b200attn_last_error()
    rc = lib.b200attn_store_kvcache(16, 1024, 16, 1024, 16, 16, 16, 4, 1020, 64, 0, None)
    assert rc == _lib.E_UNSUPPORTED
    rc = lib.b200attn_store_kvcache(16, 1024, 16, 1024, 16, 16, 16, 4, 1024, 64, 7, None)
    assert rc == _lib.E_UNSUPPORTED and b"dtype" in lib.b200attn_last_error()
    rc = lib.b200attn_store_kvcache(16, 1024, 16, 1024, 16, 16, 16, 0, 1024, 64, 0, None)
    assert rc == 0  # empty batch is a no-op
    rc = lib.b200attn_decode(16, 16, 16, 16, 4, 4, 16, 16, None, 0, 2, 16, 8, 64, 256, 8, 0.1, 1, 0,
                             None)
    assert rc == _lib.E_UNSUPPORTED and b"head_dim" in lib.b200attn_last_error()
    rc = lib.b200attn_decode(16, 16, 16, 16, 4, 4, 16, 16, None, 0, 2, 16, 5, 128, 256, 8, 0.1, 1, 0,
                             None)
    assert rc == _lib.E_BADARG  # Hq % Hkv != 0
    rc = lib.b200attn_decode(16, 16, 16, 16, 4, 4, 16, 16, None, 0, 2, 16, 8, 128, 256, 8, 0.1, 4, 0,
                             None)
    assert rc == _lib.E_WORKSPACE
    rc = lib.b200attn_prefill_varlen(16, None, 0, None, 0, None, None, None, 0, 0, 16, 16, 1, 8, 8,
                                     8, 8, 16, 16, 8, 128, 0, 0, 0.1, 0, None)
    assert rc == _lib.E_BADARG  # dense prefill without key/value
    with pytest.raises(_lib.B200AttnError) as ei:
        _lib.check("b200attn_decode", lib.b200attn_decode(
            16, 16, 16, 16, 4, 4, 16, 16, None, 0, 2, 16, 8, 64, 256, 8, 0.1, 1, 0, None))
    assert ei.value.code == _lib.E_UNSUPPORTED


def test_attention_module_surface_and_no_cpu_fallback(built_lib):
    from byox_vllm_b200.attention import Attention, store_kvcache
    from byox_vllm_b200.context import reset_context, set_context

    a = Attention(16, 128, 128 ** -0.5, 8)  # positional ctor as reference models/qwen3.py:60-65
    assert (a.num_heads, a.head_dim, a.num_kv_heads) == (16, 128, 8)
    assert a.k_cache.numel() == 0 and a.v_cache.numel() == 0  # placeholders (attention.py:65)
    assert hasattr(a, "k_cache") and hasattr(a, "v_cache")  # what ModelRunner looks for
    assert list(a.state_dict().keys()) == []  # caches are plain attributes, not buffers
    q = torch.zeros(2, 16, 128, dtype=torch.bfloat16)
    k = torch.zeros(2, 8, 128, dtype=torch.bfloat16)
    set_context(True, torch.tensor([0, 2], dtype=torch.int32), torch.tensor([0, 2], dtype=torch.int32),
                2, 2, torch.zeros(0, dtype=torch.int32), None, None)
    with pytest.raises(RuntimeError, match="no CPU path"):
        a(q, k, k)
    reset_context()
    kc = torch.zeros(1, 16, 8, 128, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU path"):
        store_kvcache(k, k, kc, kc.clone(), torch.zeros(2, dtype=torch.int32))
    with pytest.raises(AssertionError):  # reference precondition attention.py:46
        store_kvcache(k, k, kc, kc.clone(), torch.zeros(3, dtype=torch.int32))


def test_product_never_imports_oracle():
    from conftest import ROOT

    for path in (ROOT / "byox_vllm_b200").rglob("*"):
        if path.suffix in (".py", ".cu", ".cuh", ".h"):
            text = path.read_text()
            assert "oracle" not in text.replace("oracle/", "").lower() or path.name == "__init__.py", path
            assert "import triton" not in text and "flash_attn import" not in text, path


def test_install_under_a_nanovllm_package(tmp_path, monkeypatch, built_lib):
    """install() must make `from nanovllm.layers.attention import Attention` resolve to ours and
    route set_context() of nanovllm.utils.context into our forward — checked on a stand-in package
    with the reference's module layout (the real one needs CUDA-only deps to import)."""
    pkg = tmp_path / "nanovllm"
    (pkg / "utils").mkdir(parents=True)
    (pkg / "layers").mkdir()
    (pkg / "models").mkdir()
    for d in (pkg / "utils", pkg / "layers", pkg / "models"):
        (d / "__init__.py").write_text("")
    # like the real package: importing ANY nanovllm submodule runs __init__ -> llm -> ... ->
    # models.qwen3 -> `from nanovllm.layers.attention import Attention`
    (pkg / "__init__.py").write_text("from nanovllm.models import qwen3  # noqa: F401\n")
    (pkg / "utils" / "context.py").write_text(textwrap.dedent("""
        from dataclasses import dataclass
        @dataclass
        class Context:
            is_prefill: bool = False
            marker: str = "reference-context"
        _CONTEXT = Context()
        def get_context():
            return _CONTEXT
    """))
    (pkg / "layers" / "attention.py").write_text("raise ImportError('needs triton + flash_attn')\n")
    (pkg / "models" / "qwen3.py").write_text("from nanovllm.layers.attention import Attention\n")
    monkeypatch.syspath_prepend(str(tmp_path))
    for m in [m for m in sys.modules if m == "nanovllm" or m.startswith("nanovllm.")]:
        monkeypatch.delitem(sys.modules, m)
    from byox_vllm_b200 import attention as ours, install

    install.install()
    try:
        qwen3 = importlib.import_module("nanovllm.models.qwen3")
        assert qwen3.Attention is ours.Attention
        assert ours.get_context().marker == "reference-context"
    finally:
        install.uninstall()
        for m in [m for m in sys.modules if m == "nanovllm" or m.startswith("nanovllm.")]:
            sys.modules.pop(m, None)
    from byox_vllm_b200.context import get_context

    assert ours.get_context is get_context


_BLOCKED_INSTALL = r"""
import sys, importlib.abc
class Block(importlib.abc.MetaPathFinder):
    def find_spec(self, name, path=None, target=None):
        if name.split(".")[0] in ("flash_attn", "flash_attn_2_cuda", "triton"):
            raise ImportError(name + " is blocked for this test")
        return None
sys.meta_path.insert(0, Block())
sys.path.insert(0, {root!r}); sys.path.insert(0, {ref!r})
from byox_vllm_b200 import install
install.install(fuse_qkv=True, fast_meta=True, fast_rpc=True)
import nanovllm.models.qwen3 as qwen3
import nanovllm.engine.model_runner as mr
import nanovllm.utils.context as ctx
from byox_vllm_b200 import attention as ours, fast_meta, qkv_attention
assert qwen3.Attention is ours.Attention and ours.get_context is ctx.get_context
assert qwen3.Qwen3Attention.forward is qkv_attention.qwen3_attention_forward
assert mr.ModelRunner._prepare_decode is fast_meta.prepare_decode
bad = sorted(m for m in sys.modules if m.split(".")[0] in ("flash_attn", "triton"))
assert not bad, bad
print("INSTALL-OK")
"""


def test_install_never_imports_flash_attn_or_triton(built_lib):
    """The UNMODIFIED reference (baseline/_ref) with install() in a fresh interpreter where
    flash_attn and triton cannot be imported at all: the engine modules import, every rebinding is
    in place, and neither wheel is in sys.modules (install() registers our module under
    nanovllm.layers.attention BEFORE the first nanovllm import)."""
    import subprocess

    from conftest import ROOT

    ref = ROOT / "baseline" / "_ref"
    if not (ref / "nanovllm" / "layers" / "attention.py").exists():
        pytest.skip("baseline/_ref is not installed (needs /root/reference: __graft_entry__.build)")
    r = subprocess.run([sys.executable, "-c", _BLOCKED_INSTALL.format(root=str(ROOT), ref=str(ref))],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "INSTALL-OK" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


def test_cascade_chunks_heuristic_is_host_only(lib):
    """Armed only where a shared prefix can pay: batch >= 64, tables holding >= 8192 tokens, pages
    that are multiples of 128 tokens; about one prefix CTA per SM; never more chunks than tiles."""
    sm = 148
    pick = lib.b200attn_cascade_pick_chunks
    assert pick(256, 16, 8, 16, 256, sm) == 0      # bench shape c2: tables hold 4096 tokens
    assert pick(128, 32, 8, 17, 256, sm) == 0      # c3
    assert pick(32, 8, 1, 40, 256, sm) == 0        # small batch
    assert pick(256, 8, 1, 40, 64, sm) == 0        # 64-token pages
    c4 = pick(256, 8, 1, 40, 256, sm)              # config 4 per rank: 8 row-tile pairs x 1 head
    assert c4 == 16
    assert pick(256, 64, 8, 40, 256, sm) == 2      # 32B at TP=1: 8 pairs x 8 KV heads = 64 CTAs
    assert pick(256, 32, 4, 40, 256, sm) == 4      # 32 CTAs: 4 chunks = 128 CTAs, one wave
    assert 1 <= pick(512, 32, 8, 64, 256, 0) <= 16  # unknown SM count -> 148
    assert pick(0, 8, 1, 40, 256, sm) == 0 and pick(256, 8, 1, 0, 256, sm) == 0


def test_cascade_argument_errors(lib):
    from byox_vllm_b200._lib import E_BADARG

    z = 0x1000  # fake non-null, 16-byte aligned device pointers: rejected before any CUDA call
    rc = lib.b200attn_store_decode_cascade(z, None, 0, None, 0, None, z, z, z, 4, 4, z, z, None, 0,
                                           2, 8, 2, 128, 256, 8, 0.1, 1, None, 3, 0, None)
    assert rc == E_BADARG and b"plan" in lib.b200attn_last_error()
    rc = lib.b200attn_store_decode_cascade(z, None, 0, None, 0, None, z, z, z, 4, 4, z, z, None, 0,
                                           2, 8, 2, 128, 256, 8, 0.1, 1, z, 0, 0, None)
    assert rc == E_BADARG
    rc = lib.b200attn_store_decode_cascade(z, None, 0, None, 0, None, z, z, z, 4, 4, z, z, None, 0,
                                           2, 8, 2, 128, 256, 8, 0.1, 1, z, 3, 0, None)
    assert rc == _lib.E_WORKSPACE  # a cascade always needs the partial workspace
    assert lib.b200attn_cascade_plan(z, 4, 4, z, 2, 256, None, None) == E_BADARG


def test_shared_prefix_oracle_on_block_manager_style_tables():
    """The plan checker itself: tables as BlockManager hands them out for prefix hits
    (reference engine/block_manager.py:41-71 reuses the cached block ids in order)."""
    from oracle import attention_oracle as orc
    from tools import synth

    c = synth.shared_prefix_decode_case([900, 1000, 1100, 800], 3, 8, 2, seed=1)
    assert orc.shared_prefix_pages_ref(c.meta["context_lens"], c.meta["block_tables"], 256) == \
        (3, 0, 4)
    c = synth.shared_prefix_decode_case([300, 1000, 1100], 3, 8, 2, seed=1)
    assert orc.shared_prefix_pages_ref(c.meta["context_lens"], c.meta["block_tables"], 256)[0] == 1
    c = synth.shared_prefix_decode_case([512, 769], 2, 8, 2, seed=1)  # token 511 opens page 1
    assert orc.shared_prefix_pages_ref(c.meta["context_lens"], c.meta["block_tables"], 256)[0] == 1
    c = synth.shared_prefix_decode_case([900, 1000, 1100], [3, 1, 3], 8, 2, seed=1)
    assert orc.shared_prefix_pages_ref(c.meta["context_lens"], c.meta["block_tables"], 256)[0] == 1
    lens = c.meta["context_lens"].clone()
    lens[1] = 0  # the diverging row becomes a padding row
    assert orc.shared_prefix_pages_ref(lens, c.meta["block_tables"], 256) == (3, 0, 2)
    one = synth.shared_prefix_decode_case([2000], 4, 8, 2, seed=1)
    assert orc.shared_prefix_pages_ref(one.meta["context_lens"], one.meta["block_tables"], 256) == \
        (0, 0, 1)
