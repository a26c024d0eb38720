"""-m gpu: the fp16 build of the path (C-ABI dtype code 1). The reference takes its dtype from the
checkpoint (engine/model_runner.py:30) and both of its kernels accept fp16
(layers/attention.py:10-31 is a byte copy, flash-attn takes fp16 or bf16); SURVEY §8(b): "bf16 (or
fp16)". Store, decode (splits, fused scatter, cascade) and prefill (tcgen05 and the mma.sync shape
fallback) are checked against the dtype-agnostic fp32 oracle on fp16 inputs, with the same
tolerance as bf16 (2e-2 + 1e-2 |ref|); the fused q/k norm + RoPE entry points stay bf16 only and
must say so."""
import copy

import pytest
import torch

from gpu_util import assert_close, to_dev
from oracle import attention_oracle as orc
from tools import synth

pytestmark = pytest.mark.gpu
F16 = torch.float16


def _to_f16(case):
    """A synth case with every 16-bit tensor converted to fp16 (bf16 N(0,1) values are exact in
    fp16 up to its subnormal range)."""
    c = copy.copy(case)
    for name in ("q", "k", "k_cache", "v_cache"):
        setattr(c, name, getattr(case, name).to(F16))
    v = case.v
    n, Hkv, D = v.shape
    packed = torch.zeros(n, max(v.stride(0), Hkv * D) if n > 1 else Hkv * D * 3, dtype=F16)
    vv = packed[:, packed.size(1) - Hkv * D:].view(n, Hkv, D)
    vv.copy_(v)
    c.v = vv
    return c


def test_store_fp16_is_a_byte_copy(cuda_dev):
    from byox_vllm_b200.attention import store_kvcache

    c = _to_f16(synth.decode_case(37, 300, 16, 8, seed=3, fill_cache=False))
    d = to_dev(c, cuda_dev)
    store_kvcache(d["k"], d["v"], d["k_cache"], d["v_cache"], d["meta"]["slot_mapping"])
    torch.cuda.synchronize()
    orc.store_kvcache_ref(c.k, c.v, c.k_cache, c.v_cache, c.meta["slot_mapping"])
    assert torch.equal(d["k_cache"].cpu().view(torch.int16), c.k_cache.view(torch.int16))
    assert torch.equal(d["v_cache"].cpu().view(torch.int16), c.v_cache.view(torch.int16))


@pytest.mark.parametrize("splits", [None, 1, 3])
@pytest.mark.parametrize("heads", [(16, 8), (32, 8), (8, 1)], ids=["16_8", "32_8", "8_1"])
def test_decode_fp16_matches_oracle(cuda_dev, heads, splits):
    from byox_vllm_b200.attention import paged_decode

    Hq, Hkv = heads
    lens = [1, 63, 64, 65, 700, 256, 1500, 300, 2049, 17, 4000]
    c = _to_f16(synth.decode_case(len(lens), lens, Hq, Hkv, seed=21))
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    got = paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"], m["block_tables"],
                       c.scale, num_splits=splits)
    torch.cuda.synchronize()
    assert got.dtype == F16
    want = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                                c.meta["block_tables"], c.scale)
    assert_close(got, want, f"fp16 decode {Hq}/{Hkv} splits={splits}")


def test_fused_store_decode_fp16(cuda_dev):
    from byox_vllm_b200.attention import Attention
    from byox_vllm_b200.context import reset_context, set_context
    from tools.metadata_ref import context_kwargs

    lens = [513, 1, 256, 257, 90, 1024, 3000]
    c = _to_f16(synth.decode_case(len(lens), lens, 16, 8, seed=22))
    d = to_dev(c, cuda_dev)
    attn = Attention(16, 128, c.scale, 8)
    attn.k_cache, attn.v_cache = d["k_cache"], d["v_cache"]
    set_context(**context_kwargs(c.meta, cuda_dev))
    try:
        got = attn(d["q"], d["k"], d["v"])
        torch.cuda.synchronize()
    finally:
        reset_context()
    ctx = type("Ctx", (), dict(c.meta))()
    want = orc.attention_forward_ref(c.q, c.k, c.v, c.k_cache, c.v_cache, ctx, c.scale)
    assert_close(got, want, "fp16 Attention.forward decode")
    assert torch.equal(d["k_cache"].cpu().view(torch.int16), c.k_cache.view(torch.int16))
    assert torch.equal(d["v_cache"].cpu().view(torch.int16), c.v_cache.view(torch.int16))


def test_cascade_fp16_matches_plain_and_oracle(cuda_dev):
    from byox_vllm_b200.attention import cascade_plan, paged_decode

    lens = [2304 + 17 * i for i in range(70)]
    c = _to_f16(synth.shared_prefix_decode_case(lens, 8, 8, 1, seed=23))
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    plan = cascade_plan(m["context_lens"], m["block_tables"], 256)
    got = paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"], m["block_tables"],
                       c.scale, plan=plan, prefix_chunks=3)
    plain = paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"], m["block_tables"],
                         c.scale)
    torch.cuda.synchronize()
    assert int(plan[0]) == 8
    want = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                                c.meta["block_tables"], c.scale)
    assert_close(got, want, "fp16 cascade vs oracle")
    assert_close(got, plain, "fp16 cascade vs plain decode")


@pytest.mark.parametrize("case", [([300, 1, 1027, 64, 129, 700], None, 16, 8, 256),
                                  ([600, 1300, 257], [512, 1024, 0], 8, 1, 256),
                                  ([2100], None, 32, 8, 256),
                                  ([200, 131], [64, 0], 16, 8, 64)],
                         ids=["dense", "paged_prefix", "long", "mma_fallback_block64"])
def test_prefill_fp16_matches_oracle(cuda_dev, case):
    from byox_vllm_b200.attention import Attention
    from byox_vllm_b200.context import reset_context, set_context
    from tools.metadata_ref import context_kwargs

    lens, cached, Hq, Hkv, bs = case
    c = _to_f16(synth.prefill_case(lens, Hq, Hkv, cached=cached, seed=24, block_size=bs))
    d = to_dev(c, cuda_dev)
    attn = Attention(Hq, 128, c.scale, Hkv)
    attn.k_cache, attn.v_cache = d["k_cache"], d["v_cache"]
    set_context(**context_kwargs(c.meta, cuda_dev))
    try:
        got = attn(d["q"], d["k"], d["v"])
        torch.cuda.synchronize()
    finally:
        reset_context()
    assert got.dtype == F16
    ctx = type("Ctx", (), dict(c.meta))()
    want = orc.attention_forward_ref(c.q, c.k, c.v, c.k_cache, c.v_cache, ctx, c.scale)
    assert_close(got, want, f"fp16 prefill {lens}")


def test_fused_qkv_entry_points_are_bf16_only(cuda_dev):
    from byox_vllm_b200.qkv_attention import qk_norm_rope_store

    qkv = torch.zeros(4, (16 + 16) * 128, dtype=F16, device=cuda_dev)
    pos = torch.zeros(4, dtype=torch.int64, device=cuda_dev)
    cs = torch.zeros(16, 128, dtype=torch.float32, device=cuda_dev)
    with pytest.raises(RuntimeError, match="bf16"):
        qk_norm_rope_store(qkv, pos, None, None, 0.0, cs, 16, 8)
