"""-m gpu: varlen causal prefill (dense and paged-prefix) through the C-ABI vs the fp32 oracle,
the flash-attn golden vectors and size-independent properties."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN
from gpu_util import assert_close, to_dev
from oracle import attention_oracle as orc
from tools import golden_cases as gc, synth

pytestmark = pytest.mark.gpu


def _forward(dev, c, Hq, Hkv):
    """Attention.forward on the device for a prefill case; returns (out, device dict)."""
    from byox_vllm_b200.attention import Attention
    from byox_vllm_b200.context import reset_context, set_context
    from tools.metadata_ref import context_kwargs

    d = to_dev(c, dev)
    attn = Attention(Hq, 128, c.scale, Hkv)
    attn.k_cache, attn.v_cache = d["k_cache"], d["v_cache"]
    set_context(**context_kwargs(c.meta, dev))
    try:
        o = attn(d["q"], d["k"], d["v"])
        torch.cuda.synchronize()
    finally:
        reset_context()
    return o, d


@pytest.mark.parametrize("case", gc.PREFILL_CASES, ids=[c[0] for c in gc.PREFILL_CASES])
def test_prefill_matches_oracle_and_flash_attn(cuda_dev, case):
    name, lens, cached, Hq, Hkv, seed = case
    c = synth.prefill_case(lens, Hq, Hkv, cached=cached, seed=seed)
    got, d = _forward(cuda_dev, c, Hq, Hkv)
    assert got.shape == c.q.shape and got.dtype == torch.bfloat16
    ctx = type("Ctx", (), dict(c.meta))()
    want = orc.attention_forward_ref(c.q, c.k, c.v, c.k_cache, c.v_cache, ctx, c.scale)
    assert_close(got, want, f"{name} vs oracle")
    # the scatter that precedes attention left the cache bit-identical to the oracle's
    assert torch.equal(d["k_cache"].cpu().view(torch.int16), c.k_cache.view(torch.int16))
    assert torch.equal(d["v_cache"].cpu().view(torch.int16), c.v_cache.view(torch.int16))
    g = np.load(GOLDEN / f"fa_prefill_{name}.npz")
    fa = torch.from_numpy(g["out"]).view(torch.bfloat16)
    assert_close(got, fa, f"{name} vs flash-attn golden")


@pytest.mark.parametrize("Hq,Hkv", [(16, 8), (32, 8), (8, 1), (4, 1)])
def test_prefill_more_shapes_vs_oracle(cuda_dev, Hq, Hkv):
    lens = [1, 2, 64, 127, 128, 129, 300]
    c = synth.prefill_case(lens, Hq, Hkv, seed=50 + Hq)
    got, _ = _forward(cuda_dev, c, Hq, Hkv)
    ctx = type("Ctx", (), dict(c.meta))()
    want = orc.attention_forward_ref(c.q, c.k, c.v, c.k_cache, c.v_cache, ctx, c.scale)
    assert_close(got, want, f"dense {Hq}/{Hkv}")
    c = synth.prefill_case([257, 520, 1100], Hq, Hkv, cached=[256, 256, 1024], seed=60 + Hq)
    got, _ = _forward(cuda_dev, c, Hq, Hkv)
    ctx = type("Ctx", (), dict(c.meta))()
    want = orc.attention_forward_ref(c.q, c.k, c.v, c.k_cache, c.v_cache, ctx, c.scale)
    assert_close(got, want, f"paged prefix {Hq}/{Hkv}")


SCHEDULE_CASES = [
    # name, lens, cached, Hq, Hkv: the persistent kernel's item list (csrc/prefill_tc.cu)
    ("many_short_16_8", "rand:220:1:90", None, 16, 8),          # > 148 CTAs' worth of 1-tile items
    ("bench_prompts_16_8", "rand:40:100:1024", None, 16, 8),     # the reference bench's prefill mix
    ("long_and_short_32_8", [2300, 3, 700, 1, 129, 64, 1500, 31], None, 32, 8),
    ("cached_mix_8_1", [600, 300, 1300, 200, 2049], [512, 0, 1024, 0, 1792], 8, 1),
    ("mha_4_4", [513, 77, 256, 1030], None, 4, 4),               # G = 1: 128 tokens per sub-tile
    ("gqa8_one_seq_8_1", [3000], None, 8, 1),                    # single sequence: unsorted list
]


def _schedule_lens(spec, seed):
    if not isinstance(spec, str):
        return list(spec)
    import random

    _, n, lo, hi = spec.split(":")
    rng = random.Random(seed)
    return [rng.randint(int(lo), int(hi)) for _ in range(int(n))]


@pytest.mark.parametrize("case", SCHEDULE_CASES, ids=[c[0] for c in SCHEDULE_CASES])
def test_prefill_persistent_schedule(cuda_dev, case):
    """Every (sequence, pair, KV head) item is visited exactly once whatever the mix: all rows of
    all sequences match the oracle (a skipped item leaves rows of the poisoned output untouched, a
    doubled one is harmless but a wrong barrier phase is not), twice in a row (no state between
    launches)."""
    from byox_vllm_b200.attention import prefill_varlen

    name, spec, cached, Hq, Hkv = case
    lens = _schedule_lens(spec, seed=len(name))
    c = synth.prefill_case(lens, Hq, Hkv, cached=cached, seed=90 + len(name))
    got, d = _forward(cuda_dev, c, Hq, Hkv)
    ctx = type("Ctx", (), dict(c.meta))()
    want = orc.attention_forward_ref(c.q, c.k, c.v, c.k_cache, c.v_cache, ctx, c.scale)
    assert_close(got, want, f"{name} vs oracle")
    m = c.meta
    paged = m.get("block_tables") is not None
    out = torch.full_like(d["q"], float("nan"))
    for _ in range(2):
        prefill_varlen(d["q"], d["k_cache"] if paged else d["k"], d["v_cache"] if paged else d["v"],
                       m["cu_seqlens_q"].to(cuda_dev), m["cu_seqlens_k"].to(cuda_dev),
                       m["max_seqlen_q"], m["max_seqlen_k"], c.scale,
                       block_tables=m["block_tables"].to(cuda_dev) if paged else None, out=out)
        torch.cuda.synchronize()
        assert torch.equal(out.view(torch.int16), got.view(torch.int16)), name
        out.fill_(float("nan"))


def test_prefill_more_sequences_than_the_schedule_holds(cuda_dev):
    """> 1023 sequences in one call: the tcgen05 kernel's shared-memory schedule does not hold
    them, the call takes the mma.sync shape fallback (csrc/api.cu) - same results."""
    lens = [1 + (i * 7) % 5 for i in range(1100)]
    c = synth.prefill_case(lens, 16, 8, seed=77)
    got, _ = _forward(cuda_dev, c, 16, 8)
    ctx = type("Ctx", (), dict(c.meta))()
    want = orc.attention_forward_ref(c.q, c.k, c.v, c.k_cache, c.v_cache, ctx, c.scale)
    assert_close(got, want, "1100 sequences")


def test_prefill_more_items_than_the_sort_holds(cuda_dev):
    """Two sequences of 66 000 tokens at heads 8/1 (32 tokens per 256-row item): 4126 items per KV
    head, more than the 4096 sort keys the schedule holds -> the kernel walks them in descending
    index order instead of cost order (csrc/prefill_tc.cu: use_keys == 0 with several sequences).
    Sampled rows against an fp32 einsum on the device; every row written."""
    from byox_vllm_b200.attention import prefill_varlen

    dev = cuda_dev
    L, S, Hq, Hkv = 66000, 2, 8, 1
    g = torch.Generator(device=dev).manual_seed(3)
    q, k, v = synth._qkv(L * S, Hq, Hkv, 128, g, dev)
    cu = torch.arange(0, S + 1, dtype=torch.int32, device=dev) * L
    scale = 128 ** -0.5
    o = torch.full_like(q, float("nan"))
    prefill_varlen(q, k, v, cu, cu, L, L, scale, out=o)
    torch.cuda.synchronize()
    assert not torch.isnan(o.float()).any(), "an item was skipped"
    for s, r in ((0, 0), (0, 31), (0, 32), (0, 40000), (0, L - 1), (1, 0), (1, 12345), (1, L - 1)):
        i = s * L + r
        kk = k[s * L:s * L + r + 1].float().repeat_interleave(Hq // Hkv, 1)
        vv = v[s * L:s * L + r + 1].float().repeat_interleave(Hq // Hkv, 1)
        p_ = torch.softmax(torch.einsum("hd,khd->hk", q[i].float(), kk) * scale, -1)
        want = torch.einsum("hk,khd->hd", p_, vv)
        assert_close(o[i], want.cpu(), f"sequence {s} row {r}")


def test_prefill_under_cuda_graph_capture_and_replay(cuda_dev):
    """The persistent kernel (device-built schedule, hardware work queue) is captured like any other
    launch; a replay with other lengths (same total / buffers, new cu_seqlens contents) follows the
    device data, bit-identical to the eager call."""
    from byox_vllm_b200.attention import prefill_varlen

    dev = cuda_dev
    Hq, Hkv, total, S = 16, 8, 3000, 6
    g = torch.Generator(device=dev).manual_seed(9)
    q, k, v = synth._qkv(total, Hq, Hkv, 128, g, dev)
    cu = torch.zeros(S + 1, dtype=torch.int32, device=dev)
    out = torch.empty_like(q)
    scale = 128 ** -0.5

    def set_lens(lens):
        cu.copy_(torch.tensor([0] + list(__import__("itertools").accumulate(lens)),
                              dtype=torch.int32))

    set_lens([500, 500, 500, 500, 500, 500])
    prefill_varlen(q, k, v, cu, cu, total, total, scale, out=out)  # eager warm-up (attributes)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        prefill_varlen(q, k, v, cu, cu, total, total, scale, out=out)
    for lens in ([1, 2, 997, 1000, 700, 300], [3000, 0, 0, 0, 0, 0], [129, 127, 128, 1, 2614, 1]):
        set_lens(lens)
        out.fill_(float("nan"))
        graph.replay()
        torch.cuda.synchronize()
        want = prefill_varlen(q, k, v, cu, cu, total, total, scale)
        torch.cuda.synchronize()
        assert torch.equal(out.view(torch.int16), want.view(torch.int16)), lens


def test_warmup_path_without_cache(cuda_dev):
    """_warmup_model (model_runner.py:102-113): empty caches, no slots, dense attention."""
    from byox_vllm_b200.attention import Attention
    from byox_vllm_b200.context import reset_context, set_context
    from tools.metadata_ref import SeqState, context_kwargs, prepare_prefill

    dev = cuda_dev
    meta = prepare_prefill([SeqState(300), SeqState(300)], 256)
    g = torch.Generator().manual_seed(5)
    q, k, v = synth._qkv(600, 16, 8, 128, g, "cpu")
    attn = Attention(16, 128, 128 ** -0.5, 8)
    set_context(**context_kwargs(meta, dev))
    try:
        o = attn(q.to(dev), k.to(dev), v.to(dev))  # v.to() keeps the strided view's strides
        torch.cuda.synchronize()
    finally:
        reset_context()
    want = orc.prefill_varlen_ref(q, k, v, meta["cu_seqlens_q"], meta["cu_seqlens_k"], 128 ** -0.5)
    assert_close(o, want, "warm-up dense prefill")


def test_prefill_last_row_equals_decode(cuda_dev):
    """Consistency between the two kernels: the last prefill row of a sequence is a decode step
    over the same keys (bottom-right causal alignment)."""
    from byox_vllm_b200.attention import paged_decode

    c = synth.prefill_case([700, 129], 16, 8, seed=70)
    got, d = _forward(cuda_dev, c, 16, 8)
    ends = c.meta["cu_seqlens_q"][1:].long() - 1
    qd = d["q"][ends.to(cuda_dev)]
    lens = torch.tensor([700, 129], dtype=torch.int32, device=cuda_dev)
    from tools.metadata_ref import prepare_block_tables

    bt = prepare_block_tables(c.seqs).to(cuda_dev)
    od = paged_decode(qd, d["k_cache"], d["v_cache"], lens, bt, c.scale)
    torch.cuda.synchronize()
    assert_close(od, got[ends.to(cuda_dev)], "decode vs last prefill row")


def test_prefill_properties_at_warmup_size(cuda_dev):
    """4 x 4096 tokens (the reference's warm-up batch, model_runner.py:110-112), Hq 32 / Hkv 8:
    (a) causality — changing keys/values after position p leaves rows <= p bit-identical;
    (b) constant V gives the constant; (c) sampled rows match the oracle."""
    from byox_vllm_b200.attention import prefill_varlen

    dev = cuda_dev
    L, S, Hq, Hkv = 4096, 4, 32, 8
    g = torch.Generator(device=dev).manual_seed(1)
    q, k, v = synth._qkv(L * S, Hq, Hkv, 128, g, dev)
    cu = torch.arange(0, S + 1, dtype=torch.int32, device=dev) * L
    scale = 128 ** -0.5
    o = prefill_varlen(q, k, v, cu, cu, L, L, scale)
    p = 2500
    k2, v2 = k.clone(), v.contiguous().clone()
    k2[p + 1:L] = 3.0
    v2[p + 1:L] = -5.0
    o2 = prefill_varlen(q, k2, v2, cu, cu, L, L, scale)
    torch.cuda.synchronize()
    assert torch.equal(o[:p + 1].view(torch.int16), o2[:p + 1].view(torch.int16))
    assert torch.equal(o[L:].view(torch.int16), o2[L:].view(torch.int16))
    oc = prefill_varlen(q, k, torch.full_like(v2, 0.5), cu, cu, L, L, scale)
    assert (oc.float() - 0.5).abs().max().item() <= 1e-2
    rows = [0, 1, 63, 64, 2047, 4095]
    for r in rows:
        i = L + r  # second sequence
        s = torch.einsum("hd,khd->hk", q[i].float(),
                         k[L:L + r + 1].float().repeat_interleave(Hq // Hkv, 1)) * scale
        want = torch.einsum("hk,khd->hd", torch.softmax(s, -1),
                            v[L:L + r + 1].float().repeat_interleave(Hq // Hkv, 1))
        assert_close(o[i], want.cpu(), f"row {r}")
