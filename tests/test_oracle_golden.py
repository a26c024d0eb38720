"""Pin the oracle: against the reference's own Triton kernel (scatter, bit-exact) and against
flash-attn run on a B200 through the reference's call sites (decode / prefill, bf16 tolerance)."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import attention_oracle as orc
from tools import golden_cases as gc, synth

ATOL, RTOL = 2e-2, 1e-2  # north_star: 2e-2 absolute / 1e-2 relative in bf16


def _load(name):
    path = GOLDEN / name
    assert path.exists(), f"golden fixture {name} missing (tools/make_golden_*.py)"
    return np.load(path)


@pytest.mark.parametrize("case", gc.SCATTER_CASES, ids=[c[0] for c in gc.SCATTER_CASES])
def test_scatter_oracle_matches_reference_triton_kernel(case):
    name, n, Hkv, nb, bs, hole, seed = case
    g = _load(f"scatter_{name}.npz")
    k, v, k_cache, v_cache, slots = gc.scatter_inputs(n, Hkv, nb, bs, hole, seed)
    assert synth.checksum(k, v.contiguous(), k_cache, v_cache, slots) == str(g["input_sha"])
    assert n == 1 or v.stride(0) != v.shape[1] * v.shape[2]  # value really is a strided view
    orc.store_kvcache_ref(k, v, k_cache, v_cache, slots)
    assert np.array_equal(k_cache.view(torch.int16).numpy(), g["k_cache"])
    assert np.array_equal(v_cache.view(torch.int16).numpy(), g["v_cache"])


@pytest.mark.parametrize("case", gc.DECODE_CASES, ids=[c[0] for c in gc.DECODE_CASES])
def test_decode_oracle_matches_flash_attn(case):
    name, ctx, Hq, Hkv, seed = case
    g = _load(f"fa_decode_{name}.npz")
    c = synth.decode_case(len(ctx), ctx, Hq, Hkv, seed=seed)
    assert synth.checksum(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                          c.meta["block_tables"]) == str(g["input_sha"])
    want = torch.from_numpy(g["out"]).view(torch.bfloat16).float()
    got = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                               c.meta["block_tables"], c.scale)
    ok, ratio = orc.within_tolerance(got, want, ATOL, RTOL)
    assert ok, f"oracle vs flash-attn: violation ratio {ratio:.3f}"


@pytest.mark.parametrize("case", gc.PREFILL_CASES, ids=[c[0] for c in gc.PREFILL_CASES])
def test_prefill_oracle_matches_flash_attn(case):
    name, lens, cached, Hq, Hkv, seed = case
    g = _load(f"fa_prefill_{name}.npz")
    c = synth.prefill_case(lens, Hq, Hkv, cached=cached, seed=seed)
    m = c.meta
    assert synth.checksum(c.q, c.k, c.v.contiguous(), c.k_cache, c.v_cache, m["cu_seqlens_q"],
                          m["cu_seqlens_k"]) == str(g["input_sha"])
    ctx = type("Ctx", (), dict(m))()
    got = orc.attention_forward_ref(c.q, c.k, c.v, c.k_cache, c.v_cache, ctx, c.scale)
    want = torch.from_numpy(g["out"]).view(torch.bfloat16).float()
    ok, ratio = orc.within_tolerance(got, want, ATOL, RTOL)
    assert ok, f"oracle vs flash-attn: violation ratio {ratio:.3f}"


def test_oracle_edge_cases():
    # context_lens == 0 row -> zeros, and its (garbage) table row is never dereferenced
    c = synth.decode_case(3, [5, 300, 17], 4, 2, seed=3)
    lens = c.meta["context_lens"].clone()
    lens[1] = 0
    bt = c.meta["block_tables"].clone()
    bt[1] = 10 ** 6
    o = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, lens, bt, c.scale)
    assert torch.count_nonzero(o[1]) == 0 and torch.isfinite(o).all()
    # a single visible key -> output is exactly that value row (softmax of one element)
    c = synth.decode_case(1, [1], 4, 2, seed=4)
    o = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                             c.meta["block_tables"], c.scale)
    row = c.v_cache[c.meta["block_tables"][0, 0].item(), 0].float()
    assert torch.allclose(o[0], row.repeat_interleave(2, dim=0))
    # prefill of a sequence == decode of its last token over the same keys (bottom-right causal)
    p = synth.prefill_case([77], 4, 2, seed=5)
    ctx = type("Ctx", (), dict(p.meta))()
    full = orc.attention_forward_ref(p.q, p.k, p.v, p.k_cache, p.v_cache, ctx, p.scale)
    dec = orc.paged_decode_ref(p.q[-1:], p.k_cache, p.v_cache, torch.tensor([77], dtype=torch.int32),
                               p.meta["cu_seqlens_q"].new_tensor([p.seqs[0].block_table]),
                               p.scale)
    assert torch.allclose(full[-1], dec[0], atol=1e-5)
    # empty scatter / all-holes scatter leave the cache untouched
    k, v, kc, vc, slots = gc.scatter_inputs(5, 2, 2, 16, 0, 6)
    before = kc.clone()
    orc.store_kvcache_ref(k, v, kc, vc, torch.full_like(slots, -1))
    assert torch.equal(kc, before)


@pytest.mark.parametrize("case", gc.NORMROPE_CASES, ids=[c[0] for c in gc.NORMROPE_CASES])
def test_normrope_oracle_matches_reference_modules(case):
    """Bit-exact against the reference's RMSNorm / RotaryEmbedding (eager) called as
    models/qwen3.py:75-83 calls them (tools/make_golden_normrope.py)."""
    name, n, Hq, Hkv, has_norm, max_pos, base, seed = case
    g = _load(f"normrope_{name}.npz")
    qkv, positions, q_w, k_w = gc.normrope_inputs(n, Hq, Hkv, has_norm, max_pos, seed)
    assert synth.checksum(qkv, positions, *([q_w, k_w] if has_norm else [])) == str(g["input_sha"])
    cache = orc.rope_cos_sin_cache(128, max_pos, base)
    assert synth.checksum(cache) == str(g["cos_sin_sha"])
    q, k, v = orc.qk_norm_rope_ref(qkv, positions, q_w, k_w, gc.NORMROPE_EPS, cache, Hq, Hkv)
    assert np.array_equal(q.contiguous().view(torch.int16).numpy(), g["q"])
    assert np.array_equal(k.contiguous().view(torch.int16).numpy(), g["k"])
    # v is the untouched strided view of qkv (qwen3.py:79)
    assert v.data_ptr() == qkv.data_ptr() + (Hq + Hkv) * 128 * 2
