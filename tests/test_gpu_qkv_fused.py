"""-m gpu parity tests of the fused q/k RMSNorm + RoPE + KV-store path through
the C-ABI: golden fixtures made by the reference's own RMSNorm / RotaryEmbedding modules, seeded
comparisons with the oracle, equality of the one-launch decode with the two-call sequence, the
Context dispatch of qkv_attention, CUDA-graph replay and size-independent properties at full size.

Tolerance for the normalised + rotated rows: the arithmetic follows the reference's rounding
points, so results are bit-identical except where the fp32 sum of squares or the reciprocal
square root round differently (summation order; MUFU.RSQ on the GPU, as torch.rsqrt there, vs
IEEE 1/sqrt in the CPU oracle) and flip a bf16 rounding of the normalised value. Such a flip is one ulp of
xn, at most two ulps after the weight multiply and its rounding (|w| < 2), the rotation can shrink
the row maximum by sqrt(2), and the final rounding adds one more: every element must be within
FOUR bf16 ulps (4 * 2^-7) of the head row's largest element of the reference, and at most 0.5 %
of the elements may differ at all. Without
the norm (and for V rows) the result must be bit-exact. Attention outputs: north_star bound
(2e-2 abs + 1e-2 rel)."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN
from gpu_util import assert_close
from oracle import attention_oracle as orc
from tools import golden_cases as gc, synth

pytestmark = pytest.mark.gpu

D = 128
EPS = gc.NORMROPE_EPS


def _ulp_check(got: torch.Tensor, want: torch.Tensor, what: str, exact: bool = False):
    got, want = got.cpu().contiguous(), want.cpu().contiguous()
    assert got.shape == want.shape, what
    same = got.view(torch.int16) == want.view(torch.int16)
    if exact:
        assert bool(same.all()), f"{what}: not bit-exact ({(~same).sum().item()} elements differ)"
        return 0.0
    g, w = got.float(), want.float()
    # ulps of the head row's largest element: a flipped rounding of x1 moves y = x1 cos - x2 sin
    # by ulps of x1 even where y itself is small
    bound = w.abs().amax(dim=-1, keepdim=True) * (4 * 2.0 ** -7) + 1e-30
    assert bool(((g - w).abs() <= bound).all()), f"{what}: more than four bf16 ulps off"
    frac = 1.0 - same.float().mean().item()
    assert frac <= 5e-3, f"{what}: {frac:.4%} of the elements differ from the reference"
    return frac


def _normrope_dev(qkv, positions, q_w, k_w, cache, Hq, Hkv, dev, **kw):
    from byox_vllm_b200 import qkv_attention as fused

    to = lambda t: None if t is None else t.to(dev)
    qkv_d = to(qkv).clone()
    q, k, v = fused.qk_norm_rope_store(qkv_d, to(positions), to(q_w), to(k_w), EPS, to(cache), Hq,
                                       Hkv, D, **kw)
    torch.cuda.synchronize()
    return q, k, v, qkv_d


@pytest.mark.parametrize("case", gc.NORMROPE_CASES, ids=[c[0] for c in gc.NORMROPE_CASES])
def test_normrope_matches_reference_modules(cuda_dev, case):
    name, n, Hq, Hkv, has_norm, max_pos, base, seed = case
    g = np.load(GOLDEN / f"normrope_{name}.npz")
    qkv, positions, q_w, k_w = gc.normrope_inputs(n, Hq, Hkv, has_norm, max_pos, seed)
    cache = orc.rope_cos_sin_cache(D, max_pos, base)
    want_q = torch.from_numpy(g["q"]).view(torch.bfloat16)
    want_k = torch.from_numpy(g["k"]).view(torch.bfloat16)
    q, k, v, qkv_d = _normrope_dev(qkv, positions, q_w, k_w, cache.unsqueeze(1), Hq, Hkv, cuda_dev)
    _ulp_check(q, want_q, "q", exact=not has_norm)
    _ulp_check(k, want_k, "k (in place)", exact=not has_norm)
    # v and everything else in the row is untouched
    assert torch.equal(v.cpu(), qkv[:, (Hq + Hkv) * D:].view(n, Hkv, D))
    assert torch.equal(qkv_d[:, :Hq * D].cpu(), qkv[:, :Hq * D]), "q part of qkv must not change"


@pytest.mark.parametrize("n,Hq,Hkv,has_norm", [(1, 16, 8, True), (255, 16, 8, True),
                                               (4097, 32, 8, True), (333, 4, 1, True),
                                               (130, 8, 2, False), (2048, 8, 1, True),
                                               (1500, 8, 2, False), (1024, 16, 8, True)])
def test_normrope_store_vs_oracle(cuda_dev, n, Hq, Hkv, has_norm):
    """Seeded sizes the CPU oracle finishes in seconds; scatter with holes and out-of-range
    slots; write_k on/off; untouched cache rows stay untouched. n >= 1024 runs the bulk-staged
    (TMA) kernel, smaller n the direct one."""
    max_pos, bs, nb = 8192, 256, 24
    qkv, positions, q_w, k_w = gc.normrope_inputs(n, Hq, Hkv, has_norm, max_pos, seed=n)
    cache = orc.rope_cos_sin_cache(D, max_pos, 1e6)
    gen = torch.Generator().manual_seed(n)
    slots = torch.randperm(nb * bs, generator=gen)[:n].to(torch.int32)
    slots[::7] = -1
    if n > 3:
        slots[3] = nb * bs + 5  # out of range: skipped, never written out of bounds
    k_cache = torch.randn(nb, bs, Hkv, D, generator=gen).to(torch.bfloat16)
    v_cache = torch.randn(nb, bs, Hkv, D, generator=gen).to(torch.bfloat16)
    want_q, want_k, want_v = orc.qk_norm_rope_ref(qkv, positions, q_w, k_w, EPS, cache, Hq, Hkv)
    kc_ref, vc_ref = k_cache.clone(), v_cache.clone()
    ok = (slots >= 0) & (slots < nb * bs)
    orc.store_kvcache_ref(want_k[ok], want_v[ok], kc_ref, vc_ref, slots[ok])
    for write_k in (True, False):
        kc, vc = k_cache.to(cuda_dev), v_cache.to(cuda_dev)
        q, k, v, qkv_d = _normrope_dev(qkv, positions, q_w, k_w, cache, Hq, Hkv, cuda_dev,
                                       k_cache=kc, v_cache=vc, slot_mapping=slots.to(cuda_dev),
                                       write_k=write_k)
        _ulp_check(q, want_q, "q", exact=not has_norm)
        if write_k:
            _ulp_check(k, want_k, "k in place", exact=not has_norm)
        else:
            assert torch.equal(qkv_d.cpu(), qkv), "write_k=False must leave qkv untouched"
        _ulp_check(kc, kc_ref, "k_cache", exact=not has_norm)
        assert torch.equal(vc.cpu().view(torch.int16), vc_ref.view(torch.int16)), "v_cache"
        if write_k:  # the cache holds exactly the rows written in place
            live = torch.nonzero(ok).flatten()
            rows = kc.view(-1, Hkv, D)[slots[live].long().to(cuda_dev)]
            assert torch.equal(rows, k[live.to(cuda_dev)])


def test_normrope_identity_properties_full_size(cuda_dev):
    """BASELINE-size batch (16384 tokens, 32/8 heads): position 0 without norm is the identity
    (cos 1, sin 0) bit for bit; RoPE preserves the norm of every head row; with unit weights the
    normalised rows have RMS 1."""
    from byox_vllm_b200 import qkv_attention as fused

    n, Hq, Hkv, max_pos = 16384, 32, 8, 40960
    gen = torch.Generator(device=cuda_dev).manual_seed(5)
    qkv = torch.randn(n, (Hq + 2 * Hkv) * D, generator=gen, device=cuda_dev).to(torch.bfloat16)
    cache = orc.rope_cos_sin_cache(D, max_pos, 1e6).to(cuda_dev)
    zeros = torch.zeros(n, dtype=torch.int64, device=cuda_dev)
    work = qkv.clone()
    q, k, v = fused.qk_norm_rope_store(work, zeros, None, None, 0.0, cache, Hq, Hkv, D)
    assert torch.equal(q.view(n, -1), qkv[:, :Hq * D]) and torch.equal(work, qkv)
    pos = torch.randint(0, max_pos, (n,), generator=gen, device=cuda_dev)
    work = qkv.clone()
    q, k, _ = fused.qk_norm_rope_store(work, pos, None, None, 0.0, cache, Hq, Hkv, D)
    n_in = qkv[:, :Hq * D].view(n, Hq, D).float().norm(dim=-1)
    assert torch.allclose(q.float().norm(dim=-1), n_in, rtol=4e-3)
    ones = torch.ones(D, dtype=torch.bfloat16, device=cuda_dev)
    work = qkv.clone()
    q, k, _ = fused.qk_norm_rope_store(work, pos, ones, ones, EPS, cache, Hq, Hkv, D)
    rms = k.float().pow(2).mean(dim=-1).sqrt()
    assert torch.allclose(rms, torch.ones_like(rms), atol=6e-3)


def _fused_decode_case(batch, ctx, Hq, Hkv, seed, dev):
    c = synth.decode_case(batch, ctx, Hq, Hkv, seed=seed)
    gen = torch.Generator().manual_seed(seed + 1000)
    qkv = (1.5 * torch.randn(batch, (Hq + 2 * Hkv) * D, generator=gen)).to(torch.bfloat16)
    q_w = (1.0 + 0.25 * torch.randn(D, generator=gen)).to(torch.bfloat16)
    k_w = (1.0 + 0.25 * torch.randn(D, generator=gen)).to(torch.bfloat16)
    max_pos = max(4096, max(c.meta["context_lens"].tolist()) + 1)
    cache = orc.rope_cos_sin_cache(D, max_pos, 1e6)
    return c, qkv, q_w, k_w, cache


@pytest.mark.parametrize("batch,ctx,Hq,Hkv,splits", [
    (4, 512, 16, 8, None), (5, [1, 256, 257, 700, 64], 16, 8, None), (3, 2000, 32, 8, 4),
    (2, [4096, 130], 8, 1, 3), (7, 300, 4, 1, None), (64, 1024, 16, 4, None)])
def test_qkv_decode_one_launch(cuda_dev, batch, ctx, Hq, Hkv, splits):
    """b200attn_qkv_decode vs the oracle (qwen3.py:75-84 restated) and vs the two-call sequence:
    the cache must end up bit-identical (same K arithmetic in both kernels)."""
    from byox_vllm_b200 import _lib, qkv_attention as fused
    from byox_vllm_b200.attention import paged_decode

    c, qkv, q_w, k_w, cache = _fused_decode_case(batch, ctx, Hq, Hkv, 21, cuda_dev)
    m = c.meta
    ctxobj = type("Ctx", (), dict(m))()
    kc_ref, vc_ref = c.k_cache.clone(), c.v_cache.clone()
    want = orc.qkv_attention_ref(qkv, m["positions"], q_w, k_w, EPS, cache, kc_ref, vc_ref,
                                 ctxobj, c.scale, Hq, Hkv)
    dev = cuda_dev
    args = [t.to(dev) for t in (qkv, m["positions"], q_w, k_w)] + [EPS, cache.to(dev)]
    meta = [m[k].to(dev) for k in ("context_lens", "block_tables", "slot_mapping")]
    kc1, vc1 = c.k_cache.to(dev), c.v_cache.to(dev)
    n0 = _lib.load().b200attn_launch_count()
    got = fused.qkv_decode(*args, kc1, vc1, *meta, c.scale, Hq, Hkv, num_splits=splits)
    torch.cuda.synchronize()
    n_launch = _lib.load().b200attn_launch_count() - n0
    assert 1 <= n_launch <= 2, f"{n_launch} launches for one fused decode step"
    assert_close(got.unsqueeze(1), want, "qkv_decode vs oracle")
    _ulp_check(kc1, kc_ref, "k_cache vs oracle")
    assert torch.equal(vc1.cpu().view(torch.int16), vc_ref.view(torch.int16))
    # two-call sequence
    kc2, vc2 = c.k_cache.to(dev), c.v_cache.to(dev)
    q2, _, _ = fused.qk_norm_rope_store(args[0].clone(), *args[1:], Hq, Hkv, D, k_cache=kc2,
                                        v_cache=vc2, slot_mapping=meta[2], write_k=False)
    got2 = paged_decode(q2, kc2, vc2, meta[0], meta[1], c.scale, num_splits=splits)
    torch.cuda.synchronize()
    assert torch.equal(kc1, kc2) and torch.equal(vc1, vc2), "fused and two-call caches differ"
    assert_close(got, got2.float(), "one launch vs two calls")


def test_qkv_attention_dispatch(cuda_dev):
    """qkv_attention under a Context: dense prefill, paged-prefix prefill, decode — against
    qwen3.py:75-84 restated on the CPU; then the decode step that follows the prefill reads what
    the prefill stored."""
    from byox_vllm_b200 import qkv_attention as fused
    from tools import metadata_ref as md
    from byox_vllm_b200.attention import Attention
    from byox_vllm_b200.context import reset_context, set_context

    Hq, Hkv = 16, 8
    dev = cuda_dev
    for lens, cached, seed in (([300, 512, 40, 257], None, 31), ([300, 700, 513], [256, 0, 512], 32)):
        c = synth.prefill_case(lens, Hq, Hkv, cached=cached, seed=seed)
        m = c.meta
        n = m["positions"].numel()
        gen = torch.Generator().manual_seed(seed)
        qkv = (1.5 * torch.randn(n, (Hq + 2 * Hkv) * D, generator=gen)).to(torch.bfloat16)
        q_w = (1.0 + 0.25 * torch.randn(D, generator=gen)).to(torch.bfloat16)
        k_w = (1.0 + 0.25 * torch.randn(D, generator=gen)).to(torch.bfloat16)
        cache = orc.rope_cos_sin_cache(D, 4096, 1e6)
        kc_ref, vc_ref = c.k_cache.clone(), c.v_cache.clone()
        want = orc.qkv_attention_ref(qkv, m["positions"], q_w, k_w, EPS, cache, kc_ref, vc_ref,
                                     type("Ctx", (), dict(m))(), c.scale, Hq, Hkv)
        attn = Attention(Hq, D, c.scale, Hkv)
        attn.k_cache, attn.v_cache = c.k_cache.to(dev), c.v_cache.to(dev)
        set_context(**md.context_kwargs(m, dev))
        try:
            got = fused.qkv_attention(attn, qkv.to(dev), m["positions"].to(dev), q_w.to(dev),
                                      k_w.to(dev), EPS, cache.unsqueeze(1).to(dev))
            torch.cuda.synchronize()
        finally:
            reset_context()
        assert_close(got, want, f"prefill cached={cached}")
        _ulp_check(attn.k_cache, kc_ref, "k_cache after prefill")
        assert torch.equal(attn.v_cache.cpu().view(torch.int16), vc_ref.view(torch.int16))
    # warm-up shape: no cache allocated yet (model_runner.py:104-117)
    c = synth.prefill_case([64, 64], Hq, Hkv, seed=33)
    attn = Attention(Hq, D, c.scale, Hkv)
    m = dict(c.meta)
    m["slot_mapping"] = torch.empty(0, dtype=torch.int32)
    set_context(**md.context_kwargs(m, dev))
    try:
        qkv = torch.randn(128, (Hq + 2 * Hkv) * D).to(torch.bfloat16)
        got = fused.qkv_attention(attn, qkv.to(dev), m["positions"].to(dev), None, None, 0.0,
                                  orc.rope_cos_sin_cache(D, 4096, 1e4).to(dev))
        torch.cuda.synchronize()
    finally:
        reset_context()
    want = orc.qkv_attention_ref(qkv, m["positions"], None, None, 0.0,
                                 orc.rope_cos_sin_cache(D, 4096, 1e4), torch.tensor([]),
                                 torch.tensor([]), type("Ctx", (), m)(), c.scale, Hq, Hkv)
    assert_close(got, want, "warm-up prefill without cache")


def test_qkv_decode_graph_replay(cuda_dev):
    """Captured once with padding rows (context_lens 0, slot -1), replayed with new metadata and
    new qkv in the same buffers, as ModelRunner does (model_runner.py:242-254)."""
    from byox_vllm_b200 import qkv_attention as fused

    Hq, Hkv, B = 16, 8, 8
    dev = cuda_dev
    from tools import metadata_ref as md

    first = _fused_decode_case(6, [300, 17, 1000, 256, 257, 511], Hq, Hkv, 41, dev)
    c0 = first[0]
    # next step of the same sequences: one more token each, a fresh block where a page fills up
    used = {b for s in c0.seqs for b in s.block_table}
    spare = [b for b in range(c0.k_cache.size(0)) if b not in used]
    seqs2 = []
    for s in c0.seqs:
        table = list(s.block_table)
        if s.num_tokens % 256 == 0:
            table.append(spare.pop())
        seqs2.append(md.SeqState(num_tokens=s.num_tokens + 1, block_table=table))
    c1 = synth.DecodeCase(c0.q, c0.k, c0.v, c0.k_cache, c0.v_cache,
                          md.prepare_decode(seqs2, 256), seqs2, c0.scale)
    qkv2 = (1.5 * torch.randn(6, (Hq + 2 * Hkv) * D,
                              generator=torch.Generator().manual_seed(77))).to(torch.bfloat16)
    cases = [first, (c1, qkv2) + tuple(first[2:])]
    W = 8
    g_qkv = torch.zeros(B, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16, device=dev)
    g_pos = torch.zeros(B, dtype=torch.int64, device=dev)
    g_slot = torch.full((B,), -1, dtype=torch.int32, device=dev)
    g_lens = torch.zeros(B, dtype=torch.int32, device=dev)
    g_bt = torch.zeros(B, W, dtype=torch.int32, device=dev)
    g_out = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device=dev)
    kc, vc = c0.k_cache.to(dev), c0.v_cache.to(dev)
    q_w, k_w, cache = (t.to(dev) for t in cases[0][2:5])

    def run():
        fused.qkv_decode(g_qkv, g_pos, q_w, k_w, EPS, cache, kc, vc, g_lens, g_bt, g_slot,
                         c0.scale, Hq, Hkv, out=g_out)

    run()  # eager warm-up with all-padding rows
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        run()
    kc_ref, vc_ref = c0.k_cache.clone(), c0.v_cache.clone()
    for c, qkv, _, _, _ in cases:
        m = c.meta
        n = qkv.size(0)
        g_qkv[:n] = qkv.to(dev)
        g_pos[:n] = m["positions"].to(dev)
        g_slot.fill_(-1)
        g_slot[:n] = m["slot_mapping"].to(dev)
        g_lens.zero_()
        g_lens[:n] = m["context_lens"].to(dev)
        g_bt[:n, :m["block_tables"].size(1)] = m["block_tables"].to(dev)
        graph.replay()
        torch.cuda.synchronize()
        want = orc.qkv_attention_ref(qkv, m["positions"], cases[0][2], cases[0][3], EPS,
                                     cases[0][4], kc_ref, vc_ref, type("Ctx", (), dict(m))(),
                                     c.scale, Hq, Hkv)
        assert_close(g_out[:n].unsqueeze(1), want, "graph replay")
        assert torch.count_nonzero(g_out[n:]) == 0, "padding rows must give zeros"
        _ulp_check(kc, kc_ref, "k_cache under replay")


def test_qkv_attention_decode_with_cascade_armed(cuda_dev):
    """A decode batch the size heuristic arms the shared-prefix cascade for (64 rows, tables of
    32 pages): qkv_attention must run qk_norm_rope_store (q materialised for the prefix pass's
    TMA map, K/V scattered) followed by the cascade decode, and still equal qwen3.py:75-84."""
    from byox_vllm_b200 import _lib, qkv_attention as fused
    from byox_vllm_b200.attention import Attention, cascade_chunks
    from byox_vllm_b200.context import reset_context, set_context

    dev = cuda_dev
    Hq, Hkv, B, W = 8, 2, 64, 32
    assert cascade_chunks(B, Hq, Hkv, W, 256) > 0
    c = synth.shared_prefix_decode_case([1300 + 5 * i for i in range(B)], 5, Hq, Hkv, seed=71)
    m = c.meta
    gen = torch.Generator().manual_seed(72)
    qkv = (1.5 * torch.randn(B, (Hq + 2 * Hkv) * D, generator=gen)).to(torch.bfloat16)
    q_w = (1.0 + 0.25 * torch.randn(D, generator=gen)).to(torch.bfloat16)
    k_w = (1.0 + 0.25 * torch.randn(D, generator=gen)).to(torch.bfloat16)
    cache = orc.rope_cos_sin_cache(D, 4096, 1e6)
    kc_ref, vc_ref = c.k_cache.clone(), c.v_cache.clone()
    want = orc.qkv_attention_ref(qkv, m["positions"], q_w, k_w, EPS, cache, kc_ref, vc_ref,
                                 type("Ctx", (), dict(m))(), c.scale, Hq, Hkv)
    bt = torch.full((B, W), -1, dtype=torch.int32)
    bt[:, :m["block_tables"].size(1)] = m["block_tables"]
    attn = Attention(Hq, D, c.scale, Hkv)
    attn.k_cache, attn.v_cache = c.k_cache.to(dev), c.v_cache.to(dev)
    set_context(False, slot_mapping=m["slot_mapping"].to(dev),
                context_lens=m["context_lens"].to(dev), block_tables=bt.to(dev))
    try:
        n0 = _lib.load().b200attn_launch_count()
        got = fused.qkv_attention(attn, qkv.to(dev), m["positions"].to(dev), q_w.to(dev),
                                  k_w.to(dev), EPS, cache.to(dev))
        torch.cuda.synchronize()
        launches = _lib.load().b200attn_launch_count() - n0
    finally:
        reset_context()
    assert launches == 4, f"plan + norm/rope/store + prefix pass + decode, got {launches}"
    assert_close(got.unsqueeze(1), want, "qkv_attention with the cascade armed")
    _ulp_check(attn.k_cache, kc_ref, "k_cache")
    assert torch.equal(attn.v_cache.cpu().view(torch.int16), vc_ref.view(torch.int16))
