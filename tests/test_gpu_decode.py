"""-m gpu: paged decode attention through the C-ABI vs the fp32 oracle, the flash-attn golden
vectors, and size-independent properties at BASELINE sizes."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN
from gpu_util import assert_close, to_dev
from oracle import attention_oracle as orc
from tools import golden_cases as gc, synth

pytestmark = pytest.mark.gpu


def _decode(d, scale, **kw):
    from byox_vllm_b200.attention import paged_decode

    m = d["meta"]
    o = paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"], m["block_tables"],
                     scale, **kw)
    torch.cuda.synchronize()
    return o


@pytest.mark.parametrize("case", gc.DECODE_CASES, ids=[c[0] for c in gc.DECODE_CASES])
@pytest.mark.parametrize("splits", [None, 1, 2, 5])
def test_decode_matches_oracle_and_flash_attn(cuda_dev, case, splits):
    name, ctx, Hq, Hkv, seed = case
    c = synth.decode_case(len(ctx), ctx, Hq, Hkv, seed=seed)
    d = to_dev(c, cuda_dev)
    got = _decode(d, c.scale, num_splits=splits)
    assert got.dtype == torch.bfloat16 and got.shape == c.q.shape
    want = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                                c.meta["block_tables"], c.scale)
    assert_close(got, want, f"{name} vs oracle")
    g = np.load(GOLDEN / f"fa_decode_{name}.npz")
    fa = torch.from_numpy(g["out"]).view(torch.bfloat16)
    assert_close(got, fa, f"{name} vs flash-attn golden")


def test_decode_padding_rows_and_stale_tables(cuda_dev):
    """CUDA-graph replay semantics (model_runner.py:242-254): rows with context_lens == 0 give
    zeros and read no table entry; entries beyond ceil(ctx/256) are -1 / garbage and never read;
    the table may be wider than needed (row stride = ceil(max_model_len/256))."""
    c = synth.decode_case(6, [300, 1, 700, 256, 513, 90], 16, 8, seed=31)
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    lens = m["context_lens"].clone()
    lens[1] = 0
    lens[4] = 0
    W = 16
    wide = torch.full((6, W), 2 ** 30, dtype=torch.int32, device=cuda_dev)  # poison
    bt = c.meta["block_tables"]
    for b, s in enumerate(c.seqs):
        n = len(s.block_table)
        wide[b, :n] = bt[b, :n].to(cuda_dev)
    wide[1] = -1
    wide[4] = 2 ** 30
    from byox_vllm_b200.attention import paged_decode

    for splits in (1, 3, None):
        got = paged_decode(d["q"], d["k_cache"], d["v_cache"], lens, wide[:, :W], c.scale,
                           num_splits=splits)
        torch.cuda.synchronize()
        want = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, lens.cpu(), wide.cpu(), c.scale)
        assert torch.count_nonzero(got[1]) == 0 and torch.count_nonzero(got[4]) == 0
        assert_close(got, want, f"padding rows, splits={splits}")
    # a sliced (non-contiguous rows) table view, like graph_vars["block_tables"][:bs]
    big = torch.full((8, 32), -1, dtype=torch.int32, device=cuda_dev)
    big[:6, :W] = wide
    got = paged_decode(d["q"], d["k_cache"], d["v_cache"], lens, big[:6], c.scale)
    torch.cuda.synchronize()
    assert_close(got, want, "wide table view")


def test_decode_ignores_garbage_beyond_context(cuda_dev):
    """Slots past ctx inside the last page may hold anything (NaN/Inf): never leaks into out."""
    c = synth.decode_case(3, [10, 257, 600], 8, 2, seed=32)
    for b, s in enumerate(c.seqs):
        last = s.block_table[-1]
        used = s.num_tokens - (len(s.block_table) - 1) * 256
        c.k_cache[last, used:] = float("nan")
        c.v_cache[last, used:] = float("inf")
    d = to_dev(c, cuda_dev)
    want = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                                c.meta["block_tables"], c.scale)
    for splits in (1, 2):
        got = _decode(d, c.scale, num_splits=splits)
        assert torch.isfinite(got.float()).all()
        assert_close(got, want, "garbage beyond ctx")


def test_fused_store_decode_equals_store_then_decode(cuda_dev):
    from byox_vllm_b200.attention import paged_decode, store_kvcache

    c = synth.decode_case(5, [1, 256, 257, 1000, 77], 16, 8, seed=33)
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    kc2, vc2 = d["k_cache"].clone(), d["v_cache"].clone()
    o1 = paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"], m["block_tables"],
                      c.scale, key=d["k"], value=d["v"], slot_mapping=m["slot_mapping"])
    store_kvcache(d["k"], d["v"], kc2, vc2, m["slot_mapping"])
    o2 = paged_decode(d["q"], kc2, vc2, m["context_lens"], m["block_tables"], c.scale)
    torch.cuda.synchronize()
    assert torch.equal(o1.view(torch.int16), o2.view(torch.int16))
    assert torch.equal(kc2.view(torch.int16), d["k_cache"].view(torch.int16))
    # and against the oracle's forward (scatter + decode)
    ctx = type("Ctx", (), dict(c.meta))()
    want = orc.attention_forward_ref(c.q, c.k, c.v, c.k_cache, c.v_cache, ctx, c.scale)
    assert_close(o1, want.squeeze(1), "fused store+decode vs oracle forward")
    assert torch.equal(d["k_cache"].cpu().view(torch.int16), c.k_cache.view(torch.int16))
    assert torch.equal(d["v_cache"].cpu().view(torch.int16), c.v_cache.view(torch.int16))


def test_decode_under_cuda_graph_capture_and_replay(cuda_dev):
    """Mimics ModelRunner._capture_cudagraph / _run_model (model_runner.py:242-297): capture with
    all-zero metadata buffers at a rounded-up batch size, then replay with live metadata copied
    into the static buffers (slot_mapping padded with -1, context_lens with 0, stale tables)."""
    from byox_vllm_b200.attention import Attention
    from byox_vllm_b200.context import reset_context, set_context

    dev = cuda_dev
    Hq, Hkv, D, bs_graph, W = 16, 8, 128, 8, 12
    live = synth.decode_case(5, [300, 1, 700, 256, 513], Hq, Hkv, seed=34)
    d = to_dev(live, dev)
    attn = Attention(Hq, D, live.scale, Hkv)
    attn.k_cache, attn.v_cache = d["k_cache"], d["v_cache"]
    g_q = torch.zeros(bs_graph, Hq, D, dtype=torch.bfloat16, device=dev)
    g_k = torch.zeros(bs_graph, Hkv, D, dtype=torch.bfloat16, device=dev)
    g_v = torch.zeros(bs_graph, Hkv, D, dtype=torch.bfloat16, device=dev)
    g_slots = torch.zeros(bs_graph, dtype=torch.int32, device=dev)
    g_lens = torch.zeros(bs_graph, dtype=torch.int32, device=dev)
    g_bt = torch.zeros(bs_graph, W, dtype=torch.int32, device=dev)
    g_out = torch.zeros(bs_graph, 1, Hq, D, dtype=torch.bfloat16, device=dev)
    slot0_before = (d["k_cache"].view(-1, Hkv * D)[0].clone(),
                    d["v_cache"].view(-1, Hkv * D)[0].clone())
    set_context(False, slot_mapping=g_slots, context_lens=g_lens, block_tables=g_bt)
    g_out.copy_(attn(g_q, g_k, g_v))  # warm-up (writes the dummy row to slot 0, like the reference)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        g_out.copy_(attn(g_q, g_k, g_v))
    torch.cuda.synchronize()
    reset_context()
    d["k_cache"].view(-1, Hkv * D)[0].copy_(slot0_before[0])  # undo the capture-time dummy write
    d["v_cache"].view(-1, Hkv * D)[0].copy_(slot0_before[1])
    # two replays with different live batches through the same graph
    for n_live in (5, 3):
        m = d["meta"]
        g_q[:n_live] = d["q"][:n_live]
        g_k[:n_live] = d["k"][:n_live]
        g_v[:n_live] = d["v"][:n_live]
        g_slots.fill_(-1)
        g_slots[:n_live] = m["slot_mapping"][:n_live]
        g_lens.zero_()
        g_lens[:n_live] = m["context_lens"][:n_live]
        g_bt[:n_live, : m["block_tables"].size(1)] = m["block_tables"][:n_live]
        graph.replay()
        torch.cuda.synchronize()
        kc, vc = live.k_cache.clone(), live.v_cache.clone()
        ctx = type("Ctx", (), dict(live.meta))()
        want = orc.attention_forward_ref(live.q, live.k, live.v, kc, vc, ctx, live.scale)
        assert_close(g_out[:n_live], want[:n_live], f"graph replay n_live={n_live}")
        assert torch.count_nonzero(g_out[n_live:]) == 0
        rows = live.meta["slot_mapping"][:n_live].long()
        assert torch.equal(d["k_cache"].view(-1, Hkv * D)[rows.to(dev)].cpu().view(torch.int16),
                           kc.view(-1, Hkv * D)[rows].view(torch.int16))


@pytest.mark.parametrize("shape", [(64, 4096, 32, 8), (128, 4096, 4, 1), (4, 32768, 32, 8)],
                         ids=["b64_ctx4k_32_8", "b128_ctx4k_4_1", "b4_ctx32k_32_8"])
def test_decode_properties_at_baseline_sizes(cuda_dev, shape):
    """Sizes the oracle cannot finish in seconds: (a) constant V -> output equals the constant
    (softmax weights sum to 1); (b) output is invariant to the physical page permutation;
    (c) split count does not change the result beyond tolerance; (d) a sampled subset of rows
    matches the oracle."""
    from byox_vllm_b200.attention import paged_decode

    B, ctx, Hq, Hkv = shape
    dev = cuda_dev
    c = synth.decode_case(B, ctx, Hq, Hkv, seed=40, device=dev)
    m = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in c.meta.items()}
    o = paged_decode(c.q, c.k_cache, c.v_cache, m["context_lens"], m["block_tables"], c.scale)
    # (c)
    o1 = paged_decode(c.q, c.k_cache, c.v_cache, m["context_lens"], m["block_tables"], c.scale,
                      num_splits=1)
    o7 = paged_decode(c.q, c.k_cache, c.v_cache, m["context_lens"], m["block_tables"], c.scale,
                      num_splits=7)
    assert_close(o, o1, "auto splits vs 1")
    assert_close(o7, o1, "7 splits vs 1")
    # (b) move every page to a new physical block
    nb = c.k_cache.shape[0]
    perm = torch.randperm(nb, device=dev)
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(nb, device=dev)
    kc2, vc2 = c.k_cache[perm], c.v_cache[perm]  # new block i holds old block perm[i]
    bt2 = inv[m["block_tables"].clamp_min(0).long()].to(torch.int32)
    o_perm = paged_decode(c.q, kc2, vc2, m["context_lens"], bt2, c.scale, num_splits=1)
    assert torch.equal(o_perm.view(torch.int16), o1.view(torch.int16))
    del kc2, vc2
    # (a)
    vconst = torch.full_like(c.v_cache, 0.75)
    oc = paged_decode(c.q, c.k_cache, vconst, m["context_lens"], m["block_tables"], c.scale)
    assert (oc.float() - 0.75).abs().max().item() <= 1e-2
    del vconst
    # (d) rows 0 and B-1 against the oracle
    for b in (0, B - 1):
        want = orc.paged_decode_ref(c.q[b:b + 1].cpu(), c.k_cache.cpu(), c.v_cache.cpu(),
                                    c.meta["context_lens"][b:b + 1],
                                    c.meta["block_tables"][b:b + 1], c.scale)
        assert_close(o[b:b + 1], want, f"row {b} vs oracle")


@pytest.mark.parametrize("batch,ctx,Hq,Hkv,splits", [
    (1, 4096, 32, 8, 8), (3, [5000, 1, 300], 16, 8, 5), (2, [700, 64], 8, 1, 64),
    (7, 1000, 4, 1, 3), (16, 2048, 16, 4, 2), (1, 20000, 8, 1, 296), (40, 333, 16, 8, 296)])
def test_in_launch_merge_of_cut_rows(cuda_dev, batch, ctx, Hq, Hkv, splits):
    """Rows cut across CTA ranges are finished by the piece that arrives last: ONE launch, the
    arrival counters come back to zero every time, the result does not depend on which piece
    arrived last (bit-identical across repeats) and agrees with the single-CTA walk (num_splits
    = 1, no cut rows at all) and the oracle. Padding rows (ctx 0) stay zero."""
    from byox_vllm_b200 import _lib, attention as A

    lib = _lib.load()
    c = synth.decode_case(batch, ctx, Hq, Hkv, seed=61)
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    lens = m["context_lens"].clone()
    if batch >= 3:
        lens[1] = 0  # padding row of a captured batch
    args = (d["q"], d["k_cache"], d["v_cache"], lens, m["block_tables"], c.scale)
    A.ensure_split_counters(cuda_dev)
    counters = A._SPLIT_COUNTERS[cuda_dev.index or 0][-1]
    one = A.paged_decode(*args, num_splits=1)
    torch.cuda.synchronize()
    first = None
    for _ in range(3):
        n0 = lib.b200attn_launch_count()
        got = A.paged_decode(*args, num_splits=splits)
        torch.cuda.synchronize()
        assert lib.b200attn_launch_count() - n0 == 1
        assert int(counters.abs().sum()) == 0
        first = got if first is None else first
        assert torch.equal(got, first)
    assert_close(got, one, "cut rows vs single-CTA walk")
    if batch >= 3:
        assert torch.count_nonzero(got[1]) == 0
    ref = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, lens.cpu(), c.meta["block_tables"],
                               c.scale)
    assert_close(got, ref, "in-launch merge vs oracle")


def test_cut_rows_need_registered_counters(cuda_dev):
    """num_splits > 1 without the arrival counters is an error, not a silent second path."""
    from byox_vllm_b200 import _lib, attention as A

    lib = _lib.load()
    c = synth.decode_case(2, [700, 64], 8, 1, seed=62)
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    A.ensure_split_counters(cuda_dev)
    counters = A._SPLIT_COUNTERS[cuda_dev.index or 0][-1]
    try:
        assert lib.b200attn_set_split_counters(None, 0) == 0
        with pytest.raises(_lib.B200AttnError) as e:
            A.paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"],
                           m["block_tables"], c.scale, num_splits=4)
        assert e.value.code == _lib.E_WORKSPACE and "counters" in e.value.msg
        got = A.paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"],
                             m["block_tables"], c.scale, num_splits=1)  # needs none
        torch.cuda.synchronize()
    finally:
        assert lib.b200attn_set_split_counters(counters.data_ptr(), counters.numel()) == 0
    want = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                                c.meta["block_tables"], c.scale)
    assert_close(got, want, "single CTA without counters")


@pytest.mark.parametrize("shape", [(1, 131072, 32, 8), (512, 1024, 32, 8), (1, 131072, 4, 1),
                                   (1500, 300, 8, 2)],
                         ids=["b1_ctx128k_32_8", "b512_ctx1k_32_8", "b1_ctx128k_4_1_tp8",
                              "b1500_eager_chunks"])
def test_decode_config5_corners(cuda_dev, shape):
    """BASELINE config 5's corners (B 1 x 131072: one row per KV head cut across up to 296 CTAs;
    B 512 x 1024: many whole rows per CTA) and a batch above the 1024-row launch limit (eager
    decode of > 512 sequences, model_runner.py:230): sampled rows vs the oracle, constant V,
    agreement between the automatic partition and other CTA counts."""
    from byox_vllm_b200.attention import paged_decode

    B, ctx, Hq, Hkv = shape
    dev = cuda_dev
    c = synth.decode_case(B, ctx, Hq, Hkv, seed=44, device=dev)
    m = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in c.meta.items()}
    args = (c.q, c.k_cache, c.v_cache, m["context_lens"], m["block_tables"], c.scale)
    o = paged_decode(*args)
    o3 = paged_decode(*args, num_splits=3 if B > 1 else 37)
    assert_close(o, o3, "auto partition vs another CTA count")
    vconst = torch.full_like(c.v_cache, -1.25)
    oc = paged_decode(c.q, c.k_cache, vconst, m["context_lens"], m["block_tables"], c.scale)
    assert (oc.float() + 1.25).abs().max().item() <= 1e-2
    del vconst
    k_cpu, v_cpu = c.k_cache.cpu(), c.v_cache.cpu()
    for b in sorted({0, B // 2, B - 1}):
        want = orc.paged_decode_ref(c.q[b:b + 1].cpu(), k_cpu, v_cpu,
                                    c.meta["context_lens"][b:b + 1],
                                    c.meta["block_tables"][b:b + 1], c.scale)
        assert_close(o[b:b + 1], want, f"row {b} vs oracle")


def test_fused_store_with_foreign_slot_mapping(cuda_dev):
    """include/b200attn.h documents the fused store's precondition (slot_mapping[b] = slot of token
    ctx_b - 1). A caller that breaks it still gets every row stored bit-exactly at the slot it
    named; when that slot is attended by no row of the batch (here: free blocks) the attention
    output is exactly the decode over the cache without this step's rows."""
    from byox_vllm_b200.attention import paged_decode

    c = synth.decode_case(4, [300, 256, 700, 65], 16, 8, seed=63)
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    nb = d["k_cache"].shape[0]
    used = {b for s in c.seqs for b in s.block_table}
    free = [b for b in range(nb) if b not in used]
    assert len(free) >= 2
    foreign = torch.tensor([free[0] * 256 + 3, free[0] * 256 + 4, -1, free[1] * 256 + 255],
                           dtype=torch.int32, device=cuda_dev)
    kc0, vc0 = d["k_cache"].clone(), d["v_cache"].clone()
    plain = paged_decode(d["q"], kc0, vc0, m["context_lens"], m["block_tables"], c.scale)
    got = paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"], m["block_tables"],
                       c.scale, key=d["k"], value=d["v"], slot_mapping=foreign)
    torch.cuda.synchronize()
    assert torch.equal(got.view(torch.int16), plain.view(torch.int16))
    kc_ref, vc_ref = c.k_cache.clone(), c.v_cache.clone()
    orc.store_kvcache_ref(c.k, c.v, kc_ref, vc_ref, foreign.cpu())
    assert torch.equal(d["k_cache"].cpu().view(torch.int16), kc_ref.view(torch.int16))
    assert torch.equal(d["v_cache"].cpu().view(torch.int16), vc_ref.view(torch.int16))


@pytest.mark.parametrize("shape", [(24, [300, 1, 700, 256, 513, 90] * 4, 16, 8),      # whole rows
                                   (2, [9000, 5000], 32, 8),                            # cut rows
                                   (40, None, 4, 1)], ids=["rows", "cut", "ragged"])
def test_overlap_mode_back_to_back(cuda_dev, shape):
    """Overlap mode (b200attn_set_decode_overlap): every decode launch is a programmatic dependent
    of the kernel before it and starts its prologue while that kernel still runs. The data it
    reads after griddepcontrol.wait - q here, rewritten by a copy kernel right before each launch -
    must be the new data, and consecutive launches share counters and workspace. Ten launches
    back to back, eagerly and from one CUDA graph, against the same calls with overlap off."""
    from byox_vllm_b200.attention import paged_decode, set_decode_overlap

    B, ctx, Hq, Hkv = shape
    if ctx is None:
        rng = np.random.default_rng(5)
        ctx = rng.integers(1, 3000, size=B).tolist()
    c = synth.decode_case(B, ctx, Hq, Hkv, seed=77)
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    g = torch.Generator(device=cuda_dev).manual_seed(3)
    qs = [torch.empty_like(d["q"]).normal_(generator=g) for _ in range(10)]
    q = torch.empty_like(d["q"])

    def run_all(outs):
        for i in range(10):
            q.copy_(qs[i])  # a kernel that writes q immediately before the dependent launch
            paged_decode(q, d["k_cache"], d["v_cache"], m["context_lens"], m["block_tables"],
                         c.scale, out=outs[i])

    want = [torch.empty_like(q) for _ in range(10)]
    prev = set_decode_overlap(False)
    try:
        run_all(want)
        torch.cuda.synchronize()
        assert_close(want[3], orc.paged_decode_ref(qs[3].cpu(), c.k_cache, c.v_cache,
                                                   c.meta["context_lens"], c.meta["block_tables"],
                                                   c.scale), "overlap off vs oracle")
        set_decode_overlap(True)
        got = [torch.zeros_like(q) for _ in range(10)]
        run_all(got)
        torch.cuda.synchronize()
        for i in range(10):
            assert torch.equal(got[i], want[i]), f"eager launch {i} differs with overlap on"
        got = [torch.zeros_like(q) for _ in range(10)]
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            run_all(got)
        for _ in range(3):
            graph.replay()
        torch.cuda.synchronize()
        for i in range(10):
            assert torch.equal(got[i], want[i]), f"graph launch {i} differs with overlap on"
    finally:
        set_decode_overlap(prev)


def test_split_counters_grow_with_the_batch(cuda_dev, monkeypatch):
    """The registered counter buffer follows the call: rows x KV heads + 2 beyond the current
    buffer registers a larger one (an eager call with an explicit num_splits > 1 on more
    (sequence, KV head) pairs than the default 65536 used to fail with B200ATTN_E_WORKSPACE).
    Checked at a small scale by starting from a 64-slot default."""
    from byox_vllm_b200 import _lib
    from byox_vllm_b200 import attention as A

    idx = cuda_dev.index or 0
    A.ensure_split_counters(cuda_dev)
    kept = A._SPLIT_COUNTERS.pop(idx)  # stays alive: earlier graphs may hold it
    monkeypatch.setattr(A, "_SPLIT_COUNTER_SLOTS", 64)
    try:
        sizes = []
        for B, seed in ((4, 41), (24, 42), (100, 43), (24, 44)):
            c = synth.decode_case(B, [300 + 37 * (i % 9) for i in range(B)], 16, 8, seed=seed)
            d = to_dev(c, cuda_dev)
            got = _decode(d, c.scale, num_splits=2)
            want = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                                        c.meta["block_tables"], c.scale)
            assert_close(got, want, f"B={B} after counter registration")
            sizes.append(A._SPLIT_COUNTERS[idx][-1].numel())
            assert sizes[-1] >= B * 8 + 2
            assert all(int(b.abs().max()) == 0 for b in A._SPLIT_COUNTERS[idx])
        assert sizes == [64, 256, 1024, 1024]
    finally:
        A._SPLIT_COUNTERS[idx] = kept
        rc = _lib.load().b200attn_set_split_counters(kept[-1].data_ptr(), kept[-1].numel())
        _lib.check("b200attn_set_split_counters", rc)


@pytest.mark.parametrize("fused", [False, True], ids=["decode", "store_decode"])
@pytest.mark.parametrize("shape", [(32, 16, 8), (48, 4, 1), (3, 32, 8)],
                         ids=["graph32", "ragged48", "cut3"])
def test_overlap_mode_adjacent_launches_with_padding_rows(cuda_dev, shape, fused):
    """The chain bench.py's layer loop and the 1024-row chunk loop build: decode launches with
    NOTHING between them, so each one really is a programmatic dependent of the previous decode
    (in test_overlap_mode_back_to_back the q copy sits between them and serialises the pair). The
    batch is a CUDA-graph batch that is not full: rows with context_lens == 0 in front, in the
    middle and at the end, so some CTAs start on an empty row and ask for more work in their
    prologue. Ten launches on ten pre-staged q (and, fused, ten new K/V rows for the same slots),
    eagerly and from one graph: bit-equal to the same launches with overlap off."""
    from byox_vllm_b200.attention import paged_decode, set_decode_overlap

    B, Hq, Hkv = shape
    rng = np.random.default_rng(11)
    ctx = rng.integers(1, 2500, size=B).tolist() if B > 3 else [9000, 1, 5000]
    c = synth.decode_case(B, ctx, Hq, Hkv, seed=78)
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    lens = m["context_lens"].clone()
    slots = m["slot_mapping"].clone()
    if B > 3:
        pad = [0, 1, 2, B // 2, B // 2 + 1, B - 3, B - 2, B - 1]
    else:
        pad = [1]
    lens[pad] = 0
    slots[pad] = -1
    g = torch.Generator(device=cuda_dev).manual_seed(4)
    qs = [torch.empty_like(d["q"]).normal_(generator=g) for _ in range(10)]
    ks = [torch.empty_like(d["k"]).normal_(generator=g) for _ in range(10)]
    vs = [torch.empty_like(d["k"]).normal_(generator=g) for _ in range(10)]
    k0, v0 = d["k_cache"].clone(), d["v_cache"].clone()

    def run_all(outs):
        for i in range(10):
            kw = dict(key=ks[i], value=vs[i], slot_mapping=slots) if fused else {}
            paged_decode(qs[i], d["k_cache"], d["v_cache"], lens, m["block_tables"], c.scale,
                         out=outs[i], **kw)

    def fresh():
        d["k_cache"].copy_(k0)
        d["v_cache"].copy_(v0)
        torch.cuda.synchronize()
        return [torch.full_like(qs[0], float("nan")) for _ in range(10)]

    prev = set_decode_overlap(False)
    try:
        want = fresh()
        run_all(want)
        torch.cuda.synchronize()
        for i in (0, 9):
            assert torch.isfinite(want[i].float()).all()
            assert torch.count_nonzero(want[i][pad]) == 0
        if not fused:
            assert_close(want[7], orc.paged_decode_ref(qs[7].cpu(), c.k_cache, c.v_cache,
                                                       lens.cpu(), c.meta["block_tables"], c.scale),
                         "overlap off vs oracle")
        set_decode_overlap(True)
        got = fresh()
        run_all(got)
        torch.cuda.synchronize()
        for i in range(10):
            assert torch.equal(got[i], want[i]), f"eager launch {i} differs with overlap on"
        got = fresh()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            run_all(got)
        for _ in range(3):
            graph.replay()
        torch.cuda.synchronize()
        for i in range(10):
            assert torch.equal(got[i], want[i]), f"graph launch {i} differs with overlap on"
    finally:
        set_decode_overlap(prev)


def test_stream_engine_passes_the_decode_and_cascade_tests():
    """The persistent stream engine (B200ATTN_DECODE_ENGINE=stream; the engine is read once per
    process) against the same parity tests as the default rows engine."""
    import os
    import subprocess
    import sys

    if os.environ.get("B200ATTN_DECODE_ENGINE", "rows").startswith("s"):
        pytest.skip("already the stream engine")
    here = os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, B200ATTN_DECODE_ENGINE="stream")
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu",
                        os.path.join(here, "test_gpu_decode.py"),
                        os.path.join(here, "test_gpu_cascade.py"),
                        os.path.join(here, "test_gpu_qkv_fused.py")],
                       env=env, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
