"""-m gpu: shared-prefix (cascade) decode through the C-ABI. The result must equal plain paged
decode — the oracle is still paged_decode_ref — whatever the plan kernel finds; the plan itself is
checked bit-exactly against oracle.shared_prefix_pages_ref."""
import pytest
import torch

from gpu_util import assert_close, to_dev
from oracle import attention_oracle as orc
from tools import synth

pytestmark = pytest.mark.gpu


def _plan(d, block_size=256):
    from byox_vllm_b200.attention import cascade_plan

    m = d["meta"]
    plan = cascade_plan(m["context_lens"], m["block_tables"], block_size)
    torch.cuda.synchronize()
    return plan


PLAN_CASES = [
    # (name, ctx per row, shared pages per row or int)
    ("all_share_3", [900, 1000, 1100, 800], 3),
    ("limited_by_shortest", [300, 1000, 1100, 800], 3),      # row 0 has 1 page before its token
    ("page_boundary", [513, 769, 1025], 2),                  # ctx-1 = 512 -> exactly 2 full pages
    ("newest_token_opens_page", [512, 769, 1025], 2),        # ctx-1 = 511 -> only 1 page
    ("diverge_at_1", [900, 1000, 1100], [3, 1, 3]),
    ("unrelated_row", [900, 1000, 1100], [3, 3, 0]),
    ("single_row", [2000], 4),
    ("one_page_rows", [100, 200], 1),
]


@pytest.mark.parametrize("case", PLAN_CASES, ids=[c[0] for c in PLAN_CASES])
def test_plan_kernel_matches_oracle(cuda_dev, case):
    _, ctx, shared = case
    c = synth.shared_prefix_decode_case(ctx, shared, 8, 2, seed=51)
    d = to_dev(c, cuda_dev)
    got = _plan(d).cpu().tolist()
    P, b0, active = orc.shared_prefix_pages_ref(c.meta["context_lens"], c.meta["block_tables"], 256)
    assert got[0] == P and got[2] == active
    if P > 0:
        assert got[1] == b0


def test_plan_ignores_padding_rows(cuda_dev):
    """Graph replay: rows with context_lens == 0 carry stale tables and must not break sharing."""
    c = synth.shared_prefix_decode_case([900, 1000, 1100, 800], 3, 8, 2, seed=52)
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    lens = m["context_lens"].clone()
    bt = m["block_tables"].clone()
    lens[0] = 0
    bt[0] = -1
    from byox_vllm_b200.attention import cascade_plan

    plan = cascade_plan(lens, bt, 256).cpu().tolist()
    want = orc.shared_prefix_pages_ref(lens.cpu(), bt.cpu(), 256)
    assert (plan[0], plan[1], plan[2]) == want == (3, 1, 3)


CASCADE_CASES = [
    # (name, ctx, shared, Hq, Hkv)
    ("g8_tp8_share4", [1100 + 37 * i for i in range(40)], 4, 8, 1),
    ("g4_share2_ragged", [600 + 91 * i for i in range(19)], 2, 16, 4),
    ("g2_share3", [800 + 13 * i for i in range(70)], 3, 16, 8),
    ("g1_share2", [700 + 5 * i for i in range(9)], 2, 4, 4),
    ("no_sharing", [600 + 91 * i for i in range(10)], 0, 8, 2),
    ("partial_divergence", [1300, 1400, 1500, 1600, 1700], [5, 5, 2, 5, 5], 8, 1),
    ("exact_page_boundary", [1025, 1281, 1025], 4, 8, 2),
]


@pytest.mark.parametrize("case", CASCADE_CASES, ids=[c[0] for c in CASCADE_CASES])
@pytest.mark.parametrize("splits", [1, 3])
@pytest.mark.parametrize("chunks", [1, 3, 16])
def test_cascade_decode_matches_oracle_and_plain(cuda_dev, case, splits, chunks):
    from byox_vllm_b200.attention import paged_decode

    name, ctx, shared, Hq, Hkv = case
    c = synth.shared_prefix_decode_case(ctx, shared, Hq, Hkv, seed=53)
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    plan = _plan(d)
    got = paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"], m["block_tables"],
                       c.scale, num_splits=splits, plan=plan, prefix_chunks=chunks)
    plain = paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"], m["block_tables"],
                         c.scale, num_splits=splits)
    torch.cuda.synchronize()
    want = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                                c.meta["block_tables"], c.scale)
    assert torch.isfinite(got.float()).all()
    assert_close(got, want, f"{name} cascade vs oracle")
    assert_close(got, plain.float(), f"{name} cascade vs plain decode")
    if plan.cpu()[0].item() == 0 and splits == 1:
        # nothing shared: the cascade must degrade to the very same arithmetic
        assert torch.equal(got.view(torch.int16), plain.view(torch.int16))


def test_cascade_fused_store_and_padding_rows(cuda_dev):
    """Fused scatter + cascade with graph-style padding rows (ctx 0, slot -1, poisoned q)."""
    from byox_vllm_b200.attention import paged_decode

    ctx = [1500, 1400, 1300, 1600, 1450, 1550]
    c = synth.shared_prefix_decode_case(ctx, 5, 16, 4, seed=54)
    d = to_dev(c, cuda_dev)
    m = d["meta"]
    lens = m["context_lens"].clone()
    slots = m["slot_mapping"].clone()
    bt = m["block_tables"].clone()
    q = d["q"].clone()
    for b in (1, 4):
        lens[b] = 0
        slots[b] = -1
        bt[b] = 2 ** 30
        q[b] = float("nan")
    from byox_vllm_b200.attention import cascade_plan

    plan = cascade_plan(lens, bt, 256)
    for splits in (1, 2):
        kc, vc = d["k_cache"].clone(), d["v_cache"].clone()
        got = paged_decode(q, kc, vc, lens, bt, c.scale, key=d["k"], value=d["v"],
                           slot_mapping=slots, num_splits=splits, plan=plan, prefix_chunks=4)
        torch.cuda.synchronize()
        kc_ref, vc_ref = c.k_cache.clone(), c.v_cache.clone()
        orc.store_kvcache_ref(c.k, c.v, kc_ref, vc_ref, slots.cpu())
        want = orc.paged_decode_ref(q.cpu(), kc_ref, vc_ref, lens.cpu(), bt.cpu(), c.scale)
        assert torch.count_nonzero(got[1]) == 0 and torch.count_nonzero(got[4]) == 0
        keep = [0, 2, 3, 5]
        assert_close(got[keep], want[keep], f"cascade + fused store, splits={splits}")
        assert torch.equal(kc.cpu().view(torch.int16), kc_ref.view(torch.int16))
        assert torch.equal(vc.cpu().view(torch.int16), vc_ref.view(torch.int16))
    assert plan.cpu()[0].item() == 5


def test_cascade_module_under_graph_replay(cuda_dev):
    """Attention.forward arms the cascade from host-visible sizes, computes the plan once per
    Context (not per layer) and stays capturable: replay with different metadata (a batch that
    shares a prefix, then one that does not) must track the oracle both times."""
    from byox_vllm_b200 import _lib
    from byox_vllm_b200.attention import Attention, cascade_chunks
    from byox_vllm_b200.context import get_context, reset_context, set_context

    dev = cuda_dev
    Hq, Hkv, D, B, W = 8, 2, 128, 64, 32  # a shape the size heuristic arms by itself
    assert cascade_chunks(B, Hq, Hkv, W, 256) > 0
    assert cascade_chunks(B, Hq, Hkv, 16, 256) == 0 and cascade_chunks(8, Hq, Hkv, W, 256) == 0
    shared = synth.shared_prefix_decode_case([1300 + 5 * i for i in range(B)], 5, Hq, Hkv,
                                             seed=55, spare_blocks=48)
    ds = to_dev(shared, dev)
    layers = [Attention(Hq, D, shared.scale, Hkv) for _ in range(3)]
    for a in layers:
        a.k_cache, a.v_cache = ds["k_cache"], ds["v_cache"]
    g_q = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device=dev)
    g_k = torch.zeros(B, Hkv, D, dtype=torch.bfloat16, device=dev)
    g_v = torch.zeros(B, Hkv, D, dtype=torch.bfloat16, device=dev)
    g_slots = torch.full((B,), -1, dtype=torch.int32, device=dev)
    g_lens = torch.zeros(B, dtype=torch.int32, device=dev)
    g_bt = torch.zeros(B, W, dtype=torch.int32, device=dev)
    outs = [torch.zeros(B, 1, Hq, D, dtype=torch.bfloat16, device=dev) for _ in layers]
    set_context(False, slot_mapping=g_slots, context_lens=g_lens, block_tables=g_bt)
    for a, o in zip(layers, outs):
        o.copy_(a(g_q, g_k, g_v))  # warm-up
    n0 = _lib.load().b200attn_launch_count()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for a, o in zip(layers, outs):
            o.copy_(a(g_q, g_k, g_v))
    launched = _lib.load().b200attn_launch_count() - n0
    assert launched == 1 + 2 * len(layers), "one plan launch per step + (prefix, decode) per layer"
    assert getattr(get_context(), "_b200_cascade_plan", None) is not None
    reset_context()

    def replay(case):
        m = case.meta
        n = m["block_tables"].size(1)
        g_q.copy_(case.q.to(dev))
        g_slots.fill_(-1)
        g_lens.copy_(m["context_lens"].to(dev))
        g_bt.fill_(-1)
        g_bt[:, :n] = m["block_tables"].to(dev)
        graph.replay()
        torch.cuda.synchronize()
        want = orc.paged_decode_ref(case.q, shared.k_cache, shared.v_cache, m["context_lens"],
                                    m["block_tables"], case.scale)
        for o in outs:
            assert_close(o.squeeze(1), want, "graph replay")

    replay(shared)
    # same cache, rows now point at unrelated pages (no common prefix): the plan kernel re-runs
    # inside the graph and the prefix pass switches itself off
    other = synth.decode_case(B, [300 + 3 * i for i in range(B)], Hq, Hkv, seed=56,
                              fill_cache=False)
    limit = shared.k_cache.size(0)
    assert int(other.meta["block_tables"].max()) < limit
    replay(other)
    replay(shared)


def test_cascade_at_config4_size(cuda_dev):
    """BASELINE config 4's per-rank shape: Qwen3-32B at TP=8 -> heads 8/1, batch 256, ctx 8192
    with a 7680-token (30-page) shared prefix + 512 own tokens. The size heuristic must arm the
    cascade by itself; the plan must find the 30 pages; the result must equal plain paged decode
    on every row and the oracle on sampled rows; the fused store must leave the cache bit-exact."""
    from byox_vllm_b200.attention import cascade_chunks, cascade_plan, paged_decode

    B, Hq, Hkv, ctx, P = 256, 8, 1, 8192, 30
    c = synth.shared_prefix_decode_case([ctx] * B, P, Hq, Hkv, seed=57, device=cuda_dev)
    m = {k: (v.to(cuda_dev) if isinstance(v, torch.Tensor) else v) for k, v in c.meta.items()}
    W = m["block_tables"].size(1)
    chunks = cascade_chunks(B, Hq, Hkv, W, 256)
    assert chunks > 0, "the size heuristic must arm the cascade for config 4"
    plan = cascade_plan(m["context_lens"], m["block_tables"], 256)
    assert plan.cpu()[0].item() == P
    kc, vc = c.k_cache.clone(), c.v_cache.clone()
    got = paged_decode(c.q, kc, vc, m["context_lens"], m["block_tables"], c.scale, key=c.k,
                       value=c.v, slot_mapping=m["slot_mapping"], plan=plan, prefix_chunks=chunks)
    kc2, vc2 = c.k_cache.clone(), c.v_cache.clone()
    plain = paged_decode(c.q, kc2, vc2, m["context_lens"], m["block_tables"], c.scale, key=c.k,
                         value=c.v, slot_mapping=m["slot_mapping"])
    torch.cuda.synchronize()
    assert torch.equal(kc.view(torch.int16), kc2.view(torch.int16))
    assert torch.equal(vc.view(torch.int16), vc2.view(torch.int16))
    assert_close(got, plain.float(), "config 4 cascade vs plain decode")
    k_cpu, v_cpu = kc.cpu(), vc.cpu()
    for b in (0, 101, B - 1):
        want = orc.paged_decode_ref(c.q[b:b + 1].cpu(), k_cpu, v_cpu,
                                    c.meta["context_lens"][b:b + 1].cpu(),
                                    c.meta["block_tables"][b:b + 1].cpu(), c.scale)
        assert_close(got[b:b + 1], want, f"config 4 row {b} vs oracle")
