"""N>1 host logic on CPU (gloo, world_size 2): the path shards by heads with no data-path
collective — each rank's head slice of the attention equals the same slice of the full result —
and harness timings are combined as the max over ranks."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from byox_vllm_b200 import parallel
from oracle import attention_oracle as orc
from tools import synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        Hq, Hkv = 16, 8
        c = synth.decode_case(3, [300, 17, 513], Hq, Hkv, seed=7)  # replicated metadata
        sh = parallel.shard_heads(Hq, Hkv, world, rank)
        qs = c.q[:, sh.q_head_start:sh.q_head_start + sh.num_q_heads]
        kc = c.k_cache[:, :, sh.kv_head_start:sh.kv_head_start + sh.num_kv_heads].contiguous()
        vc = c.v_cache[:, :, sh.kv_head_start:sh.kv_head_start + sh.num_kv_heads].contiguous()
        local = orc.paged_decode_ref(qs, kc, vc, c.meta["context_lens"], c.meta["block_tables"],
                                     c.scale)
        parts = [torch.empty_like(local) for _ in range(world)]
        dist.all_gather(parts, local)  # test-side gather only: the path itself has no collective
        full = orc.paged_decode_ref(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                                    c.meta["block_tables"], c.scale)
        ok = torch.allclose(torch.cat(parts, dim=1), full, atol=1e-6)
        # the fused step before the path shards the same way: rank r's QKV rows hold
        # its own q / k / v heads (QKVParallelLinear, linear.py:92-123), the norm weights and the
        # cos/sin table are replicated, so the rank-local result is the head slice of the full one
        gen = torch.Generator().manual_seed(3)
        D, n = 128, 5
        qf = torch.randn(n, Hq, D, generator=gen).to(torch.bfloat16)
        kf = torch.randn(n, Hkv, D, generator=gen).to(torch.bfloat16)
        vf = torch.randn(n, Hkv, D, generator=gen).to(torch.bfloat16)
        wq = (1 + 0.2 * torch.randn(D, generator=gen)).to(torch.bfloat16)
        wk = (1 + 0.2 * torch.randn(D, generator=gen)).to(torch.bfloat16)
        pos = torch.tensor([0, 7, 4095, 300, 1])
        cache = orc.rope_cos_sin_cache(D, 4096, 1e6)
        pack = lambda a, b, c_: torch.cat([a.flatten(1), b.flatten(1), c_.flatten(1)], dim=1)
        q_all, k_all, _ = orc.qk_norm_rope_ref(pack(qf, kf, vf), pos, wq, wk, 1e-6, cache, Hq, Hkv)
        qs_, ks_ = slice(sh.q_head_start, sh.q_head_start + sh.num_q_heads), slice(
            sh.kv_head_start, sh.kv_head_start + sh.num_kv_heads)
        q_loc, k_loc, _ = orc.qk_norm_rope_ref(pack(qf[:, qs_], kf[:, ks_], vf[:, ks_]), pos, wq,
                                               wk, 1e-6, cache, sh.num_q_heads, sh.num_kv_heads)
        ok = ok and torch.equal(q_loc, q_all[:, qs_]) and torch.equal(k_loc, k_all[:, ks_])
        t = parallel.max_over_ranks(10.0 + rank)
        q.put((rank, bool(ok), t, sh.group))
    finally:
        dist.destroy_process_group()


def test_head_sharding_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [True, True]
    assert [r[2] for r in res] == [11.0, 11.0]  # max over ranks on every rank
    assert [r[3] for r in res] == [2, 2]  # GQA group is unchanged by sharding


def test_shard_heads_matches_reference_layout():
    # Qwen3-8B (32/8) and 32B (64/8) over TP 1..8
    for tp, want in ((1, (32, 8)), (2, (16, 4)), (4, (8, 2)), (8, (4, 1))):
        shards = [parallel.shard_heads(32, 8, tp, r) for r in range(tp)]
        assert all((s.num_q_heads, s.num_kv_heads) == want for s in shards)
        assert [s.q_head_start for s in shards] == [r * want[0] for r in range(tp)]
        assert all(s.q_head_start // s.group == s.kv_head_start for s in shards)
    s = parallel.shard_heads(64, 8, 8, 5)
    assert (s.num_q_heads, s.num_kv_heads, s.group) == (8, 1, 8)
    assert parallel.max_over_ranks(3.5) == 3.5  # no process group -> identity
