"""-m gpu: the drop-in Attention module driven like ModelRunner drives it — a prefill step, then
decode steps that append tokens (crossing a page boundary) — checked step by step against the
oracle's forward on CPU copies of the same cache. Cache contents must stay bit-identical."""
import random

import pytest
import torch

from gpu_util import assert_close
from oracle import attention_oracle as orc
from tools import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("Hq,Hkv", [(16, 8), (8, 1)])
def test_prefill_then_decode_steps(cuda_dev, Hq, Hkv):
    from tools import metadata_ref as md
    from byox_vllm_b200.attention import Attention
    from byox_vllm_b200.context import reset_context, set_context

    dev, D, bs = cuda_dev, 128, 256
    rng = random.Random(3)
    lens = [254, 600, 17]
    free = list(range(12))
    rng.shuffle(free)
    seqs = []
    for n in lens:
        nb = (n + bs - 1) // bs
        seqs.append(md.SeqState(n, [free.pop() for _ in range(nb)]))
    gen = torch.Generator().manual_seed(9)
    kc = torch.zeros(12, bs, Hkv, D, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    attn = Attention(Hq, D, D ** -0.5, Hkv)
    attn.k_cache, attn.v_cache = kc.to(dev), vc.to(dev)

    def step(meta, n_rows):
        q, k, v = synth._qkv(n_rows, Hq, Hkv, D, gen, "cpu")
        set_context(**md.context_kwargs(meta, dev))
        try:
            got = attn(q.to(dev), k.to(dev), v.to(dev))
            torch.cuda.synchronize()
        finally:
            reset_context()
        ctx = type("Ctx", (), dict(meta))()
        want = orc.attention_forward_ref(q, k, v, kc, vc, ctx, D ** -0.5)
        assert got.shape == want.shape
        assert_close(got, want, "module step")
        assert torch.equal(attn.k_cache.cpu().view(torch.int16), kc.view(torch.int16))
        assert torch.equal(attn.v_cache.cpu().view(torch.int16), vc.view(torch.int16))

    step(md.prepare_prefill(seqs, bs), sum(lens))
    for _ in range(4):  # 254 -> 258 crosses into a second page (may_append, block_manager.py:90-107)
        for s in seqs:
            s.num_tokens += 1
            if (s.num_tokens + bs - 1) // bs > len(s.block_table):
                s.block_table.append(free.pop())
        step(md.prepare_decode(seqs, bs), len(seqs))
