"""-m gpu: oracle parity on the exact shapes the bench records quote (VERDICT r02 "missing" #3).

c2 = BASELINE.json's headline decode step: B 256, context lengths from the reference bench's RNG
(bench.context_lengths: prompt + half of max_tokens, 169-1502), heads 16/8 and its tensor-parallel
shards 8/4, 4/2, 2/1. c3 = the target point B 128 x 4096, 32/8. The call is the one bench.py times:
the fused store + decode (reference attention.py:70-71 + :88-96), overlap mode off and on.
Checks: cache rows bit-exact vs store_kvcache_ref; every output row equal (within tolerance) to
num_splits=1; >= 8 sampled rows, the shortest and the longest among them, vs the fp32 oracle.
"""
import numpy as np
import pytest
import torch

import bench
from gpu_util import assert_close
from oracle import attention_oracle as orc
from tools import metadata_ref as md

pytestmark = pytest.mark.gpu

D, BLOCK = 128, 256


def _state(lens, Hq, Hkv, dev, seed):
    """Same construction as bench.build_state for one layer (shuffled block ids, N(0,1) cache)."""
    import random

    rng = random.Random(seed)
    need = [(c + BLOCK - 1) // BLOCK for c in lens]
    nb = sum(need) + 8
    ids = list(range(nb))
    rng.shuffle(ids)
    seqs, at = [], 0
    for c, n in zip(lens, need):
        seqs.append(md.SeqState(c, ids[at:at + n]))
        at += n
    meta = md.prepare_decode(seqs, BLOCK)
    g = torch.Generator(device=dev).manual_seed(seed)
    kc = torch.empty(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
    vc = torch.empty_like(kc).normal_(generator=g)
    B = len(lens)
    q = torch.empty(B, Hq, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
    kv = torch.empty(B, 2 * Hkv * D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
    k = kv[:, :Hkv * D].view(B, Hkv, D)   # strided views of fused rows, as qkv.split gives
    v = kv[:, Hkv * D:].view(B, Hkv, D)
    return seqs, meta, q, k, v, kc, vc


def _check(lens, Hq, Hkv, dev, seed):
    from byox_vllm_b200.attention import paged_decode, set_decode_overlap

    seqs, meta, q, k, v, kc0, vc0 = _state(lens, Hq, Hkv, dev, seed)
    lens_t = meta["context_lens"].to(dev)
    bt = meta["block_tables"].to(dev)
    slots = meta["slot_mapping"].to(dev)
    scale = D ** -0.5

    def run(splits, overlap):
        kc, vc = kc0.clone(), vc0.clone()
        prev = set_decode_overlap(overlap)
        try:
            o = paged_decode(q, kc, vc, lens_t, bt, scale, key=k, value=v, slot_mapping=slots,
                             num_splits=splits)
            torch.cuda.synchronize()
        finally:
            set_decode_overlap(prev)
        return o, kc, vc

    base, kc, vc = run(1, False)
    # a1 inside the fused launch: bit-exact vs the oracle scatter, nothing else touched
    want_k, want_v = kc0.clone(), vc0.clone()
    idx = slots.long()
    want_k.view(-1, Hkv, D)[idx] = k
    want_v.view(-1, Hkv, D)[idx] = v
    assert torch.equal(kc, want_k) and torch.equal(vc, want_v)
    rows = sorted({int(np.argmin(lens)), int(np.argmax(lens)), 0, len(lens) - 1,
                   *np.random.default_rng(seed).choice(len(lens), 6, replace=False).tolist()})
    assert len(rows) >= 8
    r = torch.tensor(rows)
    want = orc.paged_decode_ref(q[r].cpu(), kc.cpu(), vc.cpu(), meta["context_lens"][r],
                                meta["block_tables"][r], scale)
    assert_close(base[r], want, f"splits=1 rows {rows} vs oracle")
    for splits, overlap in ((None, False), (None, True), (3, True)):
        got, kc2, vc2 = run(splits, overlap)
        assert torch.equal(kc2, want_k) and torch.equal(vc2, want_v)
        assert_close(got, base, f"splits={splits} overlap={overlap} vs num_splits=1")
        assert_close(got[r], want, f"splits={splits} overlap={overlap} rows vs oracle")


@pytest.mark.parametrize("tp", [1, 2, 4, 8])
def test_c2_headline_shape_and_its_shards(cuda_dev, tp):
    w = bench.WORKLOADS["c2"]
    lens = bench.context_lengths(w)
    assert len(lens) == 256 and min(lens) == 169 and max(lens) == 1502  # VERDICT r02's numbers
    _check(lens, w["Hq"] // tp, w["Hkv"] // tp, cuda_dev, seed=100 + tp)


def test_c3_target_point(cuda_dev):
    w = bench.WORKLOADS["c3"]
    lens = bench.context_lengths(w)
    assert lens == [4096] * 128
    _check(lens, w["Hq"], w["Hkv"], cuda_dev, seed=300)
