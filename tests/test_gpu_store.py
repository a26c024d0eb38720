"""-m gpu: store_kvcache through the C-ABI is bit-exact against the oracle and the golden
vectors of the reference's own Triton kernel."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import attention_oracle as orc
from tools import golden_cases as gc

pytestmark = pytest.mark.gpu


def _run(dev, k, v, kc, vc, slots):
    from byox_vllm_b200.attention import store_kvcache

    n, Hkv, D = k.shape
    kd = k.to(dev)
    width = max(v.stride(0), Hkv * D)
    packed = torch.zeros(max(n, 1), width, dtype=v.dtype, device=dev)
    vd = packed[:n, width - Hkv * D:].view(n, Hkv, D)
    vd.copy_(v)
    kcd, vcd = kc.to(dev), vc.to(dev)
    store_kvcache(kd, vd, kcd, vcd, slots.to(dev))
    torch.cuda.synchronize()
    return kcd.cpu(), vcd.cpu()


@pytest.mark.parametrize("case", gc.SCATTER_CASES, ids=[c[0] for c in gc.SCATTER_CASES])
def test_store_matches_reference_golden(cuda_dev, case):
    name, n, Hkv, nb, bs, hole, seed = case
    g = np.load(GOLDEN / f"scatter_{name}.npz")
    k, v, kc, vc, slots = gc.scatter_inputs(n, Hkv, nb, bs, hole, seed)
    kg, vg = _run(cuda_dev, k, v, kc, vc, slots)
    assert np.array_equal(kg.view(torch.int16).numpy(), g["k_cache"])
    assert np.array_equal(vg.view(torch.int16).numpy(), g["v_cache"])


@pytest.mark.parametrize("n", [1, 128, 2047, 2048, 16384])
@pytest.mark.parametrize("Hkv", [8, 4, 2, 1])
def test_store_matches_oracle_sizes(cuda_dev, n, Hkv):
    bs = 256
    nb = (n + bs - 1) // bs + 2
    k, v, kc, vc, slots = gc.scatter_inputs(n, Hkv, nb, bs, 7 if n > 8 else 0, seed=n + Hkv)
    kg, vg = _run(cuda_dev, k, v, kc, vc, slots)
    orc.store_kvcache_ref(k, v, kc, vc, slots)
    assert torch.equal(kg.view(torch.int16), kc.view(torch.int16))
    assert torch.equal(vg.view(torch.int16), vc.view(torch.int16))


def test_store_edge_cases(cuda_dev):
    from byox_vllm_b200.attention import store_kvcache

    k, v, kc, vc, slots = gc.scatter_inputs(9, 2, 2, 16, 0, seed=1)
    # all slots -1 (graph padding rows) and other negative values: nothing is written
    for bad in (-1, -12):
        kg, vg = _run(cuda_dev, k, v, kc, vc, torch.full_like(slots, bad))
        assert torch.equal(kg.view(torch.int16), kc.view(torch.int16))
        assert torch.equal(vg.view(torch.int16), vc.view(torch.int16))
    # N == 0 is a no-op
    e = torch.zeros(0, 2, 128, dtype=torch.bfloat16, device=cuda_dev)
    kcd = kc.to(cuda_dev)
    store_kvcache(e, e, kcd, kcd.clone(), torch.zeros(0, dtype=torch.int32, device=cuda_dev))
    assert torch.equal(kcd.cpu().view(torch.int16), kc.view(torch.int16))
    # bit patterns (NaN payloads, -0, denormals) survive: it is a byte copy
    raw = torch.arange(9 * 2 * 128, dtype=torch.int32).mul(2654435761 % 65536).to(torch.int16)
    kb = raw.view(9, 2, 128).view(torch.bfloat16)
    kg, _ = _run(cuda_dev, kb, kb, kc, vc, slots)
    want = kc.clone()
    orc.store_kvcache_ref(kb, kb, want, vc.clone(), slots)
    assert torch.equal(kg.view(torch.int16), want.view(torch.int16))


def test_store_full_prefill_size_roundtrip(cuda_dev):
    """BASELINE-size property test (16384 tokens x Hkv 8 = 64 MiB): scatter through a random
    permutation then gather back through the same permutation returns the input exactly."""
    from byox_vllm_b200.attention import store_kvcache

    dev = cuda_dev
    n, Hkv, D, bs = 16384, 8, 128, 256
    g = torch.Generator(device=dev).manual_seed(0)
    k = torch.randn(n, Hkv, D, device=dev, generator=g).to(torch.bfloat16)
    qkv = torch.randn(n, 4 * Hkv * D, device=dev, generator=g).to(torch.bfloat16)
    v = qkv[:, 3 * Hkv * D:].view(n, Hkv, D)
    kc = torch.zeros(n // bs + 4, bs, Hkv, D, dtype=torch.bfloat16, device=dev)
    vc = torch.zeros_like(kc)
    slots = torch.randperm(kc.shape[0] * bs, device=dev, generator=g)[:n].to(torch.int32)
    store_kvcache(k, v, kc, vc, slots)
    idx = slots.long()
    assert torch.equal(kc.view(-1, Hkv, D)[idx], k)
    assert torch.equal(vc.view(-1, Hkv, D)[idx], v)
    untouched = torch.ones(kc.shape[0] * bs, dtype=torch.bool, device=dev)
    untouched[idx] = False
    assert torch.count_nonzero(kc.view(-1, Hkv * D)[untouched]) == 0
