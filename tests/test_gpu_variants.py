"""-m gpu: kernel variants selected by environment variables (read once per process) must agree
bit for bit on the same inputs. Each variant runs in its own subprocess and prints a checksum."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

_SCRIPT = r"""
import sys, torch
sys.path.insert(0, {root!r})
from byox_vllm_b200 import qkv_attention as fused
from byox_vllm_b200.attention import store_kvcache
from oracle import attention_oracle as orc      # checker side: cos/sin table + scatter oracle
from tools import golden_cases as gc, synth
dev = torch.device("cuda:0")
out = []
# ---- store_kvcache: prefill-size scatter with holes, strided V ----
for n, Hkv in ((16384, 8), (2048, 4), (3000, 1)):
    k, v, kc, vc, slots = gc.scatter_inputs(n, Hkv, (n + 255) // 256 + 2, 256, 7, seed=n)
    kc_ref, vc_ref = kc.clone(), vc.clone()
    orc.store_kvcache_ref(k, v, kc_ref, vc_ref, slots)
    packed = torch.zeros(n, v.stride(0), dtype=v.dtype, device=dev)
    vv = packed[:, v.stride(0) - Hkv * 128:].view(n, Hkv, 128)
    vv.copy_(v)
    kcd, vcd = kc.to(dev), vc.to(dev)
    store_kvcache(k.to(dev), vv, kcd, vcd, slots.to(dev))
    torch.cuda.synchronize()
    assert torch.equal(kcd.cpu().view(torch.int16), kc_ref.view(torch.int16))
    assert torch.equal(vcd.cpu().view(torch.int16), vc_ref.view(torch.int16))
    out.append(synth.checksum(kcd, vcd))
# ---- q/k norm + RoPE + store: both kernels on prefill- and decode-size batches ----
for n, Hq, Hkv, norm in ((4096, 32, 8, True), (2048, 16, 8, False), (300, 16, 8, True),
                         (2500, 8, 1, True)):
    qkv, pos, qw, kw = gc.normrope_inputs(n, Hq, Hkv, norm, 8192, seed=n)
    cache = orc.rope_cos_sin_cache(128, 8192, 1e6).to(dev)
    g = torch.Generator().manual_seed(n)
    nb = (n + 255) // 256 + 3
    slots = torch.randperm(nb * 256, generator=g)[:n].to(torch.int32)
    slots[::5] = -1
    kc = torch.zeros(nb, 256, Hkv, 128, dtype=torch.bfloat16, device=dev)
    vc = torch.zeros_like(kc)
    to = lambda t: None if t is None else t.to(dev)
    work = qkv.to(dev)
    q, k, v = fused.qk_norm_rope_store(work, pos.to(dev), to(qw), to(kw), 1e-6, cache, Hq, Hkv,
                                       k_cache=kc, v_cache=vc, slot_mapping=slots.to(dev))
    torch.cuda.synchronize()
    out.append(synth.checksum(q, work, kc, vc))
print("CHECKSUMS " + " ".join(out))
"""


def _run(env_extra):
    env = dict(os.environ)
    for k in ("B200ATTN_NORMROPE", "B200ATTN_NORMROPE_WARPS"):
        env.pop(k, None)
    env.update(env_extra)
    r = subprocess.run([sys.executable, "-c", _SCRIPT.format(root=str(ROOT))], env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("CHECKSUMS ")][-1]
    return line.split()[1:]


def test_kernel_variants_agree_bit_for_bit(cuda_dev):
    base = _run({})
    assert len(base) == 7
    assert _run({"B200ATTN_NORMROPE": "bulk"}) == base
    assert _run({"B200ATTN_NORMROPE": "direct"}) == base
    assert _run({"B200ATTN_NORMROPE": "bulk", "B200ATTN_NORMROPE_WARPS": "8"}) == base
