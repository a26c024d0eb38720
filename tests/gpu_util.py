"""Helpers shared by the -m gpu parity tests."""
import torch

from oracle import attention_oracle as orc

ATOL, RTOL = 2e-2, 1e-2  # north_star tolerance for bf16 attention outputs


def to_dev(case, dev):
    """Move a synth case to the device keeping v a strided view (as qkv.split gives)."""
    out = {}
    for name in ("q", "k", "k_cache", "v_cache"):
        out[name] = getattr(case, name).to(dev)
    v = case.v
    n, Hkv, D = v.shape
    width = v.stride(0) if n > 1 else Hkv * D * 3
    packed = torch.zeros(n, width, dtype=v.dtype, device=dev)
    vv = packed[:, width - Hkv * D:].view(n, Hkv, D)
    vv.copy_(v)
    out["v"] = vv
    meta = {}
    for k, val in case.meta.items():
        meta[k] = val.to(dev) if isinstance(val, torch.Tensor) else val
    out["meta"] = meta
    return out


def assert_close(got, want, what=""):
    ok, ratio = orc.within_tolerance(got, want, ATOL, RTOL)
    assert ok, f"{what}: |diff| exceeds 2e-2 + 1e-2|ref| by factor {ratio:.3f}"
    return ratio
