"""Golden vectors for the fused q/k RMSNorm + RoPE step, produced by the REFERENCE'S OWN modules.

Runs reference nanovllm/layers/layernorm.py:RMSNorm and layers/rotary_embedding.py:RotaryEmbedding
exactly as nanovllm/models/qwen3.py:75-83 calls them, on CPU tensors, for every case of
tools/golden_cases.NORMROPE_CASES, and writes tests/golden/normrope_<name>.npz = {q, k (int16
views of the bf16 results), input_sha}. TORCHDYNAMO_DISABLE=1 makes the modules' @torch.compile
decorators no-ops, so the fixtures hold the eager arithmetic (every rounding point of the Python
source). Needs /root/reference (this container only); the fixtures travel, the reference does not.

    python tools/make_golden_normrope.py
"""
import os
import sys
from pathlib import Path

os.environ["TORCHDYNAMO_DISABLE"] = "1"
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, "/root/reference")

import numpy as np  # noqa: E402
import torch  # noqa: E402

from tools import golden_cases as gc, synth  # noqa: E402


def main():
    from nanovllm.layers.layernorm import RMSNorm  # the reference's own code
    from nanovllm.layers.rotary_embedding import RotaryEmbedding

    out_dir = ROOT / "tests" / "golden"
    D = 128
    for name, n, Hq, Hkv, has_norm, max_pos, base, seed in gc.NORMROPE_CASES:
        qkv, positions, q_w, k_w = gc.normrope_inputs(n, Hq, Hkv, has_norm, max_pos, seed)
        sha = synth.checksum(qkv, positions, *([q_w, k_w] if has_norm else []))
        rope = RotaryEmbedding(D, D, max_pos, base)
        q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)  # qwen3.py:75-79
        q = q.view(-1, Hq, D)
        k = k.view(-1, Hkv, D)
        if has_norm:
            torch.set_default_dtype(torch.bfloat16)  # model_runner.py:30: params in model dtype
            q_norm, k_norm = RMSNorm(D, gc.NORMROPE_EPS), RMSNorm(D, gc.NORMROPE_EPS)
            torch.set_default_dtype(torch.float32)
            with torch.no_grad():
                q_norm.weight.copy_(q_w)
                k_norm.weight.copy_(k_w)
                q, k = q_norm(q), k_norm(k)
        with torch.no_grad():
            q, k = rope(positions, q, k)
        np.savez_compressed(out_dir / f"normrope_{name}.npz",
                            q=q.contiguous().view(torch.int16).numpy(),
                            k=k.contiguous().view(torch.int16).numpy(),
                            cos_sin_sha=np.array(synth.checksum(rope.cos_sin_cache.squeeze(1))),
                            input_sha=np.array(sha))
        print(f"normrope_{name}: N={n} heads {Hq}/{Hkv} norm={has_norm} sha={sha}")


if __name__ == "__main__":
    main()
