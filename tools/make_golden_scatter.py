"""Golden vectors for the KV scatter, produced by the REFERENCE'S OWN Triton kernel.

Runs reference nanovllm/layers/attention.py:store_kvcache (kernel :10-31, launcher :34-49) on CPU
tensors under TRITON_INTERPRET=1 — no GPU needed — for every case of tools/golden_cases.py and
writes tests/golden/scatter_<name>.npz = {k_cache, v_cache (int16 views of the bf16 results),
input_sha}. Needs /root/reference (this container only); the fixtures travel, the reference
does not.

    TRITON_INTERPRET=1 python tools/make_golden_scatter.py
"""
import os
import sys
from pathlib import Path

os.environ.setdefault("TRITON_INTERPRET", "1")
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, "/root/reference")

import numpy as np  # noqa: E402
import torch  # noqa: E402

from tools import golden_cases as gc, synth  # noqa: E402


def main():
    from nanovllm.layers.attention import store_kvcache  # the reference's own code

    out_dir = ROOT / "tests" / "golden"
    out_dir.mkdir(parents=True, exist_ok=True)
    for name, n, Hkv, nb, bs, hole, seed in gc.SCATTER_CASES:
        k, v, k_cache, v_cache, slots = gc.scatter_inputs(n, Hkv, nb, bs, hole, seed)
        sha = synth.checksum(k, v.contiguous(), k_cache, v_cache, slots)
        store_kvcache(k, v, k_cache, v_cache, slots)
        np.savez_compressed(out_dir / f"scatter_{name}.npz",
                            k_cache=k_cache.view(torch.int16).numpy(),
                            v_cache=v_cache.view(torch.int16).numpy(),
                            input_sha=np.array(sha))
        print(f"scatter_{name}: N={n} Hkv={Hkv} slots={nb * bs} sha={sha}")


if __name__ == "__main__":
    main()
