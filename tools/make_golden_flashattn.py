"""Golden vectors for decode / prefill attention, produced by flash-attn ON A B200.

The reference's attention arithmetic is flash-attn (reference nanovllm/layers/attention.py:4);
it is CUDA-only, so these fixtures are generated under gpurun on the GPU box, calling
flash_attn_with_kvcache / flash_attn_varlen_func with exactly the arguments of the reference's
call sites (attention.py:75-86 and :88-96) on the seeded cases of tools/golden_cases.py.
Writes gpurun_out/golden/fa_<kind>_<name>.npz = {out (int16 view of bf16), input_sha}; copy them to
tests/golden/ and commit.

    gpurun -- python tools/make_golden_flashattn.py
"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from tools import golden_cases as gc, synth  # noqa: E402


def main():
    import flash_attn
    from flash_attn import flash_attn_varlen_func, flash_attn_with_kvcache

    dev = torch.device("cuda:0")
    out_dir = ROOT / "gpurun_out" / "golden"
    out_dir.mkdir(parents=True, exist_ok=True)
    print("flash_attn", flash_attn.__version__, torch.cuda.get_device_name(0))

    for name, ctx, Hq, Hkv, seed in gc.DECODE_CASES:
        c = synth.decode_case(len(ctx), ctx, Hq, Hkv, seed=seed)
        sha = synth.checksum(c.q, c.k_cache, c.v_cache, c.meta["context_lens"],
                             c.meta["block_tables"])
        o = flash_attn_with_kvcache(  # attention.py:88-96
            c.q.to(dev).unsqueeze(1), c.k_cache.to(dev), c.v_cache.to(dev),
            cache_seqlens=c.meta["context_lens"].to(dev),
            block_table=c.meta["block_tables"].to(dev), softmax_scale=c.scale, causal=True)
        np.savez_compressed(out_dir / f"fa_decode_{name}.npz",
                            out=o.squeeze(1).cpu().view(torch.int16).numpy(),
                            input_sha=np.array(sha))
        print(f"fa_decode_{name}: {tuple(o.shape)} sha={sha}")

    for name, lens, cached, Hq, Hkv, seed in gc.PREFILL_CASES:
        c = synth.prefill_case(lens, Hq, Hkv, cached=cached, seed=seed)
        m = c.meta
        sha = synth.checksum(c.q, c.k, c.v.contiguous(), c.k_cache, c.v_cache, m["cu_seqlens_q"],
                             m["cu_seqlens_k"])
        q = c.q.to(dev)
        if m["block_tables"] is None:
            packed = torch.empty(c.v.shape[0], (Hq + 2 * Hkv) * 128, dtype=c.v.dtype, device=dev)
            v = packed[:, (Hq + Hkv) * 128:].view(-1, Hkv, 128)
            v.copy_(c.v)  # strided on the device too, as in the model
            k = c.k.to(dev)
            bt = None
        else:
            # prefix-cache: this step's rows are scattered first (attention.py:70-71), then K/V are
            # the caches (attention.py:73-74). The scatter here is plain torch indexing.
            kc, vc = c.k_cache.to(dev), c.v_cache.to(dev)
            slots = m["slot_mapping"].to(dev).long()
            kc.view(-1, Hkv, 128)[slots] = c.k.to(dev)
            vc.view(-1, Hkv, 128)[slots] = c.v.to(dev)
            k, v, bt = kc, vc, m["block_tables"].to(dev)
        o = flash_attn_varlen_func(  # attention.py:75-86
            q, k, v, max_seqlen_q=m["max_seqlen_q"], cu_seqlens_q=m["cu_seqlens_q"].to(dev),
            max_seqlen_k=m["max_seqlen_k"], cu_seqlens_k=m["cu_seqlens_k"].to(dev),
            softmax_scale=c.scale, causal=True, block_table=bt)
        np.savez_compressed(out_dir / f"fa_prefill_{name}.npz",
                            out=o.cpu().view(torch.int16).numpy(), input_sha=np.array(sha))
        print(f"fa_prefill_{name}: {tuple(o.shape)} sha={sha}")

    # §7.3(b): what flash-attn returns for a context_lens == 0 row (CUDA-graph padding row)
    c = synth.decode_case(2, [300, 40], 16, 8, seed=99)
    lens = torch.tensor([300, 0], dtype=torch.int32, device=dev)
    bt = c.meta["block_tables"].to(dev).clamp_min(0)
    o = flash_attn_with_kvcache(c.q.to(dev).unsqueeze(1), c.k_cache.to(dev), c.v_cache.to(dev),
                                cache_seqlens=lens, block_table=bt, softmax_scale=c.scale,
                                causal=True)
    row = o[1].float()
    print(f"ctx0_row: absmax={row.abs().max().item()} nan={bool(torch.isnan(row).any())}")


if __name__ == "__main__":
    main()
