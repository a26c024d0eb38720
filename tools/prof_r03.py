"""Launches for the round-3 ncu captures (one kind per process; `ncu -k regex:<kernel> -s 1 -c N`):

    python tools/prof_r03.py decode   # decode_stream_kernel at the c2 shapes: 16/8 (TP=1), 8/4 (TP=2),
                                      # 4/2 (TP=4); 2 launches each on different caches (cold)
    python tools/prof_r03.py store    # store_kvcache_kernel, 16384 tokens x 8 KV heads, 3 launches
    python tools/prof_r03.py qkv      # qk_norm_rope kernel, 16384 tokens x heads 8/1, 3 launches

Inputs are built exactly as bench.py's legs build them (bench.decode_inputs / extra_store /
extra_qkv); prints the algorithmic bytes per launch next to each shape.
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import bench  # noqa: E402

D, BLOCK = bench.D, bench.BLOCK
which = sys.argv[1] if len(sys.argv) > 1 else "decode"

with torch.inference_mode():
    if which == "decode":
        from byox_vllm_b200.attention import paged_decode

        for Hq, Hkv in ((16, 8), (8, 4), (4, 2)):
            lens_l = bench.context_lengths(dict(ctx="bench", batch=256))
            lens, sets, alg, _ = bench.decode_inputs(256, lens_l, Hq, Hkv, 2)
            print(f"decode B 256 bench lengths heads {Hq}/{Hkv}: algorithmic bytes per launch {alg}")
            for q, kc, vc, bt in sets:
                paged_decode(q, kc, vc, lens, bt, D ** -0.5)
            torch.cuda.synchronize()
            del sets
            torch.cuda.empty_cache()
    elif which == "store":
        from byox_vllm_b200.attention import store_kvcache

        N, Hkv = 16384, 8
        g = torch.Generator(device="cuda").manual_seed(7)
        for i in range(3):
            packed = torch.empty(N, 3 * Hkv * D, dtype=torch.bfloat16,
                                 device="cuda").normal_(generator=g)
            k = packed[:, :Hkv * D].view(N, Hkv, D)
            v = packed[:, 2 * Hkv * D:].view(N, Hkv, D)
            nb = (N + BLOCK - 1) // BLOCK * 2
            kc = torch.zeros(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda")
            vc = torch.zeros_like(kc)
            slots = torch.randperm(nb * BLOCK, generator=g, device="cuda")[:N].to(torch.int32)
            bench.l2_flush()
            store_kvcache(k, v, kc, vc, slots)
            torch.cuda.synchronize()
        print(f"store {N} x {Hkv}: algorithmic bytes per launch {N * Hkv * D * 2 * 2 * 2 + 4 * N}")
    elif which == "qkv":
        from byox_vllm_b200 import qkv_attention as fused

        N, Hq, Hkv, eps, max_pos = 16384, 8, 1, 1e-6, 40960
        g = torch.Generator(device="cuda").manual_seed(9)
        inv = 1.0 / (1e6 ** (torch.arange(0, D, 2, device="cuda").float() / D))
        fr = torch.arange(max_pos, device="cuda").float()[:, None] * inv[None]
        cache = torch.cat([fr.cos(), fr.sin()], dim=-1).contiguous()
        wq = torch.ones(D, dtype=torch.bfloat16, device="cuda")
        wk = torch.ones(D, dtype=torch.bfloat16, device="cuda")
        for i in range(3):
            qkv = torch.empty(N, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16,
                              device="cuda").normal_(generator=g)
            pos = torch.randint(0, max_pos, (N,), generator=g, device="cuda")
            nb = (N + BLOCK - 1) // BLOCK * 2
            kc = torch.zeros(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda")
            vc = torch.zeros_like(kc)
            slots = torch.randperm(nb * BLOCK, generator=g, device="cuda")[:N].to(torch.int32)
            bench.l2_flush()
            fused.qk_norm_rope_store(qkv, pos, wq, wk, eps, cache, Hq, Hkv, D, k_cache=kc,
                                     v_cache=vc, slot_mapping=slots)
            torch.cuda.synchronize()
        alg = N * ((Hq + 2 * Hkv) * 256 + Hq * 256 + Hkv * 256 + 2 * Hkv * 256 + 12)
        print(f"qkv {N} x {Hq}/{Hkv}: algorithmic bytes per launch {alg}")
    else:
        raise SystemExit(f"unknown kind {which}")
