"""Three launches of the decode kernel per shape, for ncu: the per-rank shape of BASELINE config 2
at TP=8 (B 256, bench lengths, heads 2/1), then config 3 (B 128 x ctx 4096, heads 32/8).

    ncu --set full --clock-control none --import-source on -k regex:decode_stream_kernel -s 2 -c 1 \
        -o gpurun_out/prof_decode_c2_tp8 python tools/prof_decode.py
    ncu ... -s 5 -c 1 -o gpurun_out/prof_decode_c3 python tools/prof_decode.py
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import bench  # noqa: E402
from byox_vllm_b200.attention import paged_decode  # noqa: E402

with torch.inference_mode():
    for B, ctx, Hq, Hkv in ((256, "bench", 2, 1), (128, 4096, 32, 8)):
        lens_l = bench.context_lengths(dict(ctx=ctx, batch=B))
        lens, sets, alg, _ = bench.decode_inputs(B, lens_l, Hq, Hkv, 3)
        print(f"B {B} ctx {ctx} heads {Hq}/{Hkv}: algorithmic bytes per launch {alg}")
        for q, kc, vc, bt in sets:
            paged_decode(q, kc, vc, lens, bt, 128 ** -0.5)
        torch.cuda.synchronize()
        del sets
        torch.cuda.empty_cache()
