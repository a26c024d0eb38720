"""Small-grid decode shapes (latency / per-CTA in-flight bound): 3-stage vs 6-stage ring.

    gpurun -- 'for s in 3 0; do B200ATTN_DECODE_STAGES=$s python tools/sweep_small_decode.py; done'
"""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import bench  # noqa: E402

SHAPES = [(16, 1024, 32, 8), (8, 2048, 16, 8), (4, 4096, 32, 8), (1, 4096, 32, 8), (1, 1024, 32, 8),
          (8, 16384, 32, 8), (2, 32768, 32, 8), (1, 131072, 32, 8), (16, 4096, 32, 8),
          (128, 4096, 4, 1), (64, 1024, 8, 1), (32, 1024, 32, 8)]
with torch.inference_mode():
    for ov in (False, True):
        for B, ctx, Hq, Hkv in SHAPES:
            r = bench.extra_decode("x", B, ctx, Hq, Hkv, ref=False, overlap=ov)
            print(json.dumps({"stages_env": os.environ.get("B200ATTN_DECODE_STAGES", "auto"),
                              "overlap": ov, "B": B, "ctx": ctx, "heads": f"{Hq}/{Hkv}",
                              "ctas": r["decode_ctas"], "us": round(r["ms_per_launch"] * 1e3, 2),
                              "frac": round(r["frac"], 3)}), flush=True)
