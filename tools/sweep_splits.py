"""Sweep the split count of the paged decode kernel on one B200 to calibrate
b200attn_decode_pick_splits (host-only heuristic). Prints, per shape, the time of every split
count and which one the library would pick.

    gpurun -- python tools/sweep_splits.py
"""
from __future__ import annotations

import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from tools.microbench import BLOCK, D, PEAK, decode_inputs, timed  # noqa: E402

SHAPES = [(1, 1024, 32, 8), (1, 4096, 32, 8), (4, 4096, 32, 8), (8, 2048, 16, 8), (16, 4096, 32, 8),
          (32, 2048, 16, 8), (64, 1024, 16, 8), (64, 4096, 32, 8), (128, 4096, 4, 1),
          (128, 4096, 8, 2), (8, 16384, 32, 8), (2, 32768, 32, 8), (1, 131072, 32, 8),
          (256, 8192, 8, 1), (48, 800, 16, 8), (16, 600, 16, 8)]
CANDS = [1, 2, 3, 4, 5, 6, 8, 10, 12, 16, 20, 24, 32, 37, 48, 64]  # API limit: 64 splits


def main():
    from byox_vllm_b200.attention import decode_num_splits, paged_decode

    out = open(ROOT / "gpurun_out" / "sweep_splits.jsonl", "w")
    with torch.inference_mode():
        for (B, ctx, Hq, Hkv) in SHAPES:
            q, kc, vc, lens, bt, alg = decode_inputs(B, ctx, Hq, Hkv, False)
            o = torch.empty_like(q)
            picked = decode_num_splits(B, Hkv, bt.size(1), BLOCK)
            max_tiles = bt.size(1) * BLOCK // 64
            res = {}
            for S in CANDS:
                if S > max(1, max_tiles):
                    continue
                _, t = timed(lambda: paged_decode(q, kc, vc, lens, bt, D ** -0.5, out=o,
                                                  num_splits=S), iters=8)
                res[S] = t * 1e3
            best = min(res, key=res.get)
            rec = dict(B=B, ctx=ctx, Hq=Hq, Hkv=Hkv, picked=picked, best=best, us=res,
                       picked_us=res.get(picked), best_gbs=alg / res[best] / 1e3,
                       best_frac=alg / res[best] / 1e3 / PEAK)
            out.write(json.dumps(rec) + "\n")
            out.flush()
            row = " ".join(f"{S}:{t:.1f}" for S, t in res.items())
            print(f"B={B:4d} ctx={ctx:6d} {Hq}/{Hkv} picked={picked} best={best} "
                  f"({rec['best_frac']:.2f}) | {row}", flush=True)
            del q, kc, vc, lens, bt
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
