"""Split count sweep of the rows decode engine on RAGGED batches (the per-rank shapes of BASELINE
config 2 under TP): back-to-back launches in one graph, overlap mode on/off.

    gpurun -- python tools/sweep_ragged_splits.py > gpurun_out/sweep_ragged_splits.log
"""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import bench  # noqa: E402
from byox_vllm_b200.attention import paged_decode, set_decode_overlap  # noqa: E402

SHAPES = [(256, "bench", 2, 1), (256, "bench", 4, 2), (256, "bench", 8, 4), (64, 1024, 32, 8),
          (16, 1024, 32, 8), (32, 1024, 32, 8)]

with torch.inference_mode():
    for B, ctx, Hq, Hkv in SHAPES:
        lens_l = bench.context_lengths(dict(ctx=ctx, batch=B))
        per_copy = sum(lens_l) * Hkv * 128 * 4
        n_copies = max(2, min(24, -(-(400 << 20) // per_copy)))
        lens, sets, alg, _ = bench.decode_inputs(B, lens_l, Hq, Hkv, n_copies)
        outs = [torch.empty_like(s[0]) for s in sets]
        for ov in (False, True):
            prev = set_decode_overlap(ov)
            row = {}
            for S in (None, 1, 2, 3, 4):
                def mk(i, S=S):
                    q, kc, vc, bt = sets[i]
                    return lambda: paged_decode(q, kc, vc, lens, bt, 128 ** -0.5, out=outs[i],
                                                num_splits=S)
                ms = bench.timed_graph_many(mk, n_copies)
                row[str(S)] = round(ms * 1e3, 2)
            set_decode_overlap(prev)
            pk = bench.peaks()[0]["hbm_gbs"]
            best = min(row, key=row.get)
            print(json.dumps({"B": B, "ctx": ctx, "heads": f"{Hq}/{Hkv}", "overlap": ov, "us": row,
                              "best": best, "best_frac": round(alg / row[best] / 1e3 / pk, 3),
                              "auto_frac": round(alg / row["None"] / 1e3 / pk, 3)}), flush=True)
        del sets, outs
        torch.cuda.empty_cache()
