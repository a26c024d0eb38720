"""Calibrate the tunables of the decode kernel's unit stream on one B200:
B200ATTN_DECODE_MIN_TILES (fewest units per CTA range), B200ATTN_DECODE_ROW_COST (virtual units
per row) and B200ATTN_DECODE_SNAP (boundary snapping). They are read once per process, so every combination runs in its own subprocess over
the same shapes (graph replay over independent cache sets larger than L2, CUDA events).

    gpurun -- python tools/sweep_decode_tunables.py            # grid -> gpurun_out/decode_tunables.jsonl
    python tools/sweep_decode_tunables.py --child              # one combination (env), prints JSON
"""
from __future__ import annotations

import itertools
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
if os.environ.get("B200_AB_PKG"):  # A/B against another build of the package (e.g. an older tree)
    sys.path.insert(0, str(ROOT / os.environ["B200_AB_PKG"]))

SHAPES = [  # (B, ctx or "bench", Hq, Hkv)
    (256, "bench", 2, 1), (256, "bench", 4, 2), (256, "bench", 16, 8),
    (64, 1024, 32, 8), (32, 1024, 32, 8), (16, 1024, 32, 8), (4, 4096, 32, 8), (1, 4096, 32, 8),
    (1, 131072, 32, 8), (8, 16384, 32, 8), (64, 4096, 32, 8), (128, 4096, 32, 8),
    (128, 4096, 4, 1), (512, 1024, 32, 8), (16, 4096, 32, 8), (2, 32768, 32, 8), (1, 1024, 32, 8),
    (256, 8192, 8, 1),
]
GRID = {"B200ATTN_DECODE_WAVES": ["1", "4", "8", "16"], "B200ATTN_DECODE_DYN_MIN": ["2", "4"]}
if os.environ.get("B200_SWEEP_GRID"):  # e.g. '{"B200_AB_PKG": ["", "_ab_old"]}'
    GRID = json.loads(os.environ["B200_SWEEP_GRID"])


def child():
    """One combination: ms per launch under graph replay over independent cache sets (bench.py's
    extra_decode harness: back-to-back launches like the layers of a model step)."""
    import torch

    import bench

    if os.environ.get("B200_AB_PKG"):  # bench.py put the repo root first again
        sys.path.insert(0, str(ROOT / os.environ["B200_AB_PKG"]))
    import byox_vllm_b200

    print("package:", byox_vllm_b200.__file__, flush=True)
    res = []
    with torch.inference_mode():
        for (B, ctx, Hq, Hkv) in SHAPES:
            kw = {}
            if os.environ.get("B200_OVERLAP") in ("0", "1") and not os.environ.get("B200_AB_PKG"):
                kw["overlap"] = os.environ["B200_OVERLAP"] == "1"
            r = bench.extra_decode("sweep", B, ctx, Hq, Hkv, ref=False, **kw)
            res.append(dict(B=B, ctx=ctx, Hq=Hq, Hkv=Hkv, ctas=r["decode_ctas"],
                            us=r["ms_per_launch"] * 1e3, gbs=r["achieved"], frac=r["frac"]))
    print("RESULT " + json.dumps(res), flush=True)


def main():
    if "--child" in sys.argv:
        return child()
    out = ROOT / "gpurun_out" / os.environ.get("B200_SWEEP_OUT", "decode_tunables.jsonl")
    out.parent.mkdir(exist_ok=True)
    keys = list(GRID)
    with open(out, "w") as f:
        for combo in itertools.product(*(GRID[k] for k in keys)):
            env = dict(os.environ, **dict(zip(keys, combo)))
            r = subprocess.run([sys.executable, __file__, "--child"], env=env, capture_output=True,
                               text=True, timeout=900)
            line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            if r.returncode != 0 or not line:
                print("FAILED", combo, r.stderr[-1500:], flush=True)
                continue
            print([l for l in r.stdout.splitlines() if l.startswith("package:")], flush=True)
            res = json.loads(line[-1][7:])
            f.write(json.dumps(dict(zip(keys, combo), results=res)) + "\n")
            f.flush()
            print(dict(zip(keys, combo)), " ".join(
                f"{x['B']}x{x['ctx']}/{x['Hkv']}:{x['us']:.1f}us({x['frac']:.2f})" for x in res),
                flush=True)


if __name__ == "__main__":
    main()
