"""Short driver for ncu captures of the shared-prefix cascade (BASELINE config 4 per rank):

    python tools/prof_cascade.py [B ctx shared Hq Hkv] [--chunks=N] [--iters=N]

Runs plan + (prefix pass, decode) a few times with the L2 flushed in between; prints CUDA-event
times. Meant to be wrapped by `ncu -k regex:prefill_tc|decode_split|cascade_plan|decode_merge`.
"""
from __future__ import annotations

import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from tools import microbench as mb  # noqa: E402


def main():
    pos = [int(a) for a in sys.argv[1:] if not a.startswith("--")]
    B, ctx, shared, Hq, Hkv = pos if len(pos) == 5 else (256, 8192, 7680, 8, 1)
    opt = dict(a[2:].split("=", 1) for a in sys.argv[1:] if a.startswith("--") and "=" in a)
    from byox_vllm_b200.attention import cascade_chunks, cascade_plan, paged_decode

    q, kc, vc, lens, bt, unique, logical, flops = mb.cascade_inputs(B, ctx, shared, Hq, Hkv)
    chunks = int(opt.get("chunks", cascade_chunks(B, Hq, Hkv, bt.size(1), mb.BLOCK) or 8))
    iters = int(opt.get("iters", 5))
    o = torch.empty_like(q)
    scale = mb.D ** -0.5
    plan = cascade_plan(lens, bt, mb.BLOCK)
    t = mb.timed(lambda: paged_decode(q, kc, vc, lens, bt, scale, out=o, plan=plan,
                                      prefix_chunks=chunks), iters=iters)
    t_plain = mb.timed(lambda: paged_decode(q, kc, vc, lens, bt, scale, out=o), iters=iters)
    print(f"B={B} ctx={ctx} shared={shared} {Hq}/{Hkv} chunks={chunks} plan={plan.cpu().tolist()} "
          f"cascade {t[1] * 1e3:.1f} us (min {t[0] * 1e3:.1f}) plain {t_plain[1] * 1e3:.1f} us")


if __name__ == "__main__":
    main()
