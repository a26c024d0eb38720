"""Logits of the REAL engine (nanovllm from baseline/_ref: scheduler, block manager with prefix
caching, ModelRunner with CUDA graphs, its own metadata builders) on a scripted workload, with the
reference's attention (`--impl ref`) or the B200 drop-in (`--impl ours`, byox_vllm_b200.install
with the fused-qkv / device-metadata paths on or off). tests/test_engine_parity.py runs both arms
in separate processes and compares the logits of every step.

Lock step between the arms: the sampler is replaced (identically in both) by a deterministic token
stream that does not depend on the logits, so prompts, cache hits, batch composition and graph
sizes are the same at every step and the logits can be compared one to one.

Workload (two-layer random-init Qwen3-0.6B geometry, seeded identically):
  1. generate(): 6 prompts of 40..700 tokens, 10 decode steps (prefill + graph decode, ragged);
  2. generate(): 5 prompts that share their first 512 tokens with prompt 0 of pass 1 (prefix-cache
     hits: prefill with block tables, then decode rows that share leading pages - the cascade's
     trigger when it is armed) and 10 decode steps;
  3. generate(): 1 prompt, 6 steps (graph batch size 1, padding rows in the larger graphs).

    python tools/engine_logits.py --impl ours --fuse-qkv 1 --fast-meta 1 --out /tmp/ours.pt
"""
from __future__ import annotations

import argparse
import os
import sys
from pathlib import Path
from random import Random

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "baseline" / "_ref"))
sys.path.insert(0, str(ROOT / "tools"))
os.environ.setdefault("TOKENIZERS_PARALLELISM", "false")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", choices=["ours", "ref"], required=True)
    ap.add_argument("--fuse-qkv", default="1")
    ap.add_argument("--fast-meta", default="1")
    ap.add_argument("--eager", action="store_true")
    ap.add_argument("--dtype", default="bfloat16", choices=["bfloat16", "float16"])
    ap.add_argument("--out", required=True)
    args = ap.parse_args()
    os.environ["B200ATTN_FUSE_QKV"] = args.fuse_qkv
    os.environ["B200ATTN_FAST_META"] = args.fast_meta

    import e2e_nanovllm as harness

    harness.patch_reference(args.impl)
    import logging

    import torch
    import nanovllm.engine.model_runner as mr
    from nanovllm import LLM, SamplingParams, set_global_log_level
    from nanovllm.layers.sampler import Sampler

    set_global_log_level(logging.WARNING)

    # deterministic, logits-independent "sampling": the same token stream in both arms
    counter = {"n": 0}

    def fixed_sampler(self, logits, temperatures):
        counter["n"] += 1
        n = logits.size(0)
        return (torch.arange(n, device=logits.device) * 7919 + counter["n"] * 104729) % 20000

    Sampler.forward = fixed_sampler

    record = []
    orig_run_model = mr.ModelRunner._run_model

    def run_model(self, input_ids, positions, is_prefill):
        logits = orig_run_model(self, input_ids, positions, is_prefill)
        record.append((bool(is_prefill), int(input_ids.numel()), logits.float().cpu()))
        return logits

    mr.ModelRunner._run_model = run_model

    path = harness.make_model_dir("tiny", args.dtype)
    llm = LLM(path, enforce_eager=args.eager, max_model_len=2048, gpu_memory_utilization=0.3,
              tensor_parallel_size=1, max_num_batched_tokens=4096, max_num_seqs=16)
    record.clear()
    counter["n"] = 0

    rng = Random(11)
    p1 = [[rng.randint(0, 19999) for _ in range(n)] for n in (700, 40, 257, 512, 300, 96)]
    sp = lambda n: SamplingParams(temperature=0.6, ignore_eos=True, max_tokens=n)  # noqa: E731
    llm.generate(p1, sp(10), use_tqdm=False)
    head = p1[0][:512]
    p2 = [head + [rng.randint(0, 19999) for _ in range(n)] for n in (30, 200, 1, 64, 129)]
    llm.generate(p2, sp(10), use_tqdm=False)
    llm.generate([[rng.randint(0, 19999) for _ in range(77)]], sp(6), use_tqdm=False)

    info = {"impl": args.impl, "fuse_qkv": args.fuse_qkv, "fast_meta": args.fast_meta,
            "eager": args.eager, "dtype": args.dtype, "steps": len(record),
            "prefill_steps": sum(1 for r in record if r[0])}
    if args.impl == "ours":
        from byox_vllm_b200 import _lib

        info["libb200attn_launches"] = int(_lib.load().b200attn_launch_count())
    keep = torch.bfloat16 if args.dtype == "bfloat16" else torch.float16
    torch.save({"info": info, "steps": [(p, n, l.to(keep)) for p, n, l in record]},
               args.out)
    print(info, flush=True)


if __name__ == "__main__":
    main()
