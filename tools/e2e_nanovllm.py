"""End-to-end run of the REFERENCE ENGINE (nanovllm, unmodified, from baseline/_ref) with either its
own attention (Triton store + flash-attn: `--impl ref`) or the B200 drop-in (`--impl ours`,
byox_vllm_b200.install) — same scheduler, block manager, model runner, CUDA graphs, sampler.

    gpurun -- python tools/e2e_nanovllm.py --impl ours --model 0.6b
    gpurun -- python tools/e2e_nanovllm.py --impl ref  --model 0.6b

Workload = the reference's bench (bench.py:17-42): seed(0), N prompts of randint(100, max_in) random
token ids, max_tokens = randint(100, max_out), temperature 0.6, ignore_eos; metric = sum(max_tokens)
/ wall time of generate() (prefill included). BASELINE.json config 2: 256 seqs, 100-1024 / 100-1024.

No weights and no network exist here, so (identically for both arms):
  * the model directory is synthetic: config.json with the public Qwen3 dimensions + a WordLevel
    tokenizer; parameters are seeded random (the reference's load_model finds no *.safetensors);
  * transformers 5.x hands `rope_scaling` as a dict: get_rope is wrapped to pass None for the
    default rope type (reference rotary_embedding.py:51-59 asserts None).
Prints one JSON line; also checks (`--check`) that both arms produce the same greedy-ish tokens on
a short prompt set.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time
from pathlib import Path
from random import randint, seed

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "baseline" / "_ref"))
os.environ.setdefault("TOKENIZERS_PARALLELISM", "false")

MODELS = {
    # two layers of the 0.6B geometry with a small vocabulary: engine-level parity tests
    "tiny": dict(hidden_size=1024, num_hidden_layers=2, num_attention_heads=16,
                 num_key_value_heads=8, intermediate_size=3072, tie_word_embeddings=True,
                 vocab_size=32000),
    "0.6b": dict(hidden_size=1024, num_hidden_layers=28, num_attention_heads=16,
                 num_key_value_heads=8, intermediate_size=3072, tie_word_embeddings=True),
    "8b": dict(hidden_size=4096, num_hidden_layers=36, num_attention_heads=32,
               num_key_value_heads=8, intermediate_size=12288, tie_word_embeddings=False),
    "32b": dict(hidden_size=5120, num_hidden_layers=64, num_attention_heads=64,
                num_key_value_heads=8, intermediate_size=25600, tie_word_embeddings=False),
}


def make_model_dir(name: str, dtype: str = "bfloat16") -> str:
    d = Path(tempfile.gettempdir()) / (f"synthetic-qwen3-{name}" +
                                       ("" if dtype == "bfloat16" else f"-{dtype}"))
    d.mkdir(exist_ok=True)
    cfg = dict(architectures=["Qwen3ForCausalLM"], model_type="qwen3", head_dim=128,
               hidden_act="silu", vocab_size=151936, max_position_embeddings=40960,
               rms_norm_eps=1e-6, rope_theta=1000000, attention_bias=False,
               torch_dtype=dtype, dtype=dtype, bos_token_id=151643,
               eos_token_id=151645)
    cfg.update(MODELS[name])
    if cfg["vocab_size"] < 151936:
        cfg.update(bos_token_id=1, eos_token_id=2)
    (d / "config.json").write_text(json.dumps(cfg))
    if not (d / "tokenizer.json").exists():
        from tokenizers import Tokenizer, models, pre_tokenizers

        vocab = {f"t{i}": i for i in range(20000)}
        vocab["<unk>"] = 20000
        vocab["<eos>"] = 20001
        tok = Tokenizer(models.WordLevel(vocab=vocab, unk_token="<unk>"))
        tok.pre_tokenizer = pre_tokenizers.Whitespace()
        tok.save(str(d / "tokenizer.json"))
        (d / "tokenizer_config.json").write_text(json.dumps(
            dict(tokenizer_class="PreTrainedTokenizerFast", eos_token="<eos>",
                 unk_token="<unk>", model_max_length=40960)))
    return str(d)


def patch_reference(impl: str) -> None:
    """Harness shims (synthetic weights, rope_scaling dict) applied identically to both arms, the
    drop-in for `ours`, and - for the reference arm only - the tensor-parallel transport fixes the
    unmodified reference needs to take a step at TP > 1 (ours: byox_vllm_b200.fast_rpc)."""
    if impl == "ours":
        from byox_vllm_b200 import install

        install.install(verbose=True)
    import torch
    import nanovllm.layers.rotary_embedding as rot
    import nanovllm.models.qwen3 as qwen3
    import nanovllm.engine.model_runner as mr

    orig_get_rope = rot.get_rope

    def get_rope(head_size, rotary_dim, max_position, base, rope_scaling=None):
        if isinstance(rope_scaling, dict):
            assert rope_scaling.get("rope_type", "default") == "default"
            rope_scaling = None
        return orig_get_rope(head_size, rotary_dim, max_position, float(base), rope_scaling)

    qwen3.get_rope = get_rope

    def seeded_load_model(model, path):
        rank = torch.distributed.get_rank() if torch.distributed.is_initialized() else 0
        g = torch.Generator(device="cuda").manual_seed(1234 + rank)
        for name, p in model.named_parameters():
            if name.endswith("norm.weight"):
                p.data.fill_(1.0)
            else:
                p.data.normal_(0.0, 0.02, generator=g)

    mr.load_model = seeded_load_model
    if impl == "ours" and os.environ.get("B200ATTN_FAST_RPC", "1") != "0":
        # install(fast_rpc=True) has replaced the transport (compact per-step records in a block
        # sized from the engine limits) and ModelRunner.run: no harness shim on this arm.
        return

    # TP control plane: the reference pickles whole Sequences (all token ids) into
    # a 1 MiB shared-memory block every step, which overflows at the BASELINE batch sizes. The
    # reference arm gets two harness fixes: a larger block, and decode steps ship token-free copies
    # (the workers' _prepare_decode reads only last_token / num_tokens / block_table).
    import copy
    from multiprocessing.shared_memory import SharedMemory as _SharedMemory

    class BigSharedMemory(_SharedMemory):
        def __init__(self, name=None, create=False, size=0):
            super().__init__(name=name, create=create, size=(1 << 27) if create else 0)

    mr.SharedMemory = BigSharedMemory
    orig_write = mr.ModelRunner._write_shm

    def write_shm(self, method_name, *args):
        if method_name == "run" and len(args) == 2 and args[1] is False:
            light = []
            for seq in args[0]:
                c = copy.copy(seq)
                c.token_ids = []
                light.append(c)
            args = (light, False)
        return orig_write(self, method_name, *args)

    mr.ModelRunner._write_shm = write_shm

    # The reference's ModelRunner.run evaluates `logits.shape` inside a debug f-string
    # (model_runner.py:63); logits is None on ranks > 0, so every TP > 1 run of the unmodified
    # reference dies on its first step. Same body without the two debug lines, for both arms.
    def run(self, seqs, is_prefill):
        input_ids, positions = (self._prepare_prefill(seqs) if is_prefill
                                else self._prepare_decode(seqs))
        temperatures = self._prepare_sample(seqs) if self.rank == 0 else None
        logits = self._run_model(input_ids, positions, is_prefill)
        token_ids = self.sampler(logits, temperatures).tolist() if self.rank == 0 else None
        mr.reset_context()
        return token_ids

    mr.ModelRunner.run = run


# Tensor-parallel ranks are spawned (llm_engine.py:24-31) and re-import this file as __mp_main__
# without running main(): they pick the arm up from the environment and patch at import time.
if __name__ != "__main__" and os.environ.get("E2E_NANOVLLM_IMPL"):
    patch_reference(os.environ["E2E_NANOVLLM_IMPL"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", choices=["ours", "ref"], required=True)
    ap.add_argument("--model", choices=sorted(MODELS), default="0.6b")
    ap.add_argument("--num-seqs", type=int, default=256)
    ap.add_argument("--max-in", type=int, default=1024)
    ap.add_argument("--max-out", type=int, default=1024)
    ap.add_argument("--fixed-in", type=int, default=0, help="every prompt has exactly this length")
    ap.add_argument("--fixed-out", type=int, default=0, help="every request decodes this many")
    ap.add_argument("--shared-prefix", type=int, default=0,
                    help="the first N tokens of every prompt are identical (prefix-cache hits: "
                         "BlockManager hands every sequence the same leading block ids)")
    ap.add_argument("--max-model-len", type=int, default=4096)
    ap.add_argument("--max-batched", type=int, default=16384)
    ap.add_argument("--util", type=float, default=0.9)
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--eager", action="store_true")
    ap.add_argument("--check", action="store_true",
                    help="also dump near-greedy tokens of 8 short prompts for cross-arm comparison")
    ap.add_argument("--out", default=None)
    ap.add_argument("--cprofile", action="store_true",
                    help="run the timed generate under cProfile and print the host-side hot spots")
    ap.add_argument("--profile", action="store_true",
                    help="torch.profiler table of a short generate instead of the benchmark")
    args = ap.parse_args()

    os.environ["E2E_NANOVLLM_IMPL"] = args.impl
    patch_reference(args.impl)
    import logging

    import torch
    from nanovllm import LLM, SamplingParams, set_global_log_level

    set_global_log_level(logging.WARNING)
    path = make_model_dir(args.model)
    t0 = time.time()
    llm = LLM(path, enforce_eager=args.eager, max_model_len=args.max_model_len,
              gpu_memory_utilization=args.util, tensor_parallel_size=args.tp,
              max_num_batched_tokens=args.max_batched)
    t_init = time.time() - t0

    # per-phase accounting without touching the engine: wrap LLMEngine._step
    phase = dict(prefill_s=0.0, decode_s=0.0, prefill_tokens=0, decode_tokens=0, decode_steps=0)
    orig_step = llm._step

    def timed_step():
        t = time.perf_counter()
        out, n = orig_step()
        dt = time.perf_counter() - t
        if n > 0:
            phase["prefill_s"] += dt
            phase["prefill_tokens"] += n
        elif n < 0:
            phase["decode_s"] += dt
            phase["decode_tokens"] += -n
            phase["decode_steps"] += 1
        return out, n

    llm._step = timed_step

    seed(0)
    prompts = [[randint(0, 10000) for _ in range(args.fixed_in or randint(100, args.max_in))]
               for _ in range(args.num_seqs)]
    if args.shared_prefix:
        head = [randint(0, 10000) for _ in range(args.shared_prefix)]
        prompts = [head + p[args.shared_prefix:] for p in prompts]
        assert all(len(p) > args.shared_prefix for p in prompts), "prompts must outgrow the prefix"
    sps = [SamplingParams(temperature=0.6, ignore_eos=True,
                          max_tokens=args.fixed_out or randint(100, args.max_out))
           for _ in range(args.num_seqs)]
    llm.generate(["t1 t2 t3"], SamplingParams(), use_tqdm=False)  # warm-up as bench.py:37
    if args.profile:
        from torch.profiler import ProfilerActivity, profile

        sp8 = SamplingParams(temperature=0.6, ignore_eos=True, max_tokens=8)
        llm.generate(prompts[:32], sp8, use_tqdm=False)
        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
            llm.generate(prompts[:32], sp8, use_tqdm=False)
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25,
                                        max_name_column_width=70))
        print(prof.key_averages().table(sort_by="cpu_time_total", row_limit=15,
                                        max_name_column_width=70))
        return
    # Warm pass on DIFFERENT prompts (no prefix-cache hits for the timed run) so torch.compile's
    # dynamic-shape recompiles of the reference's own @torch.compile regions and every lazy kernel
    # load happen before the clock starts — identically for both arms.
    seed(123)
    wp = [[randint(10001, 19999) for _ in range(args.fixed_in or randint(100, args.max_in))]
          for _ in range(8 if args.fixed_in else 48)]
    llm.generate(wp, SamplingParams(temperature=0.6, ignore_eos=True, max_tokens=16),
                 use_tqdm=False)
    for k in phase:
        phase[k] = 0 if isinstance(phase[k], int) else 0.0
    torch.cuda.synchronize()
    t = time.time()
    if args.cprofile:
        import cProfile
        import pstats

        prof = cProfile.Profile()
        prof.runcall(llm.generate, prompts, sps, use_tqdm=False)
        t = time.time() - t
        for key in ("tottime", "cumulative"):
            print(f"==== cProfile by {key} ====", flush=True)
            pstats.Stats(prof).sort_stats(key).print_stats(28)
    else:
        llm.generate(prompts, sps, use_tqdm=False)
        t = time.time() - t
    total = sum(sp.max_tokens for sp in sps)

    line = {
        "tool": "e2e_nanovllm", "impl": args.impl, "model": f"Qwen3-{args.model} (random init)",
        "tp": args.tp, "num_seqs": args.num_seqs, "in": args.fixed_in or [100, args.max_in],
        "out": args.fixed_out or [100, args.max_out], "max_model_len": args.max_model_len,
        "shared_prefix": args.shared_prefix, "cascade_env": os.environ.get("B200ATTN_CASCADE"),
        "fuse_qkv_env": os.environ.get("B200ATTN_FUSE_QKV"),
        "fast_meta_env": os.environ.get("B200ATTN_FAST_META"),
        "fast_rpc_env": os.environ.get("B200ATTN_FAST_RPC"),
        "cuda_graphs": not args.eager, "total_output_tokens": total, "seconds": t,
        "tokens_per_s": total / t, "init_s": t_init,
        "decode_tokens_per_s": phase["decode_tokens"] / max(phase["decode_s"], 1e-9),
        "prefill_tokens_per_s": phase["prefill_tokens"] / max(phase["prefill_s"], 1e-9),
        "decode_steps": phase["decode_steps"], "decode_s": phase["decode_s"],
        "prefill_s": phase["prefill_s"],
    }
    if args.check:
        seed(1)
        cps = [[randint(0, 10000) for _ in range(randint(20, 600))] for _ in range(8)]
        outs = llm.generate(cps, SamplingParams(temperature=1e-5, max_tokens=24, ignore_eos=True),
                            use_tqdm=False)
        line["check_tokens"] = [o["token_ids"] for o in outs]
    if args.impl == "ours":
        from byox_vllm_b200 import _lib

        line["libb200attn_launches"] = int(_lib.load().b200attn_launch_count())
        if args.tp > 1:
            from byox_vllm_b200 import fast_rpc

            st = fast_rpc.stats()
            line["rpc_calls"] = st["steps"]
            line["rpc_bytes_per_call"] = st["bytes"] / max(st["steps"], 1)
            line["rpc_host_us_per_call"] = st.get("host_s", 0.0) * 1e6 / max(st["steps"], 1)
    print(json.dumps(line), flush=True)
    if args.out:
        Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        Path(args.out).write_text(json.dumps(line) + "\n")


if __name__ == "__main__":
    main()
