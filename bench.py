#!/usr/bin/env python
"""bench.py — decode tokens/s of the attention hot path on the reference's bench shape.

One "step" = one decode step of the whole batch through the attention path of every layer:
for each layer, `Attention.forward(q, k, v)` = scatter of this step's K/V rows into the paged
cache + paged decode attention over block_tables/context_lens (reference
nanovllm/layers/attention.py:67-97), captured once in a CUDA graph and replayed ("CUDA graphs
on", as ModelRunner does: engine/model_runner.py:242-297). tokens/s = batch / step time.

Workloads (BASELINE.json configs):
  c2  (default) Qwen3-0.6B heads 16/8, 28 layers, 256 sequences whose context lengths are the
      reference bench's prompt + half of its output lengths (bench.py:17-36 RNG replayed).
  c3  Qwen3-8B heads 32/8, 36 layers, batch 128 x ctx 4096.
N > 1 (torchrun): tensor-parallel layout of the reference — every rank gets Hq/N and Hkv/N heads
of every layer and the full (replicated) metadata; no data-path collective; strong scaling.

`--impl reference` times the CPU restatement of the same path (oracle/, the reference has no
CPU attention path of its own and is pure Python + CUDA-only wheels) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

WORKLOADS = {
    # name: (layers, Hq, Hkv, batch, ctx spec)
    "c2": dict(model="Qwen3-0.6B", layers=28, Hq=16, Hkv=8, batch=256, ctx="bench"),
    "c3": dict(model="Qwen3-8B", layers=36, Hq=32, Hkv=8, batch=128, ctx=4096),
}
D, BLOCK = 128, 256


def context_lengths(w) -> list[int]:
    from tools import synth

    if w["ctx"] == "bench":
        prompts, outputs = synth.bench_shape_lengths(w["batch"], 100, 1024, seed=0)
        return [p + o // 2 for p, o in zip(prompts, outputs)]
    return [int(w["ctx"])] * w["batch"]


def algorithmic_bytes(lens, Hq, Hkv) -> int:
    """SURVEY.md §8(d) a2, per launch (one layer): K+V rows once per KV head + q in + out +
    block-table entries + context_lens."""
    toks = sum(lens)
    pages = sum((c + BLOCK - 1) // BLOCK for c in lens)
    B = len(lens)
    return toks * Hkv * D * 2 * 2 + 2 * B * Hq * D * 2 + 4 * pages + 4 * B


def measured_traffic(workload: str, world: int):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the decode kernel from the
    committed `ncu --set full` capture of this workload (profiles/traffic.json), else None."""
    p = ROOT / "profiles" / "traffic.json"
    if not p.exists():
        return None
    return json.loads(p.read_text()).get(f"{workload}_tp{world}")


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return json.loads(p.read_text()), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(index), "-lms", "100"], stdout=subprocess.PIPE,
                stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        out, _ = self.proc.communicate(timeout=10)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def build_state(w, shard, lens, dev):
    """Per-rank device state: KV cache [2, L, nb, 256, Hkv_local, 128], metadata, Attention
    modules with their cache views injected like ModelRunner._allocate_kv_cache does."""
    import random

    from byox_vllm_b200 import metadata as md
    from byox_vllm_b200.attention import Attention

    L, Hq, Hkv = w["layers"], shard.num_q_heads, shard.num_kv_heads
    rng = random.Random(0)
    need = [(c + BLOCK - 1) // BLOCK for c in lens]
    nb = sum(need) + 8
    ids = list(range(nb))
    rng.shuffle(ids)
    seqs, at = [], 0
    for c, n in zip(lens, need):
        seqs.append(md.SeqState(c, ids[at:at + n]))
        at += n
    meta = md.prepare_decode(seqs, BLOCK, pin=True)
    torch.manual_seed(1234)
    kv_cache = torch.empty(2, L, nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device=dev)
    for layer in range(L):
        kv_cache[:, layer].normal_()
    layers = []
    for layer in range(L):
        a = Attention(Hq, D, D ** -0.5, Hkv)
        a.k_cache, a.v_cache = kv_cache[0, layer], kv_cache[1, layer]
        layers.append(a)
    B = len(lens)
    # pinned host inputs of one step: per layer q [B,Hq,D] and the fused qkv rows for k / v
    host_q = torch.randn(L, B, Hq, D).to(torch.bfloat16).pin_memory()
    host_kv = torch.randn(L, B, 2 * Hkv * D).to(torch.bfloat16).pin_memory()
    return dict(meta=meta, kv_cache=kv_cache, layers=layers, host_q=host_q, host_kv=host_kv,
                seqs=seqs, nb=nb)


def run_ours(args, w, rank, world, dev):
    from byox_vllm_b200 import _lib, metadata as md, parallel
    from byox_vllm_b200.attention import paged_decode
    from byox_vllm_b200.context import reset_context, set_context

    lib = _lib.load()
    shard = parallel.shard_heads(w["Hq"], w["Hkv"], world, rank)
    Hq, Hkv, L = shard.num_q_heads, shard.num_kv_heads, w["layers"]
    lens = context_lengths(w)
    B = len(lens)
    st = build_state(w, shard, lens, dev)
    meta, layers = st["meta"], st["layers"]
    dq = st["host_q"].to(dev)
    dkv = st["host_kv"].to(dev)
    dmeta = md.context_kwargs(meta, dev)

    def step_device():
        o = None
        for layer in range(L):
            kv = dkv[layer]
            k = kv[:, :Hkv * D].view(B, Hkv, D)
            v = kv[:, Hkv * D:].view(B, Hkv, D)
            o = layers[layer](dq[layer], k, v)  # consumed by o_proj in the model
        return o

    set_context(**dmeta)
    step_device()  # eager warm-up (descriptor cache, func attributes) before capture
    torch.cuda.synchronize()
    n0 = lib.b200attn_launch_count()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step_device()
    launches_per_step = lib.b200attn_launch_count() - n0
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        graph.replay()
    barrier()
    sampler = ClockSampler(dev.index or 0) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        graph.replay()
    e1.record()
    barrier()
    ms_step = parallel.max_over_ranks(e0.elapsed_time(e1) / args.steps)

    # ---- roofline leg: the decode kernel alone, one launch per layer, CUDA events on the stream
    splits = lib.b200attn_decode_pick_splits(B, Hkv, dmeta["block_tables"].size(1), BLOCK,
                                             lib.b200attn_sm_count())
    scratch = torch.empty(B, Hq, D, dtype=torch.bfloat16, device=dev)

    def decode_only():
        for layer in range(L):
            paged_decode(dq[layer], layers[layer].k_cache, layers[layer].v_cache,
                         dmeta["context_lens"], dmeta["block_tables"], D ** -0.5, out=scratch)

    decode_only()
    torch.cuda.synchronize()
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = max(2, min(args.steps, 10))
    k0.record()
    for _ in range(reps):
        decode_only()
    k1.record()
    torch.cuda.synchronize()
    ms_kernel = k0.elapsed_time(k1) / (reps * L)
    clocks = sampler.stop() if sampler else None

    # ---- e2e leg: the user-facing call with HOST buffers; H2D of metadata + q/k/v and D2H of
    # the outputs inside the timed region, every step
    host_out = torch.empty(L, B, 1, Hq, D, dtype=torch.bfloat16).pin_memory()
    h2d = d2h = 0
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    ev_in = [torch.cuda.Event() for _ in range(L)]
    ev_out = [torch.cuda.Event() for _ in range(L)]
    sq = torch.empty_like(dq)    # device staging of the host inputs, one slice per layer
    skv = torch.empty_like(dkv)

    def step_e2e():
        """One decode step from HOST buffers: metadata + every layer's q/k/v go pinned-host ->
        device on a copy stream, each layer's Attention.forward waits only for its own inputs, and
        its output goes device -> pinned-host on a second copy stream, so PCIe traffic overlaps
        the kernels. The step ends when the last output is on the host."""
        nonlocal h2d, d2h
        h2d = d2h = 0
        cur = torch.cuda.current_stream(dev)
        m = md.context_kwargs(meta, dev)  # pinned -> device, non_blocking (model_runner.py:217-225)
        set_context(**m)
        for t in (m["slot_mapping"], m["context_lens"], m["block_tables"]):
            h2d += t.numel() * t.element_size()
        s_in.wait_stream(cur)
        with torch.cuda.stream(s_in):
            for layer in range(L):
                sq[layer].copy_(st["host_q"][layer], non_blocking=True)
                skv[layer].copy_(st["host_kv"][layer], non_blocking=True)
                ev_in[layer].record(s_in)
                h2d += sq[layer].numel() * 2 + skv[layer].numel() * 2
        outs = []
        for layer in range(L):
            cur.wait_event(ev_in[layer])
            kv = skv[layer]
            k = kv[:, :Hkv * D].view(B, Hkv, D)
            v = kv[:, Hkv * D:].view(B, Hkv, D)
            o = layers[layer](sq[layer], k, v)
            ev_out[layer].record(cur)
            outs.append(o)  # keep alive until its D2H copy has run
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_out[layer])
                host_out[layer].copy_(o, non_blocking=True)
            d2h += o.numel() * 2
        s_out.synchronize()  # the step's results are on the host
        cur.synchronize()

    for _ in range(3):
        step_e2e()
    barrier()
    e2e_steps = max(3, min(args.steps, 20))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    barrier()
    ms_e2e = parallel.max_over_ranks((time.perf_counter() - t0) * 1e3 / e2e_steps)
    reset_context()

    if rank != 0:
        return None
    pk, pk_kind = peaks()
    alg = algorithmic_bytes(lens, Hq, Hkv)
    achieved = alg / (ms_kernel * 1e-3) / 1e9
    return {
        "metric": "decode tokens/sec (bench.py shape)", "value": B / (ms_step * 1e-3),
        "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {
            "workload": f"{args.workload}: {w['model']} attention path, {L} layers, batch {B}, "
                        f"ctx {min(lens)}-{max(lens)} (mean {sum(lens) // B}), heads "
                        f"{w['Hq']}/{w['Hkv']} over TP={world} -> {Hq}/{Hkv} per rank, "
                        f"block_size {BLOCK}, CUDA graph replay",
            "l2": f"inputs larger than L2: {alg * L / 1e9:.2f} GB of KV streamed per step",
            "parallelism": f"tp{world}", "decode_splits": int(splits),
        },
        "roofline": {
            "bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
            "frac": achieved / pk["hbm_gbs"],
            "traffic": measured_traffic(args.workload, world), "peak_kind": pk_kind,
            "kernel": "decode_split_kernel (+ merge when splits > 1)",
            "algorithmic_bytes_per_launch": alg, "ms_per_launch": ms_kernel,
            "frac_of_8TBs_nominal": achieved / 8000.0,
        },
        "e2e": {"value": B / (ms_e2e * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e,
                "path": "Attention.forward per layer from pinned HOST q/k/v (copy stream) with outputs "
                        "back to pinned host (second copy stream); H2D/D2H inside the timed region"},
        "gpu_launches": int(launches_per_step * args.steps),
        "launches_per_step": int(launches_per_step),
        "clocks": clocks,
    }


def cpu_reference(w, steps, warmup, sample_seqs=None, budget_s=20.0):
    """The CPU restatement (oracle/) of store_kvcache + paged decode, timed on the host cores on a
    bounded sample: ONE layer of the workload over the first `sample_seqs` sequences; tokens/s is
    scaled by the layer count (layers are identical work)."""
    from oracle import attention_oracle as orc  # timed CPU baseline leg
    from tools import synth

    lens = context_lengths(w)
    n = min(sample_seqs or len(lens), len(lens))
    case = synth.decode_case(n, lens[:n], w["Hq"], w["Hkv"], seed=0)
    ctx = type("Ctx", (), dict(case.meta))()

    def one():
        orc.attention_forward_ref(case.q, case.k, case.v, case.k_cache, case.v_cache, ctx,
                                  case.scale)

    for _ in range(max(1, warmup)):
        one()
    times = []
    t_begin = time.perf_counter()
    for _ in range(steps):
        t0 = time.perf_counter()
        one()
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_begin > budget_s:
            break
    t = statistics.median(times)
    tok_s = n / (t * w["layers"])
    return tok_s, t, n, len(times)


def run_reference(args, w, rank, world):
    if rank != 0:
        return None
    # torchrun exports OMP_NUM_THREADS=1; the reference arm uses every host thread it can.
    torch.set_num_threads(len(os.sched_getaffinity(0)))
    tok_s, t, n, done = cpu_reference(w, args.steps, max(args.warmup, 1), budget_s=120.0)
    cores = torch.get_num_threads()
    sample = (f"1 of {w['layers']} layers x {n} of {w['batch']} sequences (store + paged decode, "
              f"fp32 torch on CPU), {done} timed passes, tokens/s scaled by the layer count")
    return {
        "impl": "reference", "metric": "decode tokens/sec (bench.py shape)", "value": tok_s,
        "unit": "tokens/s", "n_gpus": world, "steps": done, "warmup": max(args.warmup, 1),
        "ms_per_step": t * w["layers"] * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {w['model']} attention path on host cores",
                   "parallelism": "cpu"},
        "cpu_baseline": {"value": tok_s, "unit": "tokens/s", "cores": cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": tok_s, "unit": "tokens/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    args = ap.parse_args()
    w = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        line = run_reference(args, w, rank, world)
        if line is not None:
            print(json.dumps(line), flush=True)
        return

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.distributed.init_process_group("nccl", device_id=dev)
    try:
        with torch.inference_mode():
            line = run_ours(args, w, rank, world, dev)
        if rank == 0 and world > 1:
            line["cpu_baseline"] = None  # timed at N=1 only
            print(json.dumps(line), flush=True)
        elif rank == 0:
            tok_s, t, n, done = cpu_reference(w, 5, 1, sample_seqs=64, budget_s=20.0)
            line["cpu_baseline"] = {
                "value": tok_s, "unit": "tokens/s", "cores": torch.get_num_threads(),
                "kind": "port",
                "sample": f"1 of {w['layers']} layers x {n} of {w['batch']} sequences, {done} "
                          f"passes, median {t * 1e3:.1f} ms/layer, scaled by the layer count"}
            print(json.dumps(line), flush=True)
    finally:
        if world > 1:
            torch.distributed.barrier()
            torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
