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

What one line carries besides the contract's keys:
  roofline      the decode kernel alone (CUDA events), algorithmic bytes / measured HBM peak
  e2e           the same step from pinned HOST buffers: one slab per layer group host->device,
                the group's captured graph, its outputs device->host; copies inside the timing
  flash_attn    the reference's OWN attention module (nanovllm/layers/attention.py from
                baseline/_ref: Triton store_kvcache + flash_attn_with_kvcache, flash-attn's own
                split heuristic) on the SAME per-rank state, captured in a CUDA graph the same
                way and timed after ours: the bar north_star names. `vs_flash_attn` =
                flash-attn step time / ours. A comparison leg: never inside ours' timed region;
                the product package never imports flash_attn.
  extra         (N=1) kernel-level rooflines the headline shape does not show: c3 = batch 128 x
                ctx 4096 (the "batch >= 64, ctx >= 4k" target), c4 per-rank shape with the shared
                prefix (cascade), c5 corners, the TP=8 per-rank shape of c2, varlen prefill
                (tensor pipe), store_kvcache (HBM), the fused q/k norm + RoPE + store pass (HBM),
                each beside flash-attn / the reference's Triton scatter / its norm + RoPE chain.
  engine_e2e    (N=1) the unmodified reference engine (baseline/_ref) on BASELINE config 2 end to
                end (tools/e2e_nanovllm.py): reference attention vs the drop-in at the frozen
                Attention.forward boundary vs the drop-in with every (f) row enabled.
  cpu_baseline  the CPU restatement (oracle/) on a bounded sample, same sample as --impl reference

`--impl reference` times the CPU restatement of the same path (oracle/, the reference has no
CPU attention path of its own and is pure Python + CUDA-only wheels) on the host cores.
`--workload prefill|store|c4|c5|tp8|qkv` print the corresponding extra block alone.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
import types
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
T_START = time.time()

import torch  # noqa: E402

WORKLOADS = {
    # name: (layers, Hq, Hkv, batch, ctx spec)
    "c2": dict(model="Qwen3-0.6B", layers=28, Hq=16, Hkv=8, batch=256, ctx="bench"),
    "c3": dict(model="Qwen3-8B", layers=36, Hq=32, Hkv=8, batch=128, ctx=4096),
}
EXTRA_ONLY = ("prefill", "store", "c4", "c5", "tp8", "qkv")
D, BLOCK = 128, 256
# layers per host->device slab / captured graph of the e2e leg. Measured at N=1: 4 -> e2e/value
# 0.879, 2 -> 0.895, 1 -> 0.869 (below 2 the per-group host work shows). Under tensor parallelism
# a rank's group is N times shorter on the GPU while the host work per group stays: 4 there.
GROUP_LAYERS = int(os.environ.get("B200_BENCH_GROUP", "0"))
CPU_SAMPLE_SEQS = 64      # sequences of the CPU legs' sample (both legs use the same one)


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


def measured_traffic(key: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the kernel from the COMMITTED
    `ncu --set full` capture named in profiles/traffic.json (not re-measured by this run)."""
    p = ROOT / "profiles" / "traffic.json"
    if not p.exists():
        return None, None
    v = json.loads(p.read_text()).get(key)
    if isinstance(v, dict):
        return v.get("bytes"), v.get("source")
    return v, ("profiles/traffic.json" if v is not None else None)


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


# ------------------------------------------------------------------------------------------------
# the reference's own attention module, loaded from baseline/_ref without the rest of the engine
# ------------------------------------------------------------------------------------------------
def load_reference_attention():
    """((nanovllm.layers.attention, nanovllm.utils.context), None) of the UNMODIFIED reference
    under baseline/_ref, or (None, reason). Only the two files of the hot path are executed:
    `nanovllm` / `nanovllm.layers` / `nanovllm.utils` are registered as bare packages so that
    nanovllm/__init__.py (the whole engine, transformers, ...) is not imported."""
    ref = ROOT / "baseline" / "_ref" / "nanovllm"
    if not (ref / "layers" / "attention.py").exists():
        return None, ("baseline/_ref is not installed (run __graft_entry__.build() where "
                      "/root/reference exists)")
    try:
        import importlib

        if "nanovllm" not in sys.modules:
            for name, sub in (("nanovllm", ref), ("nanovllm.layers", ref / "layers"),
                              ("nanovllm.utils", ref / "utils")):
                pkg = types.ModuleType(name)
                pkg.__path__ = [str(sub)]
                sys.modules[name] = pkg
        ctx = importlib.import_module("nanovllm.utils.context")
        att = importlib.import_module("nanovllm.layers.attention")
        return (att, ctx), None
    except Exception as e:  # flash_attn / triton missing on this box
        return None, f"reference attention not importable: {type(e).__name__}: {e}"


def flash_attn_version():
    try:
        import flash_attn

        return getattr(flash_attn, "__version__", "unknown")
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------
# per-rank state
# ------------------------------------------------------------------------------------------------
def build_state(w, shard, lens, dev):
    """Per-rank device state: KV cache [2, L, nb, 256, Hkv_local, 128], metadata, Attention
    modules with their cache views injected like ModelRunner._allocate_kv_cache does, and ONE
    pinned host slab with every layer's inputs: layer l = q [B,Hq,D] | fused k|v rows [B,2*Hkv*D]."""
    import random

    from byox_vllm_b200.attention import Attention
    from tools import metadata_ref as md

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
    row = B * (Hq + 2 * Hkv) * D
    host_slab = torch.randn(L, row).to(torch.bfloat16).pin_memory()
    return dict(meta=meta, kv_cache=kv_cache, layers=layers, host_slab=host_slab, seqs=seqs,
                nb=nb)


def slab_views(slab_layer, B, Hq, Hkv):
    """q [B,Hq,D] contiguous, k [B,Hkv,D] / v [B,Hkv,D] strided views of the fused k|v rows (as
    qkv.split gives the reference: models/qwen3.py:76-79)."""
    q = slab_layer[:B * Hq * D].view(B, Hq, D)
    kv = slab_layer[B * Hq * D:].view(B, 2 * Hkv * D)
    return q, kv[:, :Hkv * D].view(B, Hkv, D), kv[:, Hkv * D:].view(B, Hkv, D)


def capture(fn):
    """Eager warm-up (lazy module state, descriptor caches, Triton JIT), then capture."""
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    torch.cuda.synchronize()
    return g


def capture_out(fn):
    """capture(), keeping the tensor fn returns (it lives in the graph's pool: replay refills it)."""
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        out = fn()
    torch.cuda.synchronize()
    return g, out


def time_replays(graph, steps, warmup, barrier):
    for _ in range(warmup):
        graph.replay()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        graph.replay()
    e1.record()
    barrier()
    return e0.elapsed_time(e1) / steps


def run_ours(args, w, rank, world, dev):
    from byox_vllm_b200 import _lib, parallel
    from byox_vllm_b200.attention import paged_decode
    from byox_vllm_b200.context import reset_context, set_context
    from tools import metadata_ref as md

    lib = _lib.load()
    # as under install(): decode launches are programmatic dependents of the kernel before them
    # (metadata is written once per step, never right before an attention call)
    from byox_vllm_b200.attention import set_decode_overlap

    set_decode_overlap(os.environ.get("B200_BENCH_OVERLAP", "1") != "0")
    shard = parallel.shard_heads(w["Hq"], w["Hkv"], world, rank)
    Hq, Hkv, L = shard.num_q_heads, shard.num_kv_heads, w["layers"]
    lens = context_lengths(w)
    B = len(lens)
    st = build_state(w, shard, lens, dev)
    meta, layers = st["meta"], st["layers"]
    dslab = st["host_slab"].to(dev)
    dmeta = md.context_kwargs(meta, dev)
    warmup = max(args.warmup, 3)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def step_device():
        o = None
        for layer in range(L):
            q, k, v = slab_views(dslab[layer], B, Hq, Hkv)
            o = layers[layer](q, k, v)  # consumed by o_proj in the model
        return o

    set_context(**dmeta)
    step_device()  # eager warm-up (descriptor cache, func attributes, counters) before capture
    torch.cuda.synchronize()
    n0 = lib.b200attn_launch_count()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        last_out = step_device()
    launches_per_step = lib.b200attn_launch_count() - n0
    torch.cuda.synchronize()

    sampler = ClockSampler(dev.index or 0) if rank == 0 else None
    ms_step = parallel.max_over_ranks(time_replays(graph, args.steps, warmup, barrier))
    ours_last = last_out.clone()  # last layer's output of the timed graph, for the parity check

    # ---- roofline leg: the decode kernel alone, one launch per layer, CUDA events on the stream
    n_ctas = lib.b200attn_decode_pick_splits(B, Hkv, dmeta["block_tables"].size(1), BLOCK,
                                             lib.b200attn_sm_count())
    scratch = torch.empty(B, Hq, D, dtype=torch.bfloat16, device=dev)

    def decode_only():
        for layer in range(L):
            q, _, _ = slab_views(dslab[layer], B, Hq, Hkv)
            paged_decode(q, layers[layer].k_cache, layers[layer].v_cache,
                         dmeta["context_lens"], dmeta["block_tables"], D ** -0.5, out=scratch)

    g_dec = capture(decode_only)
    reps = max(2, min(args.steps, 10))
    ms_kernel = time_replays(g_dec, reps, 2, torch.cuda.synchronize) / L
    clocks = sampler.stop() if sampler else None

    # ---- the same step with overlap mode OFF (plain stream-ordered launches). In the engine the
    # kernel before a decode launch is the qkv GEMM and the one after it (o_proj) is launched
    # without the attribute, so the attention-only chain above overstates what PDL buys there;
    # this is the conservative figure.
    overlap_was_on = set_decode_overlap(False)
    ms_step_plain = ms_step
    if overlap_was_on:
        g_plain = capture(step_device)
        ms_step_plain = parallel.max_over_ranks(time_replays(g_plain, args.steps, warmup, barrier))
        del g_plain
    set_decode_overlap(overlap_was_on)

    # ---- e2e leg: the user-facing call with HOST buffers. Per step: metadata H2D, then per group
    # of GROUP_LAYERS layers ONE pinned slab H2D on a copy stream, the group's captured graph
    # (Attention.forward per layer + its output gathered into an output slab) once its slab has
    # landed, ONE D2H of the group's outputs on a second copy stream. The step ends when the last
    # output is on the host.
    group_layers = GROUP_LAYERS or (2 if world == 1 else 4)
    groups = [(g0, min(g0 + group_layers, L)) for g0 in range(0, L, group_layers)]
    eslab = torch.empty_like(dslab)                       # device staging of the host inputs
    dout = torch.empty(L, B, 1, Hq, D, dtype=torch.bfloat16, device=dev)
    host_out = torch.empty(L, B, 1, Hq, D, dtype=torch.bfloat16).pin_memory()
    emeta = {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in dmeta.items()}
    set_context(**emeta)  # static metadata buffers: the graphs read them, every step refills them

    def group_fn(g0, g1):
        def fn():
            for layer in range(g0, g1):
                q, k, v = slab_views(eslab[layer], B, Hq, Hkv)
                dout[layer].copy_(layers[layer](q, k, v))
        return fn

    eslab.copy_(dslab)
    ggraphs = [capture(group_fn(g0, g1)) for g0, g1 in groups]
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    ev_in = [torch.cuda.Event() for _ in groups]
    ev_out = [torch.cuda.Event() for _ in groups]
    meta_keys = ("slot_mapping", "context_lens", "block_tables")
    h2d = sum(meta[k].numel() * meta[k].element_size() for k in meta_keys) + \
        st["host_slab"].numel() * 2
    d2h = host_out.numel() * 2

    def step_e2e():
        cur = torch.cuda.current_stream(dev)
        for k in meta_keys:  # pinned -> device, non_blocking (model_runner.py:217-225)
            emeta[k].copy_(meta[k], non_blocking=True)
        s_in.wait_stream(cur)
        with torch.cuda.stream(s_in):
            for i, (g0, g1) in enumerate(groups):
                eslab[g0:g1].copy_(st["host_slab"][g0:g1], non_blocking=True)
                ev_in[i].record(s_in)
        for i, (g0, g1) in enumerate(groups):
            cur.wait_event(ev_in[i])
            ggraphs[i].replay()
            ev_out[i].record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_out[i])
                host_out[g0:g1].copy_(dout[g0:g1], non_blocking=True)
        s_out.synchronize()  # the step's results are on the host

    for _ in range(3):
        step_e2e()
    barrier()
    e2e_steps = max(3, min(args.steps, 20))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    barrier()
    ms_e2e = parallel.max_over_ranks((time.perf_counter() - t0) * 1e3 / e2e_steps)
    set_context(**dmeta)

    # ---- flash-attn leg: the reference's own Attention module on the same per-rank state
    fa = flash_attn_leg(args, st, dslab, dmeta, B, Hq, Hkv, L, warmup, barrier)
    reset_context()
    # parity of the timed step itself: last layer's output of our graph vs the reference module's
    # on the same state (north_star tolerance 2e-2 + 1e-2 |ref|), worst rank
    parity = None
    fa_last = fa.pop("_last_out", None)
    if fa_last is not None:
        diff = (ours_last.view(B, Hq, D).float() - fa_last.float()).abs()
        excess = (diff - (2e-2 + 1e-2 * fa_last.float().abs())).max().item()
        parity = {"max_abs_diff": parallel.max_over_ranks(diff.max().item()),
                  "within_tolerance": parallel.max_over_ranks(excess) <= 0.0,
                  "tolerance": "2e-2 + 1e-2*|ref|", "rows": B * Hq, "layer": L - 1,
                  "finite": bool(torch.isfinite(ours_last.float()).all().item())}

    if rank != 0:
        return None
    pk, pk_kind = peaks()
    alg = algorithmic_bytes(lens, Hq, Hkv)
    achieved = alg / (ms_kernel * 1e-3) / 1e9
    traffic, traffic_src = measured_traffic(f"{args.workload}_tp{world}")
    return {
        "metric": "decode tokens/sec (bench.py shape)", "value": B / (ms_step * 1e-3),
        "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {
            "workload": f"{args.workload}: {w['model']} attention path, {L} layers, batch {B}, "
                        f"ctx {min(lens)}-{max(lens)} (mean {sum(lens) // B}), heads "
                        f"{w['Hq']}/{w['Hkv']} over TP={world} -> {Hq}/{Hkv} per rank, "
                        f"block_size {BLOCK}, CUDA graph replay",
            "l2": f"inputs larger than L2: {alg * L / 1e9:.2f} GB of KV streamed per step",
            "parallelism": f"tp{world}", "decode_splits": int(n_ctas),
            "decode_engine": os.environ.get("B200ATTN_DECODE_ENGINE", "rows"),
            "decode_overlap": os.environ.get("B200_BENCH_OVERLAP", "1") != "0",
        },
        "roofline": {
            "bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
            "frac": achieved / pk["hbm_gbs"], "traffic": traffic,
            "traffic_source": (f"committed ncu --set full capture ({traffic_src}), not "
                               f"re-measured by this run") if traffic is not None else None,
            "peak_kind": pk_kind,
            "kernel": "decode_stream_kernel<rows engine> (CTA per KV head x split x sequence, "
                      "in-launch merge), graph replay of one launch per layer",
            "algorithmic_bytes_per_launch": alg, "ms_per_launch": ms_kernel,
            "frac_of_8TBs_nominal": achieved / 8000.0,
            # numbers the record should keep whole (the driver's parse keeps this object)
            "vs_flash_attn": (fa["ms_per_step"] / ms_step) if "ms_per_step" in fa else None,
            "vs_flash_attn_overlap_off": (fa["ms_per_step"] / ms_step_plain)
            if "ms_per_step" in fa else None,
            "ms_per_step_overlap_off": ms_step_plain,
            "value_overlap_off": B / (ms_step_plain * 1e-3),
            "max_abs_diff_vs_flash_attn": parity["max_abs_diff"] if parity else None,
            "within_tolerance_vs_flash_attn": parity["within_tolerance"] if parity else None,
        },
        "parity_vs_flash_attn": parity,
        "e2e": {"value": B / (ms_e2e * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e,
                "path": f"pinned HOST q/k/v slab per {group_layers}-layer group -> device (copy "
                        f"stream), captured graph of Attention.forward per group, outputs back to "
                        f"pinned host (second copy stream); H2D/D2H inside the timed region"},
        "gpu_launches": int(launches_per_step * args.steps),
        "launches_per_step": int(launches_per_step),
        "clocks": clocks,
        "flash_attn": fa,
        "vs_flash_attn": (fa["ms_per_step"] / ms_step) if "ms_per_step" in fa else None,
    }


def flash_attn_leg(args, st, dslab, dmeta, B, Hq, Hkv, L, warmup, barrier):
    """The reference's Attention.forward (decode branch: Triton store_kvcache +
    flash_attn_with_kvcache) per layer on the same caches / tables / lens / inputs, captured in a
    CUDA graph like ModelRunner captures it, timed exactly like ours. Runs on every rank (the
    timing is a max over ranks); returns a dict on every rank."""
    from byox_vllm_b200 import parallel

    mods, why = load_reference_attention()
    ok = 1.0 if mods else 0.0
    if parallel.max_over_ranks(-ok) != -1.0:  # some rank cannot run it: nobody does
        return {"unavailable": why or "unavailable on another rank"}
    att, ctx = mods
    kv = st["kv_cache"]
    ref_layers = []
    for layer in range(L):
        a = att.Attention(Hq, D, D ** -0.5, Hkv)
        a.k_cache, a.v_cache = kv[0, layer], kv[1, layer]
        ref_layers.append(a)
    ctx.set_context(False, slot_mapping=dmeta["slot_mapping"], context_lens=dmeta["context_lens"],
                    block_tables=dmeta["block_tables"])

    def step_ref():
        o = None
        for layer in range(L):
            q, k, v = slab_views(dslab[layer], B, Hq, Hkv)
            o = ref_layers[layer](q, k, v)
        return o

    try:
        g, out = capture_out(step_ref)
        ms = parallel.max_over_ranks(time_replays(g, args.steps, warmup, barrier))
        last = out.clone()
    finally:
        ctx.reset_context()
    return {"_last_out": last.view(B, Hq, D), "ms_per_step": ms, "tokens_per_s": B / (ms * 1e-3), "version": flash_attn_version(),
            "path": "baseline/_ref nanovllm.layers.attention.Attention.forward per layer "
                    "(Triton store_kvcache + flash_attn_with_kvcache), same caches / tables / "
                    "lens / inputs, CUDA graph replay, same steps and warm-up"}


# ------------------------------------------------------------------------------------------------
# extras: kernel-level rooflines beside the reference's kernels (single GPU)
# ------------------------------------------------------------------------------------------------
_flush_buf = None


def l2_flush():
    global _flush_buf
    if _flush_buf is None:
        _flush_buf = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    _flush_buf.zero_()


def l2_flush_clean():
    """l2_flush(), then a 512 MiB READ: the timed kernel's data is as cold as after l2_flush(), but
    L2 holds clean lines. After the write-flush alone L2 is full of the flush buffer's DIRTY lines,
    and a kernel that writes (the KV scatter) pays their write-back inside its own timed region:
    ncu, whose cache control leaves L2 clean, times store_kvcache_kernel at 22 us where the
    write-flushed timing says 28.7 (profiles/r03_store_full.txt)."""
    l2_flush()
    _flush_buf.view(torch.int32).max()


def timed_call(fn, iters=8, warm=3, clean=False):
    """Median ms of ONE call, L2 flushed (512 MiB written; clean=True: then 512 MiB read) before
    every timed call."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        (l2_flush_clean if clean else l2_flush)()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)


def timed_graph_many(make_fn, n_copies, reps=5):
    """ms per launch with the launches of `n_copies` independent input sets captured into ONE graph
    (back-to-back launches like the layers of a model step; the sets together exceed L2)."""
    fns = [make_fn(i) for i in range(n_copies)]

    def all_():
        for f in fns:
            f()

    g = capture(all_)
    return time_replays(g, reps, 2, torch.cuda.synchronize) / n_copies


def decode_inputs(B, lens_l, Hq, Hkv, n_copies, shared_pages=0, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    lens = torch.tensor(lens_l, dtype=torch.int32, device="cuda")
    pages = [(c + BLOCK - 1) // BLOCK for c in lens_l]
    W = max(pages)
    own = [p - min(shared_pages, p) for p in pages]
    nb = shared_pages + sum(own) + 4
    sets = []
    for _ in range(n_copies):
        perm = torch.randperm(nb, generator=g, device="cuda").to(torch.int32).cpu()
        bt = torch.full((B, W), -1, dtype=torch.int32)
        at = shared_pages
        for b, (p, o) in enumerate(zip(pages, own)):
            s = p - o
            bt[b, :s] = perm[:s]
            bt[b, s:p] = perm[at:at + o]
            at += o
        bt = bt.cuda()
        kc = torch.empty(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
        vc = torch.empty(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
        q = torch.empty(B, Hq, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
        sets.append((q, kc, vc, bt))
    unique_tokens = shared_pages * BLOCK + sum(c - min(shared_pages * BLOCK, c) for c in lens_l)
    io = 2 * B * Hq * D * 2 + 4 * sum(pages) + 4 * B
    return lens, sets, unique_tokens * Hkv * D * 4 + io, sum(lens_l) * Hkv * D * 4 + io


def extra_decode(name, B, ctx, Hq, Hkv, shared_tokens=0, ref=True, overlap=None):
    from byox_vllm_b200.attention import (cascade_chunks, cascade_plan, decode_num_splits,
                                          paged_decode)

    if overlap is not None:  # overlap mode as under install(): launches are PDL dependents
        from byox_vllm_b200.attention import set_decode_overlap

        prev = set_decode_overlap(overlap)
        try:
            rec = extra_decode(name, B, ctx, Hq, Hkv, shared_tokens, ref)
        finally:
            set_decode_overlap(prev)
        rec["overlap"] = bool(overlap)
        return rec

    lens_l = context_lengths(dict(ctx=ctx, batch=B))
    per_copy = sum(lens_l) * Hkv * D * 4
    n_copies = max(2, min(24, -(-(400 << 20) // max(per_copy, 1))))
    sp = shared_tokens // BLOCK
    lens, sets, alg, logical = decode_inputs(B, lens_l, Hq, Hkv, n_copies, sp)
    scale = D ** -0.5
    W = sets[0][3].size(1)
    outs = [torch.empty_like(s[0]) for s in sets]
    chunks = cascade_chunks(B, Hq, Hkv, W, BLOCK) if sp else 0
    plans = [cascade_plan(lens, s[3], BLOCK) if chunks else None for s in sets]

    def ours(i):
        q, kc, vc, bt = sets[i]
        return lambda: paged_decode(q, kc, vc, lens, bt, scale, out=outs[i], plan=plans[i],
                                    prefix_chunks=chunks)

    ms = timed_graph_many(ours, n_copies)
    pk, pk_kind = peaks()
    rec = {"workload": name, "batch": B, "ctx": f"{min(lens_l)}-{max(lens_l)}",
           "heads": f"{Hq}/{Hkv}", "bound": "hbm", "ms_per_launch": ms,
           "algorithmic_bytes_per_launch": alg, "achieved": alg / ms / 1e6, "peak": pk["hbm_gbs"],
           "unit": "GB/s", "frac": alg / ms / 1e6 / pk["hbm_gbs"],
           "frac_of_8TBs_nominal": alg / ms / 1e6 / 8000.0, "peak_kind": pk_kind,
           "decode_ctas": decode_num_splits(B, Hkv, W, BLOCK),
           "timing": f"{n_copies} independent cache sets ({n_copies * per_copy / 1e9:.2f} GB > L2) "
                     f"captured back to back in one graph, CUDA events"}
    if sp:
        rec.update(shared_prefix_tokens=sp * BLOCK, cascade_chunks=chunks,
                   logical_bytes_per_launch=logical, effective_gbs=logical / ms / 1e6,
                   note="achieved = unique-page bytes (each physical page once); effective = "
                        "per-row bytes (may exceed the peak: shared pages are attended once)")
    try:
        if not ref:
            raise ImportError("reference leg not requested")
        from flash_attn import flash_attn_with_kvcache

        def fa(i):
            q, kc, vc, bt = sets[i]
            q4 = q.unsqueeze(1)
            return lambda: flash_attn_with_kvcache(q4, kc, vc, cache_seqlens=lens, block_table=bt,
                                                   softmax_scale=scale, causal=True)

        ms_fa = timed_graph_many(fa, n_copies)
        ref = flash_attn_with_kvcache(sets[0][0].unsqueeze(1), sets[0][1], sets[0][2],
                                      cache_seqlens=lens, block_table=sets[0][3],
                                      softmax_scale=scale, causal=True).squeeze(1)
        ours(0)()
        torch.cuda.synchronize()
        rec.update(flash_attn_ms=ms_fa, vs_flash_attn=ms_fa / ms,
                   max_abs_diff_vs_flash_attn=(outs[0].float() - ref.float()).abs().max().item())
    except ImportError as e:
        rec["flash_attn_unavailable"] = str(e)
    del sets, outs
    torch.cuda.empty_cache()
    return rec


def prefill_flops(lens, cached, Hq):
    fl = 0
    for n, c in zip(lens, cached):
        lq = n - c
        fl += 4 * Hq * D * (lq * c + lq * (lq + 1) // 2)
    return fl


def extra_prefill(name, lens, cached, Hq, Hkv):
    import random

    from byox_vllm_b200.attention import prefill_varlen
    from tools import metadata_ref as md

    if lens is None:  # the reference bench's prefill step: ~24 prompts of 100-1024 tokens
        rng = random.Random(0)
        lens = [rng.randint(100, 1024) for _ in range(24)]
    cached = cached or [0] * len(lens)
    g = torch.Generator(device="cuda").manual_seed(1)
    need = [(n + BLOCK - 1) // BLOCK for n in lens]
    nb = sum(need) + 2
    ids = torch.randperm(nb, generator=g, device="cuda").tolist()
    seqs, at = [], 0
    for n, nbk, c in zip(lens, need, cached):
        seqs.append(md.SeqState(n, ids[at:at + nbk], c))
        at += nbk
    m = md.prepare_prefill(seqs, BLOCK)
    total_q = sum(n - c for n, c in zip(lens, cached))
    q = torch.empty(total_q, Hq, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
    cu_q, cu_k = m["cu_seqlens_q"].cuda(), m["cu_seqlens_k"].cuda()
    scale = D ** -0.5
    if any(cached):
        kc = torch.empty(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
        vc = torch.empty_like(kc).normal_(generator=g)
        k, v, bt = kc, vc, m["block_tables"].cuda()
    else:
        packed = torch.empty(total_q, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16,
                             device="cuda").normal_(generator=g)
        k = packed[:, Hq * D:(Hq + Hkv) * D].contiguous().view(total_q, Hkv, D)
        v = packed[:, (Hq + Hkv) * D:].view(total_q, Hkv, D)
        bt = None
    o = torch.empty_like(q)
    fl = prefill_flops(lens, cached, Hq)
    ms = timed_call(lambda: prefill_varlen(q, k, v, cu_q, cu_k, m["max_seqlen_q"],
                                           m["max_seqlen_k"], scale, block_tables=bt, out=o))
    pk, pk_kind = peaks()
    rec = {"workload": name, "heads": f"{Hq}/{Hkv}", "total_q": total_q, "bound": "tensor",
           "ms_per_launch": ms, "algorithmic_flops_per_launch": fl, "achieved": fl / ms / 1e9,
           "peak": pk["bf16_tflops"], "unit": "TFLOP/s", "frac": fl / ms / 1e9 / pk["bf16_tflops"],
           "peak_kind": pk_kind + " (burst: kernel timed alone)",
           "timing": "one call, L2 flushed before each, CUDA events, median of 8"}
    try:
        from flash_attn import flash_attn_varlen_func

        fa = lambda: flash_attn_varlen_func(  # noqa: E731
            q, k, v, max_seqlen_q=m["max_seqlen_q"], cu_seqlens_q=cu_q,
            max_seqlen_k=m["max_seqlen_k"], cu_seqlens_k=cu_k, softmax_scale=scale, causal=True,
            block_table=bt)
        ms_fa = timed_call(fa)
        rec.update(flash_attn_ms=ms_fa, vs_flash_attn=ms_fa / ms,
                   max_abs_diff_vs_flash_attn=(o.float() - fa().float()).abs().max().item())
    except ImportError as e:
        rec["flash_attn_unavailable"] = str(e)
    del q, k, v, o
    torch.cuda.empty_cache()
    return rec


def extra_store(N, Hq, Hkv):
    from byox_vllm_b200.attention import store_kvcache

    row = Hkv * D
    g = torch.Generator(device="cuda").manual_seed(N)
    k = torch.empty(N, Hkv, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
    packed = torch.empty(N, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16,
                         device="cuda").normal_(generator=g)
    v = packed[:, (Hq + Hkv) * D:].view(N, Hkv, D)
    nb = max(64, (N + BLOCK - 1) // BLOCK * 2)
    kc = torch.zeros(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros_like(kc)
    slots = torch.randperm(nb * BLOCK, generator=g, device="cuda")[:N].to(torch.int32)
    alg = N * row * 2 * 2 * 2 + 4 * N
    ms = timed_call(lambda: store_kvcache(k, v, kc, vc, slots), iters=12)
    ms_clean = timed_call(lambda: store_kvcache(k, v, kc, vc, slots), iters=12, clean=True)
    pk, pk_kind = peaks()
    rec = {"workload": f"store_kvcache {N} tokens x Hkv {Hkv}", "bound": "hbm", "ms_per_launch": ms,
           "ms_per_launch_clean_l2": ms_clean,
           "frac_clean_l2": alg / ms_clean / 1e6 / pk["hbm_gbs"],
           "algorithmic_bytes_per_launch": alg, "achieved": alg / ms / 1e6, "peak": pk["hbm_gbs"],
           "unit": "GB/s", "frac": alg / ms / 1e6 / pk["hbm_gbs"], "peak_kind": pk_kind,
           "timing": "one call, L2 flushed before each, CUDA events, median of 12"}
    mods, why = load_reference_attention()
    if mods:
        kc2, vc2 = torch.zeros_like(kc), torch.zeros_like(vc)
        ms_ref = timed_call(lambda: mods[0].store_kvcache(k, v, kc2, vc2, slots), iters=12)
        ms_ref_clean = timed_call(lambda: mods[0].store_kvcache(k, v, kc2, vc2, slots), iters=12,
                                  clean=True)
        rec.update(reference_triton_ms=ms_ref, vs_reference=ms_ref / ms,
                   reference_triton_ms_clean_l2=ms_ref_clean,
                   vs_reference_clean_l2=ms_ref_clean / ms_clean,
                   bit_equal=bool(torch.equal(kc, kc2) and torch.equal(vc, vc2)))
    else:
        rec["reference_unavailable"] = why
    return rec


def extra_qkv(N, Hq, Hkv):
    """SURVEY §8f-4, prefill-size pass: q/k RMSNorm + RoPE + KV scatter on the raw qkv_proj output
    in ONE launch, beside the reference's chain (its own RMSNorm x 2 and RotaryEmbedding modules,
    torch.compile'd as shipped, + its Triton store_kvcache: four launches)."""
    from byox_vllm_b200 import qkv_attention as fused

    eps, max_pos = 1e-6, 40960
    g = torch.Generator(device="cuda").manual_seed(N + Hq)
    qkv = torch.empty(N, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16,
                      device="cuda").normal_(generator=g)
    pos = torch.randint(0, max_pos, (N,), generator=g, device="cuda")
    nb = (N + BLOCK - 1) // BLOCK * 2
    kc = torch.zeros(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros_like(kc)
    slots = torch.randperm(nb * BLOCK, generator=g, device="cuda")[:N].to(torch.int32)
    wq = torch.empty(D, dtype=torch.bfloat16, device="cuda").normal_(1.0, 0.25, generator=g)
    wk = torch.empty(D, dtype=torch.bfloat16, device="cuda").normal_(1.0, 0.25, generator=g)
    inv = 1.0 / (1e6 ** (torch.arange(0, D, 2, device="cuda").float() / D))
    fr = torch.arange(max_pos, device="cuda").float()[:, None] * inv[None]
    cache = torch.cat([fr.cos(), fr.sin()], dim=-1).contiguous()  # rotary_embedding.py:27-35
    work = qkv.clone()
    # algorithmic bytes: row read + q_out + k in place (dense prefill reads it) + K/V cache rows
    alg = N * ((Hq + 2 * Hkv) * 256 + Hq * 256 + Hkv * 256 + 2 * Hkv * 256 + 12)
    ms = timed_call(lambda: fused.qk_norm_rope_store(work, pos, wq, wk, eps, cache, Hq, Hkv, D,
                                                     k_cache=kc, v_cache=vc, slot_mapping=slots),
                    iters=10)
    pk, pk_kind = peaks()
    rec = {"workload": f"q/k norm + RoPE + KV store, {N} tokens x heads {Hq}/{Hkv}", "bound": "hbm",
           "ms_per_launch": ms, "algorithmic_bytes_per_launch": alg, "achieved": alg / ms / 1e6,
           "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": alg / ms / 1e6 / pk["hbm_gbs"],
           "peak_kind": pk_kind, "timing": "one call, L2 flushed before each, CUDA events, median of 10"}
    mods, why = load_reference_attention()
    try:
        if not mods:
            raise ImportError(why)
        import importlib

        ln = importlib.import_module("nanovllm.layers.layernorm")
        ro = importlib.import_module("nanovllm.layers.rotary_embedding")
        torch.set_default_device("cuda")
        try:
            rope = ro.RotaryEmbedding(D, D, max_pos, 1e6)
            torch.set_default_dtype(torch.bfloat16)
            qn, kn = ln.RMSNorm(D, eps), ln.RMSNorm(D, eps)
        finally:
            torch.set_default_dtype(torch.float32)
            torch.set_default_device("cpu")
        qn.weight.data.copy_(wq)
        kn.weight.data.copy_(wk)
        kc2, vc2 = torch.zeros_like(kc), torch.zeros_like(vc)

        def ref():
            q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
            q, k, v = q.view(-1, Hq, D), k.view(-1, Hkv, D), v.view(-1, Hkv, D)
            q, k = rope(pos, qn(q), kn(k))
            mods[0].store_kvcache(k, v, kc2, vc2, slots)
            return q

        ms_ref = timed_call(ref, iters=10)
        rec.update(reference_chain_ms=ms_ref, vs_reference=ms_ref / ms,
                   v_cache_bit_equal=bool(torch.equal(vc, vc2)))
    except Exception as e:  # torch.compile / triton not usable on this box
        rec["reference_unavailable"] = f"{type(e).__name__}: {str(e)[:200]}"
    del qkv, work, kc, vc
    torch.cuda.empty_cache()
    return rec


def run_extras(which):
    out = {}
    with torch.inference_mode():
        if "c3" in which:
            out["c3_decode_b128_ctx4096"] = extra_decode(
                "c3: batch 128 x ctx 4096, 32/8", 128, 4096, 32, 8)
        if "c4" in which:
            out["c4_decode_tp8_rank"] = extra_decode(
                "c4 per rank: batch 256 x ctx 8192, 7680 shared, 8/1 (cascade)", 256, 8192, 8, 1,
                shared_tokens=7680)
        if "c5" in which:
            for B, ctx in ((64, 4096), (16, 4096), (1, 131072), (512, 1024), (16, 1024)):
                out[f"c5_decode_b{B}_ctx{ctx}"] = extra_decode(
                    f"c5: batch {B} x ctx {ctx}, 32/8", B, ctx, 32, 8)
        if "tp8" in which:
            for Hq, Hkv, tp in ((2, 1, 8), (4, 2, 4), (8, 4, 2)):
                out[f"c2_decode_tp{tp}_rank"] = extra_decode(
                    f"c2 per rank at TP={tp}: batch 256, bench lengths, {Hq}/{Hkv}", 256, "bench",
                    Hq, Hkv)
        if "prefill" in which:
            for name, lens, cached, Hq, Hkv in (
                    ("bench24", None, None, 16, 8), ("16x1024", [1024] * 16, None, 16, 8),
                    ("4x4096", [4096] * 4, None, 32, 8), ("1x16384", [16384], None, 32, 8),
                    ("32x(7680+512)", [8192] * 32, [7680] * 32, 8, 1)):
                out[f"prefill_{name}"] = extra_prefill(name, lens, cached, Hq, Hkv)
        if "qkv" in which:
            out["qkv_prep_16384_32_8"] = extra_qkv(16384, 32, 8)
            out["qkv_prep_16384_8_1"] = extra_qkv(16384, 8, 1)
        if "store" in which:
            out["store_16384x8"] = extra_store(16384, 32, 8)
            out["store_16384x1"] = extra_store(16384, 4, 1)
    return out


# ------------------------------------------------------------------------------------------------
# engine leg: the unmodified reference engine end to end (single GPU)
# ------------------------------------------------------------------------------------------------
def run_engine_e2e(budget_s=300, tp=1):
    """BASELINE config 2 through baseline/_ref's LLM.generate (tools/e2e_nanovllm.py) in
    subprocesses: reference attention; at TP=1 also the drop-in at the frozen Attention.forward
    boundary (no fused QKV, no device-side metadata); the drop-in with every SURVEY §8(f) row
    enabled. tp > 1 (the engine spawns its own ranks, llm_engine.py:24-31): the ref and full arms,
    both with --check (near-greedy tokens of 8 short prompts) -> tokens_equal_frac."""
    if not (ROOT / "baseline" / "_ref" / "nanovllm").exists():
        return {"skipped": "baseline/_ref is not installed"}
    arms = (("ref", "ref", {}),
            ("ours_frozen", "ours", {"B200ATTN_FUSE_QKV": "0", "B200ATTN_FAST_META": "0"}),
            ("ours_full", "ours", {}))
    if tp > 1:
        arms = (arms[0], arms[2])
    res = {"workload": "Qwen3-0.6B random-init, 256 seqs, in/out 100-1024 (reference "
                       f"bench.py:17-42), CUDA graphs on, TP={tp}; tokens/s = sum(max_tokens) / "
                       "wall of generate()"}
    check = {}
    t_begin = time.time()
    for key, impl, env_extra in arms:
        if tp > 1 and time.time() - t_begin > 1.5 * budget_s:
            res[key] = {"skipped": "engine budget of this N spent"}
            continue
        env = dict(os.environ, **env_extra)
        for k in list(env):  # torchrun's rendezvous / per-rank variables are not the engine's
            if k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT",
                     "LOCAL_WORLD_SIZE", "GROUP_RANK", "ROLE_RANK", "ROLE_WORLD_SIZE",
                     "OMP_NUM_THREADS", "GROUP_WORLD_SIZE", "ROLE_NAME") \
                    or k.startswith("TORCHELASTIC_"):
                env.pop(k)
        t0 = time.time()
        cmd = [sys.executable, str(ROOT / "tools" / "e2e_nanovllm.py"), "--impl", impl,
               "--model", "0.6b", "--tp", str(tp)] + (["--check"] if tp > 1 else [])
        try:
            r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=budget_s)
        except subprocess.TimeoutExpired:
            res[key] = {"error": f"timeout after {budget_s} s"}
            continue
        line = [ln for ln in r.stdout.splitlines() if ln.startswith('{"tool"')]
        if r.returncode != 0 or not line:
            res[key] = {"error": (r.stderr or r.stdout)[-400:]}
            continue
        j = json.loads(line[-1])
        res[key] = {"tok_s": j["tokens_per_s"], "decode_tok_s": j["decode_tokens_per_s"],
                    "prefill_tok_s": j["prefill_tokens_per_s"],
                    "wall_s": round(time.time() - t0, 1)}
        if "check_tokens" in j:
            check[key] = j["check_tokens"]
    if tp > 1 and all("tok_s" in res.get(k, {}) for k in ("ref", "ours_full")):
        res.update(ref_tok_s=res["ref"]["tok_s"], ours_full_tok_s=res["ours_full"]["tok_s"],
                   full_vs_ref=res["ours_full"]["tok_s"] / res["ref"]["tok_s"])
        if "ref" in check and "ours_full" in check:
            pairs = [(x, y) for a, b in zip(check["ref"], check["ours_full"]) for x, y in zip(a, b)]
            res["tokens_equal_frac"] = sum(x == y for x, y in pairs) / max(len(pairs), 1)
            res["check_prompts_identical"] = sum(a == b for a, b in
                                                 zip(check["ref"], check["ours_full"]))
    if all("tok_s" in res.get(k, {}) for k in ("ref", "ours_frozen", "ours_full")):
        res.update(ref_tok_s=res["ref"]["tok_s"], ours_frozen_tok_s=res["ours_frozen"]["tok_s"],
                   ours_full_tok_s=res["ours_full"]["tok_s"],
                   frozen_vs_ref=res["ours_frozen"]["tok_s"] / res["ref"]["tok_s"],
                   full_vs_ref=res["ours_full"]["tok_s"] / res["ref"]["tok_s"])
    return res


# ------------------------------------------------------------------------------------------------
# CPU legs
# ------------------------------------------------------------------------------------------------
def cpu_reference(w, steps, warmup, budget_s=20.0):
    """The CPU restatement (oracle/) of store_kvcache + paged decode, timed on the host cores on a
    bounded sample: ONE layer of the workload over its first CPU_SAMPLE_SEQS sequences; tokens/s
    is scaled by the layer count (layers are identical work). Both CPU legs use this sample."""
    from oracle import attention_oracle as orc  # timed CPU baseline leg
    from tools import synth

    lens = context_lengths(w)
    n = min(CPU_SAMPLE_SEQS, len(lens))
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
    sample = (f"1 of {w['layers']} layers x first {n} of {w['batch']} sequences (store + paged "
              f"decode, fp32 torch on CPU), {len(times)} timed passes, median {t * 1e3:.1f} "
              f"ms/pass; tokens/s = {n} / (pass time x {w['layers']} layers)")
    return tok_s, t, n, len(times), sample


def run_reference(args, w, rank, world):
    if rank != 0:
        return None
    # torchrun exports OMP_NUM_THREADS=1; the reference arm uses every host thread it can.
    torch.set_num_threads(len(os.sched_getaffinity(0)))
    tok_s, t, n, done, sample = cpu_reference(w, args.steps, max(args.warmup, 1), budget_s=120.0)
    cores = torch.get_num_threads()
    return {
        "impl": "reference", "metric": "decode tokens/sec (bench.py shape)", "value": tok_s,
        "unit": "tokens/s", "n_gpus": world, "steps": done, "warmup": max(args.warmup, 1),
        # extrapolated: one full-batch step of all layers at the sampled rate (the timed pass is
        # 1 layer x n sequences; the driver's wall clock sees only the sampled passes)
        "ms_per_step": w["batch"] / tok_s * 1e3, "ms_per_sampled_pass": t * 1e3,
        "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {w['model']} attention path on host cores",
                   "parallelism": "cpu"},
        "cpu_baseline": {"value": tok_s, "unit": "tokens/s", "cores": cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": tok_s, "unit": "tokens/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def emit_extras(line, extra):
    """The kernel-level blocks go out as their OWN json line first (the final line stays short
    enough for a log tail); the numbers the claims rest on are copied into `roofline`."""
    print(json.dumps({"extra": extra}), flush=True)
    keep = ("ms_per_launch", "frac", "frac_of_8TBs_nominal", "vs_flash_attn",
            "max_abs_diff_vs_flash_attn", "achieved", "unit", "ms_per_launch_clean_l2",
            "frac_clean_l2", "vs_reference", "vs_reference_clean_l2")
    for src, dst in (("c3_decode_b128_ctx4096", "target_c3"), ("prefill_bench24", "prefill_bench24"),
                     ("prefill_1x16384", "prefill_1x16384"), ("store_16384x8", "store_16384x8"),
                     ("c2_decode_tp8_rank", "c2_tp8_rank")):
        blk = extra.get(src)
        if isinstance(blk, dict):
            line["roofline"][dst] = {k: blk[k] for k in keep if k in blk}
    line["extra_keys"] = sorted(extra)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS) + list(EXTRA_ONLY))
    ap.add_argument("--extras", default=None,
                    help="comma list of extra blocks (c3,c4,c5,tp8,prefill,store,qkv) or 'none'; "
                         "default: all at N=1, none at N>1")
    ap.add_argument("--engine", default=None, choices=["on", "off"],
                    help="run the reference-engine e2e arms (default: on at N=1; "
                         "B200_BENCH_ENGINE=0 turns it off)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.workload in EXTRA_ONLY:
        assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
        torch.cuda.set_device(local)
        if rank == 0:
            print(json.dumps({"extra": run_extras({args.workload})}), flush=True)
        return
    w = WORKLOADS[args.workload]

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
        engine = args.engine or ("off" if os.environ.get("B200_BENCH_ENGINE") == "0" else "on")
        if world > 1:
            # north_star's target at TP=N: the unmodified reference engine against the drop-in,
            # run by rank 0 as subprocesses (the engine spawns its own N ranks) once every rank of
            # this job has released its GPU memory. The other ranks wait on the HOST (the job's
            # key-value store), not in an NCCL kernel that would sit on the GPUs being measured.
            import gc
            from datetime import timedelta

            gc.collect()
            torch.cuda.empty_cache()
            torch.distributed.barrier()
            torch.cuda.synchronize()
            store = torch.distributed.distributed_c10d._get_default_store()
            if rank == 0:
                try:
                    if engine == "on" and time.time() - T_START < 420:
                        eng = run_engine_e2e(budget_s=200, tp=world)
                        line["engine_e2e"] = eng
                        for k in ("ref_tok_s", "ours_full_tok_s", "full_vs_ref",
                                  "tokens_equal_frac"):
                            if k in eng:
                                line["e2e"]["engine_" + k] = eng[k]
                    elif engine == "on":
                        line["engine_e2e"] = {"skipped": "time budget of this N spent"}
                finally:
                    store.set("b200_bench_engine_done", "1")
            else:
                store.wait(["b200_bench_engine_done"], timedelta(seconds=900))
        if rank == 0 and world > 1:
            line["cpu_baseline"] = None  # timed at N=1 only
            if args.extras and args.extras != "none":
                emit_extras(line, run_extras(set(args.extras.split(","))))
            print(json.dumps(line), flush=True)
        elif rank == 0:
            extras = args.extras or "c3,c4,c5,tp8,prefill,store,qkv"
            if extras != "none":
                emit_extras(line, run_extras(set(extras.split(","))))
            if engine == "on":
                torch.cuda.empty_cache()
                eng = run_engine_e2e()
                line["engine_e2e"] = eng
                for k in ("ref_tok_s", "ours_frozen_tok_s", "ours_full_tok_s", "frozen_vs_ref",
                          "full_vs_ref"):  # the e2e object survives the driver's parse whole
                    if k in eng:
                        line["e2e"]["engine_" + k] = eng[k]
            # same passes and warm-up as the --impl reference arm takes by default
            tok_s, t, n, done, sample = cpu_reference(w, args.steps, max(args.warmup, 1),
                                                      budget_s=20.0)
            line["cpu_baseline"] = {"value": tok_s, "unit": "tokens/s",
                                    "cores": torch.get_num_threads(), "kind": "port",
                                    "sample": sample}
            print(json.dumps(line), flush=True)
    finally:
        if world > 1:
            torch.distributed.barrier()
            torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
