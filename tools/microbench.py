"""Kernel-level microbenchmarks on one B200: ours vs the reference's kernels on the same inputs.

    gpurun -- python tools/microbench.py [decode] [store] [prefill] [--quick]

Writes one JSON object per line to gpurun_out/microbench.jsonl. The reference bars are the kernels
the reference calls (flash_attn_with_kvcache / flash_attn_varlen_func of flash-attn 2.8.3, and a
Triton scatter equivalent to the reference's store_kvcache_kernel) — they are measured beside
ours, never used by the product path. SURVEY.md §7.3(e), §8(d) config C5.

Timing: CUDA events on the current stream around ONE call, L2 flushed (512 MiB memset) before
every timed call, 3 warm-ups, `iters` timed calls, min and median reported. Bytes are the
algorithmic bytes of SURVEY.md §8(d).
"""
from __future__ import annotations

import json
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

D, BLOCK = 128, 256
PEAK = 6543.4
_p = ROOT / "MEASURED_PEAKS.json"
if _p.exists():
    PEAK = json.loads(_p.read_text())["hbm_gbs"]
TF_PEAK = json.loads(_p.read_text())["bf16_tflops"] if _p.exists() else 1590.0

_flush = None


def timed(fn, iters=10, warm=3):
    global _flush
    if _flush is None:
        _flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        _flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), statistics.median(ts)


def decode_inputs(B, ctx, Hq, Hkv, ragged=False, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    if ragged:
        lens = torch.randint(ctx // 2, ctx + 1, (B,), generator=g, device="cuda", dtype=torch.int32)
    else:
        lens = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
    pages = (lens + BLOCK - 1) // BLOCK
    W = int(pages.max())
    nb = B * W + 4
    perm = torch.randperm(nb, generator=g, device="cuda")[:B * W].to(torch.int32).view(B, W)
    ar = torch.arange(W, device="cuda").unsqueeze(0)
    bt = torch.where(ar < pages.unsqueeze(1), perm, torch.full_like(perm, -1))
    kc = torch.empty(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
    vc = torch.empty(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
    q = torch.empty(B, Hq, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
    toks = int(lens.sum())
    alg = toks * Hkv * D * 2 * 2 + 2 * B * Hq * D * 2 + 4 * int(pages.sum()) + 4 * B
    return q, kc, vc, lens, bt, alg


DECODE_SHAPES = [  # (B, ctx, Hq, Hkv)
    (128, 4096, 32, 8), (64, 4096, 32, 8), (128, 4096, 4, 1), (256, 8192, 8, 1),
    (256, 640, 16, 8), (1, 131072, 32, 8), (8, 16384, 32, 8), (512, 1024, 32, 8),
    (1, 1024, 32, 8), (64, 8192, 32, 8), (16, 4096, 32, 8), (4, 4096, 32, 8),
    (32, 2048, 16, 8), (1, 4096, 16, 8), (128, 4096, 16, 4), (128, 4096, 8, 2),
    (2, 32768, 32, 8), (256, 1024, 16, 8),
]
DECODE_QUICK = [(128, 4096, 32, 8), (1, 131072, 32, 8), (256, 640, 16, 8), (8, 16384, 32, 8),
                (1, 1024, 32, 8), (128, 4096, 4, 1)]


def bench_decode(out, quick):
    from flash_attn import flash_attn_with_kvcache

    from byox_vllm_b200.attention import decode_num_splits, paged_decode

    for (B, ctx, Hq, Hkv) in (DECODE_QUICK if quick else DECODE_SHAPES):
        for ragged in ((False,) if quick else (False, True)):
            q, kc, vc, lens, bt, alg = decode_inputs(B, ctx, Hq, Hkv, ragged)
            scale = D ** -0.5
            o = torch.empty_like(q)
            splits = decode_num_splits(B, Hkv, bt.size(1), BLOCK)
            t_min, t_med = timed(lambda: paged_decode(q, kc, vc, lens, bt, scale, out=o))
            q4 = q.unsqueeze(1)
            btc = bt.clamp_min(0)
            f_min, f_med = timed(lambda: flash_attn_with_kvcache(
                q4, kc, vc, cache_seqlens=lens, block_table=btc, softmax_scale=scale, causal=True))
            ref = flash_attn_with_kvcache(q4, kc, vc, cache_seqlens=lens, block_table=btc,
                                          softmax_scale=scale, causal=True).squeeze(1)
            err = (o.float() - ref.float()).abs().max().item()
            rec = dict(kind="decode", B=B, ctx=ctx, Hq=Hq, Hkv=Hkv, ragged=ragged, splits=splits,
                       alg_bytes=alg, ours_us=t_med * 1e3, ours_min_us=t_min * 1e3,
                       ours_gbs=alg / t_med / 1e6, ours_frac=alg / t_med / 1e6 / PEAK,
                       fa2_us=f_med * 1e3, fa2_gbs=alg / f_med / 1e6, speedup=f_med / t_med,
                       max_abs_diff_vs_fa2=err)
            out.write(json.dumps(rec) + "\n")
            out.flush()
            print(f"decode B={B:4d} ctx={ctx:6d} {Hq}/{Hkv} rag={int(ragged)} S={splits:3d}: ours "
                  f"{t_med * 1e3:8.1f} us {rec['ours_gbs']:7.0f} GB/s ({rec['ours_frac']:.2f}) | "
                  f"fa2 {f_med * 1e3:8.1f} us {rec['fa2_gbs']:7.0f} GB/s | x{rec['speedup']:.2f} "
                  f"err {err:.4f}", flush=True)
            del q, kc, vc, lens, bt
            torch.cuda.empty_cache()


def bench_sweep(out, quick):
    """BASELINE config 5: B in 1..512 x ctx in 1k..128k (Hq 32 / Hkv 8), capped at B*ctx <= 4 Mi
    tokens (16 GiB of KV) so the whole grid runs in a couple of minutes."""
    from flash_attn import flash_attn_with_kvcache

    from byox_vllm_b200.attention import decode_num_splits, paged_decode

    Hq, Hkv = 32, 8
    no_ref = "--no-ref" in sys.argv
    batches = (1, 8, 64, 512) if quick else (1, 2, 4, 8, 16, 32, 64, 128, 256, 512)
    ctxs = (1024, 8192, 131072) if quick else (1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072)
    for B in batches:
        for ctx in ctxs:
            if B * ctx > (4 << 20):
                continue
            q, kc, vc, lens, bt, alg = decode_inputs(B, ctx, Hq, Hkv, False)
            scale = D ** -0.5
            o = torch.empty_like(q)
            splits = decode_num_splits(B, Hkv, bt.size(1), BLOCK)
            t_min, t_med = timed(lambda: paged_decode(q, kc, vc, lens, bt, scale, out=o), iters=6)
            rec = dict(kind="sweep", B=B, ctx=ctx, Hq=Hq, Hkv=Hkv, splits=splits, alg_bytes=alg,
                       ours_us=t_med * 1e3, ours_min_us=t_min * 1e3, ours_gbs=alg / t_med / 1e6,
                       ours_frac=alg / t_med / 1e6 / PEAK)
            if not no_ref:
                q4 = q.unsqueeze(1)
                f_min, f_med = timed(lambda: flash_attn_with_kvcache(
                    q4, kc, vc, cache_seqlens=lens, block_table=bt, softmax_scale=scale,
                    causal=True), iters=6)
                rec.update(fa2_us=f_med * 1e3, fa2_gbs=alg / f_med / 1e6, speedup=f_med / t_med)
            out.write(json.dumps(rec) + "\n")
            out.flush()
            print(f"sweep B={B:4d} ctx={ctx:6d} S={splits:3d}: ours {t_med * 1e3:8.1f} us "
                  f"{rec['ours_gbs']:7.0f} GB/s ({rec['ours_frac']:.2f})"
                  + (f" | fa2 {rec['fa2_us']:8.1f} us x{rec['speedup']:.2f}" if not no_ref else ""),
                  flush=True)
            del q, kc, vc, lens, bt
            torch.cuda.empty_cache()


CASCADE_SHAPES = [  # (B, ctx, shared tokens, Hq, Hkv)  -- BASELINE config 4 per rank first
    (256, 8192, 7680, 8, 1), (256, 8192, 7680, 64, 8), (128, 4096, 3840, 32, 8),
    (64, 8192, 4096, 8, 1), (256, 8192, 0, 8, 1), (256, 2048, 1024, 16, 8),
    (256, 8192, 7680, 16, 2), (256, 8192, 7680, 32, 4), (64, 4096, 2048, 32, 8),
    (512, 2048, 1536, 32, 8), (32, 16384, 15360, 32, 8), (256, 1024, 512, 16, 8),
]


def cascade_inputs(B, ctx, shared, Hq, Hkv, seed=0):
    """Rows share the first `shared` tokens (whole pages, same physical block ids) and own the
    rest; pages are scattered. Returns unique-page and per-row (logical) algorithmic bytes."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    W = (ctx + BLOCK - 1) // BLOCK
    P = shared // BLOCK
    own = W - P
    nb = P + B * own + 4
    perm = torch.randperm(nb, generator=g, device="cuda").to(torch.int32)
    bt = torch.empty(B, W, dtype=torch.int32, device="cuda")
    bt[:, :P] = perm[:P].unsqueeze(0)
    bt[:, P:] = perm[P:P + B * own].view(B, own)
    lens = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
    kc = torch.empty(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
    vc = torch.empty(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
    q = torch.empty(B, Hq, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
    row = Hkv * D * 2 * 2
    io = 2 * B * Hq * D * 2 + 4 * B * W + 4 * B
    unique = (shared + B * (ctx - shared)) * row + io
    logical = B * ctx * row + io
    flops = 4 * B * Hq * D * ctx
    return q, kc, vc, lens, bt, unique, logical, flops


def bench_cascade(out, quick):
    from flash_attn import flash_attn_with_kvcache

    from byox_vllm_b200.attention import (cascade_chunks, cascade_plan, decode_num_splits,
                                          paged_decode)

    no_ref = "--no-ref" in sys.argv
    for (B, ctx, shared, Hq, Hkv) in (CASCADE_SHAPES[:2] if quick else CASCADE_SHAPES):
        q, kc, vc, lens, bt, unique, logical, flops = cascade_inputs(B, ctx, shared, Hq, Hkv)
        scale = D ** -0.5
        o, o2 = torch.empty_like(q), torch.empty_like(q)
        splits = decode_num_splits(B, Hkv, bt.size(1), BLOCK)
        auto = cascade_chunks(B, Hq, Hkv, bt.size(1), BLOCK)
        p_min, p_med = timed(lambda: paged_decode(q, kc, vc, lens, bt, scale, out=o2))
        plan = cascade_plan(lens, bt, BLOCK)
        pl_min, pl_med = timed(lambda: cascade_plan(lens, bt, BLOCK))
        rec = dict(kind="cascade", B=B, ctx=ctx, shared=shared, Hq=Hq, Hkv=Hkv, splits=splits,
                   auto_chunks=auto, plan=plan.cpu().tolist(), unique_bytes=unique,
                   logical_bytes=logical, flops=flops, plain_us=p_med * 1e3,
                   plain_logical_gbs=logical / p_med / 1e6, plan_us=pl_med * 1e3)
        best = None
        for chunks in sorted({max(auto, 1), 1, 2, 3, 4, 5, 6, 8, 10, 12, 16, 24, 32}):
            c_min, c_med = timed(lambda: paged_decode(q, kc, vc, lens, bt, scale, out=o, plan=plan,
                                                      prefix_chunks=chunks))
            rec[f"cascade_us_c{chunks}"] = c_med * 1e3
            if chunks == max(auto, 1):
                rec["cascade_us"] = c_med * 1e3
            if best is None or c_med < best[1]:
                best = (chunks, c_med)
        err_plain = (o.float() - o2.float()).abs().max().item()
        rec.update(best_chunks=best[0], best_us=best[1] * 1e3,
                   cascade_unique_gbs=unique / (rec["cascade_us"] * 1e-6) / 1e9,
                   cascade_logical_gbs=logical / (rec["cascade_us"] * 1e-6) / 1e9,
                   cascade_tflops=flops / (rec["cascade_us"] * 1e-6) / 1e12,
                   speedup_vs_plain=p_med * 1e3 / rec["cascade_us"],
                   max_abs_diff_vs_plain=err_plain)
        if not no_ref:
            q4 = q.unsqueeze(1)
            f_min, f_med = timed(lambda: flash_attn_with_kvcache(
                q4, kc, vc, cache_seqlens=lens, block_table=bt, softmax_scale=scale, causal=True))
            ref = flash_attn_with_kvcache(q4, kc, vc, cache_seqlens=lens, block_table=bt,
                                          softmax_scale=scale, causal=True).squeeze(1)
            rec.update(fa2_us=f_med * 1e3, speedup_vs_fa2=f_med * 1e3 / rec["cascade_us"],
                       max_abs_diff_vs_fa2=(o.float() - ref.float()).abs().max().item())
        out.write(json.dumps(rec) + "\n")
        out.flush()
        print(f"cascade B={B} ctx={ctx} shared={shared} {Hq}/{Hkv} P={rec['plan'][0]} "
              f"chunks={auto}: cascade {rec['cascade_us']:.1f} us (best c{best[0]} "
              f"{best[1] * 1e3:.1f}) | plain {p_med * 1e3:.1f} us | fa2 "
              f"{rec.get('fa2_us', float('nan')):.1f} us | plan {pl_med * 1e3:.1f} us | "
              f"x{rec['speedup_vs_plain']:.2f} vs plain, err {err_plain:.4f}", flush=True)
        del q, kc, vc, lens, bt
        torch.cuda.empty_cache()


def bench_store(out, quick):
    import triton
    import triton.language as tl

    from byox_vllm_b200.attention import store_kvcache

    # A Triton scatter with the same program structure as the reference's kernel (one program per
    # token, one row of Hkv*D elements per program; reference layers/attention.py:10-31), written
    # here only as the measured bar.
    @triton.jit
    def bar_kernel(k_ptr, ks, v_ptr, vs, kc_ptr, vc_ptr, slot_ptr, ROW: tl.constexpr):
        i = tl.program_id(0)
        slot = tl.load(slot_ptr + i)
        if slot == -1:
            return
        offs = tl.arange(0, ROW)
        tl.store(kc_ptr + slot * ROW + offs, tl.load(k_ptr + i * ks + offs))
        tl.store(vc_ptr + slot * ROW + offs, tl.load(v_ptr + i * vs + offs))

    for N in ((128, 16384) if quick else (1, 128, 256, 512, 2048, 16384)):
        for Hkv, Hq in ((8, 16), (8, 32), (1, 4)):
            row = Hkv * D
            g = torch.Generator(device="cuda").manual_seed(N)
            k = torch.empty(N, Hkv, D, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
            packed = torch.empty(N, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16,
                                 device="cuda").normal_(generator=g)
            v = packed[:, (Hq + Hkv) * D:].view(N, Hkv, D)
            nb = max(64, (N + BLOCK - 1) // BLOCK * 2)
            kc = torch.zeros(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda")
            vc = torch.zeros_like(kc)
            kc2, vc2 = torch.zeros_like(kc), torch.zeros_like(kc)
            slots = torch.randperm(nb * BLOCK, generator=g, device="cuda")[:N].to(torch.int32)
            alg = N * row * 2 * 2 * 2 + 4 * N
            t_min, t_med = timed(lambda: store_kvcache(k, v, kc, vc, slots), iters=20)
            f_min, f_med = timed(lambda: bar_kernel[(N,)](k, k.stride(0), v, v.stride(0), kc2, vc2,
                                                          slots, row), iters=20)
            same = torch.equal(kc, kc2) and torch.equal(vc, vc2)
            rec = dict(kind="store", N=N, Hkv=Hkv, Hq=Hq, alg_bytes=alg, ours_us=t_med * 1e3,
                       ours_min_us=t_min * 1e3, ours_gbs=alg / t_med / 1e6,
                       triton_us=f_med * 1e3, triton_min_us=f_min * 1e3,
                       triton_gbs=alg / f_med / 1e6, bit_equal=same)
            out.write(json.dumps(rec) + "\n")
            out.flush()
            print(f"store N={N:6d} Hkv={Hkv}: ours {t_med * 1e3:7.1f} us {rec['ours_gbs']:7.0f} GB/s"
                  f" | triton {f_med * 1e3:7.1f} us {rec['triton_gbs']:7.0f} GB/s equal={same}",
                  flush=True)


PREFILL_SHAPES = [  # (lens list spec, cached, Hq, Hkv)
    ("4x4096", [4096] * 4, None, 32, 8), ("4x4096", [4096] * 4, None, 16, 8),
    ("16x1024", [1024] * 16, None, 16, 8), ("1x16384", [16384], None, 32, 8),
    ("2x8192", [8192] * 2, None, 8, 1), ("bench24", None, None, 16, 8),
    ("32x(7680+512)", [8192] * 32, [7680] * 32, 8, 1), ("8x(3840+256)", [4096] * 8, [3840] * 8, 32, 8),
    # short query blocks (max_seqlen_q < 256): remainders behind a cached prefix, short prompts
    ("short32x(1024+100)", [1124] * 32, [1024] * 32, 16, 8), ("short64x128", [128] * 64, None, 16, 8),
    ("short16x(2048+200)", [2248] * 16, [2048] * 16, 32, 8), ("short64x(256+60)", [316] * 64, [256] * 64, 16, 8),
    ("short128x200", [200] * 128, None, 16, 8),
    ("short64x(512+100)", [612] * 64, [512] * 64, 16, 8), ("short64x(512+60)", [572] * 64, [512] * 64, 16, 8),
    ("short64x(1024+30)", [1054] * 64, [1024] * 64, 16, 8), ("short128x(512+16)", [528] * 128, [512] * 128, 32, 8),
    ("short32x(768+150)", [918] * 32, [768] * 32, 32, 8), ("short256x(256+20)", [276] * 256, [256] * 256, 16, 8),
    ("short64x160", [160] * 64, None, 32, 8), ("short8x100", [100] * 8, None, 16, 8),
]


def prefill_flops(lens, cached, Hq):
    fl = 0
    for n, c in zip(lens, cached):
        lq = n - c
        fl += 4 * Hq * D * (lq * c + lq * (lq + 1) // 2)
    return fl


def bench_prefill(out, quick):
    import random

    from flash_attn import flash_attn_varlen_func

    from tools import metadata_ref as md
    from byox_vllm_b200.attention import prefill_varlen

    shapes = PREFILL_SHAPES[:3] if quick else PREFILL_SHAPES
    only = [a.split("=", 1)[1] for a in sys.argv if a.startswith("--only=")]
    no_ref = "--no-ref" in sys.argv  # ours alone (ncu captures)
    for name, lens, cached, Hq, Hkv in shapes:
        if only and not any(o in name for o in only):
            continue
        if lens is None:
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
            kc = torch.empty(nb, BLOCK, Hkv, D, dtype=torch.bfloat16,
                             device="cuda").normal_(generator=g)
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
        t_min, t_med = timed(lambda: prefill_varlen(q, k, v, cu_q, cu_k, m["max_seqlen_q"],
                                                    m["max_seqlen_k"], scale, block_tables=bt,
                                                    out=o), iters=3 if no_ref else 8)
        fa = lambda: flash_attn_varlen_func(  # noqa: E731
            q, k, v, max_seqlen_q=m["max_seqlen_q"], cu_seqlens_q=cu_q,
            max_seqlen_k=m["max_seqlen_k"], cu_seqlens_k=cu_k, softmax_scale=scale, causal=True,
            block_table=bt)
        if no_ref:
            f_med, err = float("nan"), float("nan")
        else:
            f_min, f_med = timed(fa, iters=8)
            err = (o.float() - fa().float()).abs().max().item()
        rec = dict(kind="prefill", name=name, Hq=Hq, Hkv=Hkv, total_q=total_q, flops=fl,
                   ours_us=t_med * 1e3, ours_tflops=fl / t_med / 1e9,
                   ours_frac=fl / t_med / 1e9 / TF_PEAK, fa2_us=f_med * 1e3,
                   fa2_tflops=fl / f_med / 1e9, speedup=f_med / t_med, max_abs_diff_vs_fa2=err)
        out.write(json.dumps(rec) + "\n")
        out.flush()
        print(f"prefill {name:>14s} {Hq}/{Hkv}: ours {t_med * 1e3:9.1f} us {rec['ours_tflops']:7.1f}"
              f" TF/s ({rec['ours_frac']:.3f}) | fa2 {f_med * 1e3:9.1f} us {rec['fa2_tflops']:7.1f}"
              f" TF/s | x{rec['speedup']:.2f} err {err:.4f}", flush=True)
        del q, k, v, o
        torch.cuda.empty_cache()


def timed_graph(fn, iters=10, warm=3):
    """Capture fn once into a CUDA graph and time replays (decode runs under graphs in the engine:
    model_runner.py:242-254), L2 flushed before every replay."""
    fn()
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    return timed(g.replay, iters=iters, warm=warm)


def bench_qkv(out, quick):
    """the steps between qkv_proj and o_proj (models/qwen3.py:75-84). Ours: one pass
    (prefill sizes) / one launch (decode). Reference chain: its own RMSNorm x2 + RotaryEmbedding
    (torch.compile'd, from baseline/_ref) + Triton store_kvcache (+ flash_attn_with_kvcache)."""
    sys.path.insert(0, str(ROOT / "baseline" / "_ref"))
    from nanovllm.layers.attention import store_kvcache as ref_store
    from nanovllm.layers.layernorm import RMSNorm
    from nanovllm.layers.rotary_embedding import RotaryEmbedding
    from flash_attn import flash_attn_with_kvcache

    from byox_vllm_b200 import qkv_attention as fused
    from byox_vllm_b200.attention import paged_decode

    eps, max_pos = 1e-6, 40960
    torch.set_default_device("cuda")
    rope = RotaryEmbedding(D, D, max_pos, 1e6)
    torch.set_default_dtype(torch.bfloat16)
    qn, kn = RMSNorm(D, eps), RMSNorm(D, eps)
    torch.set_default_dtype(torch.float32)
    torch.set_default_device("cpu")
    g = torch.Generator(device="cuda").manual_seed(0)
    qn.weight.data.normal_(1.0, 0.25, generator=g)
    kn.weight.data.normal_(1.0, 0.25, generator=g)
    cache = rope.cos_sin_cache

    def ref_prep(qkv, pos, Hq, Hkv):
        q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        q, k, v = q.view(-1, Hq, D), k.view(-1, Hkv, D), v.view(-1, Hkv, D)
        q, k = qn(q), kn(k)
        q, k = rope(pos, q, k)
        return q, k, v

    # ---- prefill-size pass: norm + rope + scatter -------------------------------------------
    for N, Hq, Hkv in ([(16384, 32, 8)] if quick else
                       [(16384, 32, 8), (16384, 16, 8), (16384, 8, 1), (4096, 32, 8), (1024, 16, 8)]):
        qkv = torch.empty(N, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16,
                          device="cuda").normal_(generator=g)
        pos = torch.randint(0, max_pos, (N,), generator=g, device="cuda")
        nb = (N + BLOCK - 1) // BLOCK * 2
        kc = torch.zeros(nb, BLOCK, Hkv, D, dtype=torch.bfloat16, device="cuda")
        vc, kc2, vc2 = torch.zeros_like(kc), torch.zeros_like(kc), torch.zeros_like(kc)
        slots = torch.randperm(nb * BLOCK, generator=g, device="cuda")[:N].to(torch.int32)
        work = qkv.clone()

        def ours():
            return fused.qk_norm_rope_store(work, pos, qn.weight, kn.weight, eps, cache, Hq, Hkv,
                                            D, k_cache=kc, v_cache=vc, slot_mapping=slots)

        def ref():
            q, k, v = ref_prep(qkv, pos, Hq, Hkv)
            ref_store(k, v, kc2, vc2, slots)
            return q, k

        work.copy_(qkv)
        q1, k1, _ = ours()
        q2, k2 = ref()
        torch.cuda.synchronize()
        dq = (q1.float() - q2.float()).abs().max().item()
        mism = (q1.view(torch.int16) != q2.view(torch.int16)).float().mean().item()
        v_same = torch.equal(vc, vc2)
        # algorithmic bytes (k written in place too, as dense prefill needs it)
        alg = N * ((Hq + 2 * Hkv) * 256 + Hq * 256 + Hkv * 256 + 2 * Hkv * 256 + 12)
        t_min, t_med = timed(ours, iters=10)
        f_min, f_med = timed(ref, iters=10)
        rec = dict(kind="qkv_prep", N=N, Hq=Hq, Hkv=Hkv, alg_bytes=alg, ours_us=t_med * 1e3,
                   ours_min_us=t_min * 1e3, ours_gbs=alg / t_med / 1e6,
                   ours_frac=alg / t_med / 1e6 / PEAK, ref_us=f_med * 1e3, speedup=f_med / t_med,
                   q_max_abs_diff=dq, q_mismatch_frac=mism, v_cache_equal=v_same)
        out.write(json.dumps(rec) + "\n")
        out.flush()
        print(f"qkv_prep N={N:6d} {Hq}/{Hkv}: ours {t_med * 1e3:7.1f} us {rec['ours_gbs']:6.0f} GB/s"
              f" ({rec['ours_frac']:.2f}) | ref chain {f_med * 1e3:7.1f} us | x{rec['speedup']:.2f}"
              f" q diff {dq:.4f} mismatch {mism:.2e} v equal={v_same}", flush=True)
        del qkv, work, kc, vc, kc2, vc2
        torch.cuda.empty_cache()

    # ---- decode step of one layer, graph-replayed -------------------------------------------
    for B, ctx, Hq, Hkv in ([(256, 1024, 16, 8)] if quick else
                            [(256, 1024, 16, 8), (128, 4096, 32, 8), (256, 8192, 8, 1),
                             (64, 4096, 32, 8), (16, 2048, 16, 8), (1, 4096, 32, 8)]):
        _, kc, vc, lens, bt, alg = decode_inputs(B, ctx, Hq, Hkv, ragged=True, seed=B)
        qkv = torch.empty(B, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16,
                          device="cuda").normal_(generator=g)
        pos = (lens - 1).long()
        last = (lens - 1)
        slots = (bt.gather(1, (last // BLOCK).long().unsqueeze(1)).squeeze(1) * BLOCK +
                 last % BLOCK).to(torch.int32)
        scale = D ** -0.5
        o1 = torch.empty(B, Hq, D, dtype=torch.bfloat16, device="cuda")
        o2 = torch.empty_like(o1)

        def ours():
            fused.qkv_decode(qkv, pos, qn.weight, kn.weight, eps, cache, kc, vc, lens, bt, slots,
                             scale, Hq, Hkv, out=o1)

        def ours_boundary():  # frozen Attention.forward boundary: reference prep + our fused call
            q, k, v = ref_prep(qkv, pos, Hq, Hkv)
            paged_decode(q, kc, vc, lens, bt, scale, key=k, value=v, slot_mapping=slots, out=o2)

        def ref():
            q, k, v = ref_prep(qkv, pos, Hq, Hkv)
            ref_store(k, v, kc, vc, slots)
            return flash_attn_with_kvcache(q.unsqueeze(1), kc, vc, cache_seqlens=lens,
                                           block_table=bt, softmax_scale=scale, causal=True)

        ours()
        o_ref = ref()
        torch.cuda.synchronize()
        err = (o1.float() - o_ref.squeeze(1).float()).abs().max().item()
        t_min, t_med = timed_graph(ours)
        b_min, b_med = timed_graph(ours_boundary)
        f_min, f_med = timed_graph(ref)
        alg2 = alg + B * (Hq + 2 * Hkv) * 256  # decode bytes + the qkv rows
        rec = dict(kind="qkv_decode", B=B, ctx=ctx, Hq=Hq, Hkv=Hkv, alg_bytes=alg2,
                   ours_us=t_med * 1e3, ours_gbs=alg2 / t_med / 1e6,
                   ours_frac=alg2 / t_med / 1e6 / PEAK, ours_boundary_us=b_med * 1e3,
                   ref_us=f_med * 1e3, speedup_vs_ref=f_med / t_med,
                   speedup_vs_boundary=b_med / t_med, max_abs_diff_vs_ref=err)
        out.write(json.dumps(rec) + "\n")
        out.flush()
        print(f"qkv_decode B={B:4d} ctx={ctx:6d} {Hq}/{Hkv}: one launch {t_med * 1e3:7.1f} us "
              f"{rec['ours_gbs']:6.0f} GB/s ({rec['ours_frac']:.2f}) | ref prep + ours "
              f"{b_med * 1e3:7.1f} us | ref chain {f_med * 1e3:7.1f} us | x{rec['speedup_vs_ref']:.2f}"
              f" err {err:.4f}", flush=True)
        del kc, vc
        torch.cuda.empty_cache()


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    quick = "--quick" in sys.argv
    kinds = args or ["decode", "store", "prefill"]
    out_dir = ROOT / "gpurun_out"
    out_dir.mkdir(exist_ok=True)
    tag = "_".join(kinds)
    with torch.inference_mode(), open(out_dir / f"microbench_{tag}.jsonl", "w") as out:
        out.write(json.dumps(dict(kind="env", gpu=torch.cuda.get_device_name(0), hbm_peak=PEAK,
                                  tf_peak=TF_PEAK, torch=torch.__version__)) + "\n")
        for k in kinds:
            {"decode": bench_decode, "store": bench_store, "prefill": bench_prefill,
             "sweep": bench_sweep, "cascade": bench_cascade, "qkv": bench_qkv}[k](out, quick)


if __name__ == "__main__":
    main()
