"""Regenerates BASELINE.md §5 (round-2 driver-shaped run) from the committed JSON lines under
profiles/: r02_bench_c2_default.json, r02_bench_c2_reference_arm.json, r02_bench_c2_n{2,4,8}.json.

    python tools/make_baseline_r02.py        # rewrites the section in place
"""
import json
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
MARK = "## 5. Round 2: the driver-shaped run"


def line(path):
    return json.loads((ROOT / "profiles" / path).read_text().strip().splitlines()[-1])


def main():
    d, ref = line("r02_bench_c2_default.json"), line("r02_bench_c2_reference_arm.json")
    rows = []
    for k, v in d["extra"].items():
        rate = f"{v['achieved']:.0f} GB/s" if v["bound"] == "hbm" else f"{v['achieved']:.0f} TFLOP/s"
        vs = v.get("vs_flash_attn") or v.get("vs_reference") or float("nan")
        other = (v.get("flash_attn_ms") or v.get("reference_triton_ms") or
                 v.get("reference_chain_ms") or float("nan"))
        rows.append(f"| `{k}` | {v.get('workload', k)} | {v['ms_per_launch'] * 1e3:.1f} | {rate} | "
                    f"{v['frac']:.2f} | {other * 1e3:.1f} | {vs:.2f}× |")
    e = d["engine_e2e"]
    scale = []
    for n in (2, 4, 8):
        s = line(f"r02_bench_c2_n{n}.json")
        scale.append(f"| {n} | {s['value']:.0f} | {s['value'] / d['value'] / n:.2f} | {s['e2e']['value']:.0f} | "
                     f"{s['e2e']['value'] / d['e2e']['value'] / n:.2f} | {s['roofline']['frac']:.2f} | "
                     f"{s['vs_flash_attn']:.2f}× |")
    txt = f"""{MARK} (one call on one B200: `pytest -m gpu`, `smoke()`, then `python bench.py`)

Source: `profiles/r02_bench_c2_default.json` (the JSON line `bench.py --steps 20 --warmup 3` printed, same
`gpurun` call as `profiles/r02_gputests.log`: 222 passed) and `profiles/r02_bench_c2_reference_arm.json`
(`bench.py --impl reference`). Peaks of that pod: HBM {d['roofline']['peak']:.1f} GB/s, bf16 burst 1628.7 TF/s
(`MEASURED_PEAKS.json`). flash-attn = the reference's own module from `baseline/_ref` / flash-attn 2.8.3 on the
same tensors in the same process, timed the same way, never inside ours' timed region. This section is generated
by `tools/make_baseline_r02.py` from the committed JSON lines.

| key of the JSON line | value |
|---|---|
| `value` (device-resident, 28-layer c2 step, graph replay) | **{d['value']:.0f} tok/s**, {d['ms_per_step']:.3f} ms/step |
| `e2e` (pinned host slab per 4 layers in, outputs out, inside the timing) | **{d['e2e']['value']:.0f} tok/s** ({d['e2e']['h2d_bytes_per_step'] / 1e6:.1f} MB in, {d['e2e']['d2h_bytes_per_step'] / 1e6:.1f} MB out per step) |
| `roofline` (decode kernel alone) | {d['roofline']['achieved']:.0f} GB/s = **{d['roofline']['frac']:.2f}** of the measured copy peak, {d['roofline']['frac_of_8TBs_nominal']:.2f} of 8 TB/s; ncu traffic 897 MB vs 861 MB algorithmic |
| `flash_attn` leg (reference module, same state, graph replay) | {d['flash_attn']['tokens_per_s']:.0f} tok/s → `vs_flash_attn` **{d['vs_flash_attn']:.2f}×** |
| `engine_e2e` reference / frozen boundary / all (f) rows | {e['ref_tok_s']:.0f} / **{e['ours_frozen_tok_s']:.0f} ({e['frozen_vs_ref']:.2f}×)** / **{e['ours_full_tok_s']:.0f} ({e['full_vs_ref']:.2f}×)** tok/s (decode {e['ref']['decode_tok_s']:.0f} / {e['ours_frozen']['decode_tok_s']:.0f} / {e['ours_full']['decode_tok_s']:.0f}; prefill {e['ref']['prefill_tok_s']:.0f} / {e['ours_frozen']['prefill_tok_s']:.0f} / {e['ours_full']['prefill_tok_s']:.0f}) |
| `cpu_baseline` (oracle, {d['cpu_baseline']['cores']} host threads, 1 layer × 64 sequences scaled) | {d['cpu_baseline']['value']:.1f} tok/s; `--impl reference` {ref['value']:.1f} tok/s on the same sample — its `ms_per_step` ({ref['ms_per_step']:.0f} ms) is ONE sampled pass extrapolated to 28 layers × 256 sequences, which is why the driver's `fits_in_driver_run` is false for that arm by construction |

`extra` block (kernel-level, L2 flushed or > L2 of independent inputs; "other" = flash-attn 2.8.3, for the store the
reference's Triton kernel, for the qkv pass the reference's norm + RoPE + store chain):

| key | workload | ours µs | achieved | of measured peak | other µs | ours ÷ other |
|---|---|---|---|---|---|---|
""" + "\n".join(rows) + f"""

Scaling on one 8×B200 node (torchrun, the driver's command line; `profiles/r02_bench_c2_n{{2,4,8}}.json`):

| N | value tok/s | efficiency | e2e tok/s | e2e efficiency | decode roofline frac | vs flash-attn module |
|---|---|---|---|---|---|---|
| 1 | {d['value']:.0f} | 1.00 | {d['e2e']['value']:.0f} | 1.00 | {d['roofline']['frac']:.2f} | {d['vs_flash_attn']:.2f}× |
""" + "\n".join(scale) + """

Engine end to end at TP > 1 (unmodified `baseline/_ref`, `tools/e2e_nanovllm.py`, same `gpurun` call per row; the
ours arms run on `install()` alone — `fast_rpc` is the product's transport — the reference arm needs the harness's
128 MiB block, token-free decode copies and the `run` fix to take a step at all): config 2 (Qwen3-0.6B) TP=2
reference 15 059 / frozen boundary 17 236 / all (f) rows 19 358 tok/s; TP=4 15 628 / — / 20 459; TP=8 15 161 / — /
19 020; config 3
(Qwen3-8B, 128 × (4032 + 64)) TP=1 868 / 1020 / 1064 tok/s, decode 5370 / 7087 / 7260 (1.35×), prefill 65.3k /
75.2k / 78.7k; TP=2 1427 / — / 1713 tok/s, decode 7971 / — / 10 343; TP=8 2548 / — / 3104 tok/s, decode 11 831 / — / 14 724; config 4's batch at TP=1
(Qwen3-32B, 256 × (7680 shared + 512) + 64 out, same box) **776 / — / 1478 tok/s, decode 2086 / — / 7330 (3.5×, the
shared-prefix cascade), prefill 157.8k / — / 238.5k**; config 4 as BASELINE.json names it, **TP=8: 4825 tok/s,
decode 12 899, prefill 1.02 M tok/s on `install()` alone** (ours arm only: the GPU budget of the round did not
cover the reference arm at 8 GPUs) (`profiles/r02_e2e_*`).

Round-2 changes behind these numbers: persistent prefill with hardware work stealing, compile-time MMA operands
and an item ring (DESIGN §3.3; bench24 100 → 74 µs, 1×16384 1767 → 1707 µs), decode overlap mode with pre-wait tile
requests (DESIGN §3.2; TP=8 per-rank shape 0.74 → 0.77–0.80 of HBM), slab-staged e2e leg (e2e/value 0.86 → 0.90,
N=8 e2e efficiency 0.26 → 0.60), fp16 (tests only; every BASELINE model is bf16).
"""
    p = ROOT / "BASELINE.md"
    s = p.read_text()
    if MARK in s:
        s = s[:s.index(MARK)]
    p.write_text(s.rstrip("\n") + "\n\n" + txt)


if __name__ == "__main__":
    main()
