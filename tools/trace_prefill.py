"""Phase timeline of the tcgen05 prefill kernel (debug build selected by b200attn_debug_prefill_trace).

The first 8 CTAs stamp their SM clock at the phase boundaries of every tile step, per role:
  softmax group i : S ready | S row loaded | P computed + stored | P signalled
  issuer          : tiles ready | P_0 ready | PV_0 issued | S_0 of the next tile issued
This prints, per shape, the median clocks of each phase and of the step period, i.e. where the
S -> softmax -> PV chain spends its time.

    gpurun -- python tools/trace_prefill.py > gpurun_out/trace_prefill.log
"""
from __future__ import annotations

import ctypes
import json
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

CTAS, STREAMS, STEPS = 8, 6, 128
SHAPES = [  # name, lens, Hq, Hkv
    ("1x16384 32/8", [16384], 32, 8),
    ("4x4096 32/8", [4096] * 4, 32, 8),
    ("16x1024 16/8", [1024] * 16, 16, 8),
    ("bench24 16/8", None, 16, 8),
]


def med(xs):
    xs = [float(x) for x in xs]
    return round(statistics.median(xs), 0) if xs else None


def main():
    import random

    from byox_vllm_b200 import _lib
    from byox_vllm_b200.attention import prefill_varlen

    lib = _lib.load()
    rows = []
    with torch.inference_mode():
        for name, lens, Hq, Hkv in SHAPES:
            if lens is None:
                rng = random.Random(0)
                lens = [rng.randint(100, 1024) for _ in range(24)]
            total = sum(lens)
            g = torch.Generator(device="cuda").manual_seed(1)
            q = torch.empty(total, Hq, 128, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
            k = torch.empty(total, Hkv, 128, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
            v = torch.empty(total, Hkv, 128, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
            cu = torch.tensor([0] + list(__import__("itertools").accumulate(lens)), dtype=torch.int32,
                              device="cuda")
            o = torch.empty_like(q)
            scale = 128 ** -0.5
            for _ in range(2):
                prefill_varlen(q, k, v, cu, cu, max(lens), max(lens), scale, out=o)
            torch.cuda.synchronize()
            trace = torch.zeros(CTAS * STREAMS * STEPS * 4, dtype=torch.int64, device="cuda")
            lib.b200attn_debug_prefill_trace(ctypes.c_void_p(trace.data_ptr()))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            prefill_varlen(q, k, v, cu, cu, max(lens), max(lens), scale, out=o)
            e1.record()
            torch.cuda.synchronize()
            lib.b200attn_debug_prefill_trace(ctypes.c_void_p(0))
            t = trace.view(CTAS, STREAMS, STEPS, 4).cpu().numpy()
            rec = {"shape": name, "event_us": round(e0.elapsed_time(e1) * 1e3, 1)}
            for i in (0, 1):
                a = t[:, i]
                ok = (a[:, :, 3] > 0) & (a[:, :, 0] > 0)
                rec[f"softmax{i}"] = {
                    "steps": int(ok.sum()),
                    "ld_S": med((a[:, :, 1] - a[:, :, 0])[ok]),
                    "max_exp_st": med((a[:, :, 2] - a[:, :, 1])[ok]),
                    "signal": med((a[:, :, 3] - a[:, :, 2])[ok]),
                    "busy": med((a[:, :, 3] - a[:, :, 0])[ok]),
                    # from the P signal of step n to S ready of step n + 1 (PV + next S + latencies)
                    "wait_next_S": med((a[:, 1:, 0] - a[:, :-1, 3])[ok[:, 1:] & ok[:, :-1]]),
                    "period": med((a[:, 1:, 0] - a[:, :-1, 0])[ok[:, 1:] & ok[:, :-1]]),
                }
            m = t[:, 2]
            ok = (m[:, :, 3] > 0) & (m[:, :, 0] > 0) & (m[:, :, 1] > 0) & (m[:, :, 2] > 0)
            rec["issuer"] = {
                "steps": int(ok.sum()),
                "wait_P0": med((m[:, :, 1] - m[:, :, 0])[ok]),
                "issue_PV0_8mma": med((m[:, :, 2] - m[:, :, 1])[ok]),
                "issue_S0_8mma_commit": med((m[:, :, 3] - m[:, :, 2])[ok]),
                "period": med((m[:, 1:, 0] - m[:, :-1, 0])[ok[:, 1:] & ok[:, :-1]]),
            }
            # stagger between the two softmax groups: S ready of group 1 minus S ready of group 0
            both = (t[:, 0, :, 0] > 0) & (t[:, 1, :, 0] > 0)
            rec["stagger_S1_minus_S0"] = med((t[:, 1, :, 0] - t[:, 0, :, 0])[both])
            # first 12 steps of CTA 0, relative clocks (S ready / P signalled per group, issuer step)
            base = int(t[0, :, 0, 0][t[0, :, 0, 0] > 0].min()) if (t[0, :, 0, 0] > 0).any() else 0
            rec["cta0_first_steps"] = [
                {"s0": [int(x - base) for x in t[0, 0, n]], "s1": [int(x - base) for x in t[0, 1, n]],
                 "mma": [int(x - base) for x in t[0, 2, n]]} for n in range(10)]
            rec["cta0_items"] = [
                {"issuer": [int(x - base) for x in t[0, 3, n]], "producer": [int(x - base) for x in t[0, 4, n]],
                 "softmax0": [int(x - base) for x in t[0, 5, n]]} for n in range(8)]
            # per traced CTA: prologue (entry -> roles start) and busy span (roles start -> last row
            # stored by softmax group 0), in clocks; the kernel itself took event_us
            spans = []
            for cta in range(CTAS):
                entry, start = int(t[cta, 4, STEPS - 1, 0]), int(t[cta, 4, STEPS - 1, 1])
                last = int(t[cta, 5, :, 3].max())
                n_items = int((t[cta, 5, :STEPS - 1, 3] > 0).sum())
                spans.append({"prologue": start - entry, "busy": last - start, "items": n_items})
            rec["cta_spans"] = spans
            rows.append(rec)
            print(json.dumps({k2: v2 for k2, v2 in rec.items() if k2 not in ("cta0_first_steps", "cta0_items", "cta_spans")}), flush=True)
    (ROOT / "gpurun_out").mkdir(exist_ok=True)
    (ROOT / "gpurun_out" / "trace_prefill.json").write_text(json.dumps(rows, indent=1))


if __name__ == "__main__":
    main()
