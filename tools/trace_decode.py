"""Phase timeline of the decode kernel (debug build selected by b200attn_debug_decode_trace).

Every CTA stamps globaltimer at entry / exit and its SM clock at the phase boundaries listed in
csrc/decode_attn.cu (kTr*). This prints, per shape, the median / max over CTAs of each phase in
microseconds, so that fixed costs (scan, first descriptor, first TMA, first tile, epilogue, tail)
can be told apart from streaming time.

    gpurun -- python tools/trace_decode.py > gpurun_out/trace_decode.log
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

SHAPES = [  # (B, ctx or "bench", Hq, Hkv)
    (256, "bench", 16, 8), (256, "bench", 2, 1), (16, 1024, 32, 8), (1, 1024, 32, 8),
    (128, 4096, 32, 8), (64, 4096, 32, 8), (1, 131072, 32, 8),
]
NAMES = ["entry_ns", "entry", "scan", "piece0", "tma0", "tile0", "tiles_done0", "epi0", "prod_exit",
         "cons_exit", "pieces", "tiles", "sm", "wait", "exit_ns", "epi"]


def main():
    import bench
    from byox_vllm_b200 import _lib
    from byox_vllm_b200.attention import decode_num_splits, paged_decode

    lib = _lib.load()
    D = bench.D
    mhz = torch.cuda.clock_rate() if hasattr(torch.cuda, "clock_rate") else 1900
    out_rows = []
    with torch.inference_mode():
        for (B, ctx, Hq, Hkv) in SHAPES:
            lens_l = bench.context_lengths(dict(ctx=ctx, batch=B))
            lens, sets, alg, _ = bench.decode_inputs(B, lens_l, Hq, Hkv, 3)
            W = sets[0][3].size(1)
            n_cta = decode_num_splits(B, Hkv, W, bench.BLOCK)
            outs = [torch.empty_like(s[0]) for s in sets]
            scale = D ** -0.5
            for i in range(3):  # warm (product kernel)
                q, kc, vc, bt = sets[i]
                paged_decode(q, kc, vc, lens, bt, scale, out=outs[i])
            torch.cuda.synchronize()
            trace = torch.zeros(n_cta * 16, dtype=torch.int64, device="cuda")
            lib.b200attn_debug_decode_trace(ctypes.c_void_p(trace.data_ptr()))
            runs = []
            for rep in range(4):
                i = rep % 3
                q, kc, vc, bt = sets[i]
                bench.l2_flush()
                trace.zero_()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                paged_decode(q, kc, vc, lens, bt, scale, out=outs[i])
                e1.record()
                torch.cuda.synchronize()
                runs.append((e0.elapsed_time(e1) * 1e3, trace.view(n_cta, 16).cpu().numpy().copy()))
            lib.b200attn_debug_decode_trace(ctypes.c_void_p(0))
            us_evt, t = runs[-1]
            t0 = t[:, 0].min()
            span_ns = t[:, 14].max() - t0
            # SM clock rate from the CTAs themselves: clocks per ns over each CTA's life
            life_clk = (t[:, 9] - t[:, 1]).astype(float)
            life_ns = (t[:, 14] - t[:, 0]).astype(float)
            ghz = statistics.median((life_clk / life_ns.clip(1)).tolist())

            def us(col_a, col_b):
                d = (t[:, col_b] - t[:, col_a]).astype(float) / ghz / 1e3
                return {"med": round(statistics.median(d.tolist()), 2), "max": round(float(d.max()), 2),
                        "min": round(float(d.min()), 2)}

            rec = {
                "shape": f"{B}x{ctx} {Hq}/{Hkv}", "ctas": n_cta, "event_us": round(us_evt, 1),
                "span_us": round(span_ns / 1e3, 1), "clk_ghz": round(ghz, 3),
                "start_skew_us": round(float(t[:, 0].max() - t0) / 1e3, 2),
                "exit_spread_us": {"min": round(float(t[:, 14].min() - t0) / 1e3, 1),
                                   "med": round(float(statistics.median((t[:, 14] - t0).tolist())) / 1e3, 1),
                                   "max": round(float(t[:, 14].max() - t0) / 1e3, 1)},
                "scan": us(1, 2), "scan_to_piece0": us(2, 3), "piece0_to_tma0": us(3, 4),
                "tma0_to_tile0": us(4, 5), "entry_to_tile0": us(1, 5),
                "first_piece_tiles": us(5, 6), "first_epilogue": us(6, 7),
                "prod_exit_to_cons_exit": us(8, 9), "life": us(1, 9),
                "pieces_per_cta": {"med": int(statistics.median(t[:, 10].tolist())), "max": int(t[:, 10].max())},
                "tiles_per_cta": {"med": int(statistics.median(t[:, 11].tolist())), "max": int(t[:, 11].max()),
                                  "min": int(t[:, 11].min())},
                "wait_for_tiles_us": {"med": round(statistics.median((t[:, 13] / ghz / 1e3).tolist()), 2),
                                      "max": round(float((t[:, 13] / ghz / 1e3).max()), 2)},
                "epilogues_us": {"med": round(statistics.median((t[:, 15] / ghz / 1e3).tolist()), 2),
                                 "max": round(float((t[:, 15] / ghz / 1e3).max()), 2)},
                "alg_gbs_over_span": round(alg / max(span_ns, 1), 1),
            }
            out_rows.append(rec)
            print(json.dumps(rec), flush=True)
            del sets, outs
            torch.cuda.empty_cache()
    (ROOT / "gpurun_out").mkdir(exist_ok=True)
    (ROOT / "gpurun_out" / "trace_decode.json").write_text(json.dumps(out_rows, indent=1))


if __name__ == "__main__":
    main()
