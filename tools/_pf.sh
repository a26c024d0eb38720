timeout 300 python -m pytest tests/test_gpu_decode.py tests/test_gpu_cascade.py tests/test_gpu_guard_bands.py -x -q 2>&1 | tail -2
for d in 0 3 6 12; do echo "== prefetch $d"; B200ATTN_DECODE_PREFETCH=$d timeout 200 python tools/sweep_ragged_splits.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: continue
    print(r['B'], r['ctx'], r['heads'], 'ov' if r['overlap'] else '--', r['us']['None'], r['auto_frac'])
"; done
