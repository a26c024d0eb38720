for cfg in "rows 2" "stream 2" "stream 1"; do set -- $cfg; echo "== engine $1 ctas/sm $2"; B200ATTN_DECODE_ENGINE=$1 B200ATTN_DECODE_CTAS_PER_SM=$2 timeout 200 python tools/sweep_ragged_splits.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: continue
    print(r['B'], r['ctx'], r['heads'], 'ov' if r['overlap'] else '--', r['us']['None'], r['auto_frac'])
"; done
