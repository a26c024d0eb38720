timeout 300 python -m pytest tests/test_gpu_store.py tests/test_gpu_variants.py tests/test_gpu_guard_bands.py -x -q 2>&1 | tail -2
for v in "16 4" "16 8" "32 4" "32 8"; do set -- $v; echo "== vec $1 tok $2"; B200ATTN_STORE_VEC=$1 B200ATTN_STORE_TOK=$2 timeout 100 python bench.py --workload store 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())['extra']
for k, v in d.items(): print(k, 'us %.2f frac %.3f vs_ref %.2f bit_equal %s' % (v['ms_per_launch']*1e3, v['frac'], v.get('vs_reference', 0), v.get('bit_equal')))
"; done
