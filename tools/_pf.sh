timeout 600 python -m pytest tests/test_gpu_prefill.py tests/test_gpu_cascade.py tests/test_gpu_attention_module.py -x -q 2>&1 | tail -5
for m in clc static grid; do
  B200ATTN_PREFILL_SCHED=$m timeout 200 python bench.py --workload prefill > gpurun_out/r02_prefill_sched_$m.json 2> gpurun_out/err_$m.txt; tail -2 gpurun_out/err_$m.txt
  python - <<EOF
import json
d=json.loads(open("gpurun_out/r02_prefill_sched_$m.json").read())["extra"]
print("== $m")
for k,v in d.items(): print(k, "ms %.4f frac %.3f vsFA %.2f diff %.4f"%(v["ms_per_launch"],v["frac"],v.get("vs_flash_attn",0),v.get("max_abs_diff_vs_flash_attn",-1)))
EOF
done
