for m in 1 2; do
  B200ATTN_EXP_EMU=$m timeout 200 python bench.py --workload prefill > gpurun_out/r02_prefill_v6_emu_$m.json 2> gpurun_out/err_$m.txt; tail -2 gpurun_out/err_$m.txt
  python - <<EOF
import json
d=json.loads(open("gpurun_out/r02_prefill_v6_emu_$m.json").read())["extra"]
print("== emu $m")
for k,v in d.items(): print(k, "ms %.4f frac %.3f vsFA %.2f diff %.4f"%(v["ms_per_launch"],v["frac"],v.get("vs_flash_attn",0),v.get("max_abs_diff_vs_flash_attn",-1)))
EOF
done
