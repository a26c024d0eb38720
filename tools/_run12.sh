mkdir -p gpurun_out
for w in 4 6 8; do
( B200ATTN_NORMROPE=bulk B200ATTN_NORMROPE_WARPS=$w timeout 300 python tools/microbench.py qkv > gpurun_out/microbench_qkv_w$w.log 2>&1 )
echo "warps $w"; grep qkv_prep gpurun_out/microbench_qkv_w$w.log | cut -c1-80
done
