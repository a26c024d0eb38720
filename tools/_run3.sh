mkdir -p gpurun_out
( timeout 400 python tools/microbench.py qkv > gpurun_out/microbench_qkv.log 2>&1; echo "exit $?" >> gpurun_out/microbench_qkv.log )
( timeout 300 ncu --set full --clock-control none --import-source on -k regex:qk_norm_rope -c 1 -o gpurun_out/qk_norm_rope_full python tools/microbench.py qkv --quick > gpurun_out/ncu_qkv.log 2>&1 )
( timeout 400 python tools/e2e_nanovllm.py --impl ours --model 8b --num-seqs 128 --fixed-in 4032 --fixed-out 64 --max-model-len 4224 --out gpurun_out/e2e_8b_fused.json > gpurun_out/e2e_8b_fused.log 2>&1 )
( B200ATTN_FUSE_QKV=0 timeout 400 python tools/e2e_nanovllm.py --impl ours --model 8b --num-seqs 128 --fixed-in 4032 --fixed-out 64 --max-model-len 4224 --out gpurun_out/e2e_8b_unfused.json > gpurun_out/e2e_8b_unfused.log 2>&1 )
tail -13 gpurun_out/microbench_qkv.log; tail -c 400 gpurun_out/e2e_8b_fused.log; tail -c 400 gpurun_out/e2e_8b_unfused.log
