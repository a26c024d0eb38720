mkdir -p gpurun_out
( timeout 500 python tools/e2e_nanovllm.py --impl ours --model 32b --num-seqs 256 --fixed-in 8192 --fixed-out 64 --max-model-len 8448 --shared-prefix 7680 --out gpurun_out/e2e_32b_c4_fused.json > gpurun_out/e2e_32b_c4_fused.log 2>&1 )
( timeout 500 python tools/e2e_nanovllm.py --impl ref --model 32b --num-seqs 256 --fixed-in 8192 --fixed-out 64 --max-model-len 8448 --shared-prefix 7680 --out gpurun_out/e2e_32b_c4_ref.json > gpurun_out/e2e_32b_c4_ref.log 2>&1 )
tail -c 600 gpurun_out/e2e_32b_c4_fused.log; tail -c 500 gpurun_out/e2e_32b_c4_ref.log
