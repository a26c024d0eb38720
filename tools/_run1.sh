mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log )
( timeout 400 python tools/microbench.py qkv > gpurun_out/microbench_qkv.log 2>&1; echo "exit $?" >> gpurun_out/microbench_qkv.log )
( timeout 300 python tools/e2e_nanovllm.py --impl ours --model 0.6b --check --out gpurun_out/e2e_06b_fused.json > gpurun_out/e2e_06b_fused.log 2>&1 )
( B200ATTN_FUSE_QKV=0 timeout 300 python tools/e2e_nanovllm.py --impl ours --model 0.6b --check --out gpurun_out/e2e_06b_unfused.json > gpurun_out/e2e_06b_unfused.log 2>&1 )
( timeout 300 python tools/e2e_nanovllm.py --impl ours --model 0.6b --profile > gpurun_out/e2e_06b_profile.log 2>&1 )
tail -3 gpurun_out/pytest_gpu.log; tail -15 gpurun_out/microbench_qkv.log; tail -c 600 gpurun_out/e2e_06b_fused.log; tail -c 600 gpurun_out/e2e_06b_unfused.log
