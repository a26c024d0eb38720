mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_qkv_fused.py tests/test_gpu_variants.py -m gpu -x -q > gpurun_out/pytest_qkv.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_qkv.log )
( timeout 300 python tools/microbench.py qkv > gpurun_out/microbench_qkv.log 2>&1 )
( B200ATTN_NORMROPE_TOKENS=1 B200ATTN_NORMROPE=bulk timeout 300 python tools/microbench.py qkv > gpurun_out/microbench_qkv_t1.log 2>&1 )
tail -3 gpurun_out/pytest_qkv.log; grep qkv_prep gpurun_out/microbench_qkv.log | cut -c1-110; echo T1; grep qkv_prep gpurun_out/microbench_qkv_t1.log | cut -c1-110
