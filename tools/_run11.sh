mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_qkv_fused.py -m gpu -x -q > gpurun_out/pytest_qkv.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_qkv.log )
( timeout 300 python tools/microbench.py qkv > gpurun_out/microbench_qkv_bulk.log 2>&1; echo "exit $?" >> gpurun_out/microbench_qkv_bulk.log )
( B200ATTN_NORMROPE=direct timeout 300 python tools/microbench.py qkv --quick > gpurun_out/microbench_qkv_direct.log 2>&1 )
tail -5 gpurun_out/pytest_qkv.log; grep qkv_prep gpurun_out/microbench_qkv_bulk.log gpurun_out/microbench_qkv_direct.log
