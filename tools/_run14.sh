mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log )
( timeout 300 python tools/microbench.py qkv > gpurun_out/microbench_qkv.log 2>&1; echo "exit $?" >> gpurun_out/microbench_qkv.log )
( timeout 300 ncu --set full --clock-control none --import-source on -k regex:qk_norm_rope -c 1 -o gpurun_out/qk_norm_rope_bulk_full python tools/microbench.py qkv --quick > gpurun_out/ncu_qkv.log 2>&1 )
tail -3 gpurun_out/pytest_gpu.log; grep qkv_ gpurun_out/microbench_qkv.log | cut -c1-150; tail -2 gpurun_out/ncu_qkv.log
