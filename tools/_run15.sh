mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log )
( timeout 300 python tools/microbench.py store > gpurun_out/microbench_store_bulk.log 2>&1 )
( B200ATTN_STORE=direct timeout 300 python tools/microbench.py store --quick > gpurun_out/microbench_store_direct.log 2>&1 )
tail -3 gpurun_out/pytest_gpu.log; grep "N=  2048\|N= 16384" gpurun_out/microbench_store_bulk.log | cut -c1-150; echo direct; grep "N= 16384" gpurun_out/microbench_store_direct.log | cut -c1-150
