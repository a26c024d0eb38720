mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log )
( timeout 300 python tools/e2e_nanovllm.py --impl ours --model 0.6b --check --out gpurun_out/e2e_06b_fastmeta.json > gpurun_out/e2e_06b_fastmeta.log 2>&1 )
( B200ATTN_FAST_META=0 timeout 300 python tools/e2e_nanovllm.py --impl ours --model 0.6b --check --out gpurun_out/e2e_06b_nofastmeta.json > gpurun_out/e2e_06b_nofastmeta.log 2>&1 )
tail -4 gpurun_out/pytest_gpu.log; tail -c 300 gpurun_out/e2e_06b_fastmeta.log; tail -c 300 gpurun_out/e2e_06b_nofastmeta.log
