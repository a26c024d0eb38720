mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log )
( timeout 400 python tools/sweep_splits.py > gpurun_out/sweep_splits.log 2>&1; echo "exit $?" >> gpurun_out/sweep_splits.log )
( timeout 500 python tools/microbench.py sweep > gpurun_out/sweep_c5.log 2>&1; echo "exit $?" >> gpurun_out/sweep_c5.log )
tail -4 gpurun_out/pytest_gpu.log; cat gpurun_out/sweep_splits.log | cut -c1-330
