mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log )
for i in 1 2; do
( B200ATTN_FAST_META=0 timeout 300 python tools/e2e_nanovllm.py --impl ours --model 0.6b --num-seqs 64 --check --out gpurun_out/e2e_det_off$i.json > gpurun_out/e2e_det_off$i.log 2>&1 )
done
( timeout 300 python tools/e2e_nanovllm.py --impl ours --model 0.6b --num-seqs 64 --check --out gpurun_out/e2e_det_on.json > gpurun_out/e2e_det_on.log 2>&1 )
tail -4 gpurun_out/pytest_gpu.log
