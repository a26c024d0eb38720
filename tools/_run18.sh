mkdir -p gpurun_out
( timeout 400 python tools/e2e_nanovllm.py --impl ours --model 0.6b --tp 4 --out gpurun_out/e2e_06b_tp4_ours.json > gpurun_out/e2e_06b_tp4_ours.log 2>&1 )
( timeout 400 python tools/e2e_nanovllm.py --impl ref --model 0.6b --tp 4 --out gpurun_out/e2e_06b_tp4_ref.json > gpurun_out/e2e_06b_tp4_ref.log 2>&1 )
tail -c 400 gpurun_out/e2e_06b_tp4_ours.log; tail -c 400 gpurun_out/e2e_06b_tp4_ref.log
