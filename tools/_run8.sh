mkdir -p gpurun_out
( timeout 500 python tools/e2e_nanovllm.py --impl ours --model 0.6b --tp 8 --out gpurun_out/e2e_06b_tp8_ours.json > gpurun_out/e2e_06b_tp8_ours.log 2>&1 )
( timeout 500 python tools/e2e_nanovllm.py --impl ref --model 0.6b --tp 8 --out gpurun_out/e2e_06b_tp8_ref.json > gpurun_out/e2e_06b_tp8_ref.log 2>&1 )
tail -c 500 gpurun_out/e2e_06b_tp8_ours.log; tail -c 500 gpurun_out/e2e_06b_tp8_ref.log
