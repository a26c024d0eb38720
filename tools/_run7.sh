mkdir -p gpurun_out
( timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "exit $?" >> gpurun_out/smoke.log )
( timeout 400 python tools/e2e_nanovllm.py --impl ours --model 0.6b --tp 2 --out gpurun_out/e2e_06b_tp2_ours.json > gpurun_out/e2e_06b_tp2_ours.log 2>&1 )
( timeout 400 python tools/e2e_nanovllm.py --impl ref --model 0.6b --tp 2 --out gpurun_out/e2e_06b_tp2_ref.json > gpurun_out/e2e_06b_tp2_ref.log 2>&1 )
tail -3 gpurun_out/smoke.log; tail -c 700 gpurun_out/e2e_06b_tp2_ours.log; tail -c 500 gpurun_out/e2e_06b_tp2_ref.log
