mkdir -p gpurun_out
( timeout 300 python tools/e2e_nanovllm.py --impl ours --model 0.6b --cprofile > gpurun_out/e2e_06b_cprofile.log 2>&1 )
( timeout 300 python tools/e2e_nanovllm.py --impl ref --model 0.6b --out gpurun_out/e2e_06b_ref.json > gpurun_out/e2e_06b_ref.log 2>&1 )
( timeout 300 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err )
tail -c 1500 gpurun_out/bench_c2.json; tail -c 500 gpurun_out/e2e_06b_ref.log
