mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log )
( timeout 300 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err )
( timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err )
( timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1 )
( timeout 300 ncu --set full --clock-control none --import-source on -k regex:decode_split -s 30 -c 1 -o gpurun_out/prof_decode_c2 python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_full.log 2>&1 )
tail -3 gpurun_out/pytest_gpu.log; cut -c1-700 gpurun_out/bench_c2.json; cut -c1-500 gpurun_out/bench_ref.json; tail -2 gpurun_out/ncu_full.log
