N=8
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r02_bench_c2_n$N.json 2> gpurun_out/err_n$N.txt; tail -c 300 gpurun_out/err_n$N.txt
CUDA_VISIBLE_DEVICES=0,1,2,3 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r02_bench_c2_n4.json 2> gpurun_out/err_n4.txt
for arm in ref full; do
  case $arm in
    ref) impl=ref; env="";;
    full) impl=ours; env="";;
  esac
  env $env timeout 400 python tools/e2e_nanovllm.py --impl $impl --model 0.6b --tp $N --out gpurun_out/r02_e2e_06b_tp${N}_$arm.json > gpurun_out/e2e_tp${N}_$arm.log 2>&1
  grep '^{"tool"' gpurun_out/e2e_tp${N}_$arm.log | tail -1 | cut -c1-100 || tail -3 gpurun_out/e2e_tp${N}_$arm.log
done
