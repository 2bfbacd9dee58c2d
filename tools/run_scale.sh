# weak-scaling bench lines on one 8-GPU box: N = 2, 4, 8 (N = 1 is the default bench run)
for n in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/bench_${n}gpu.json 2> gpurun_out/bench_${n}gpu.err
  cut -c1-200 gpurun_out/bench_${n}gpu.json
done
