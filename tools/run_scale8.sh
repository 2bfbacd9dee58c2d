# on an 8-GPU box: weak N=8 / 4 / 2, strong 1080p spp 8 and 3840x2160 spp 8 at N=8 (the N=1 lines come from 1-GPU runs)
run() { n=$1; name=$2; shift 2; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+RANDOM%200)) bench.py --gpus $n --steps 20 --warmup 3 --no-cpu-baseline --no-gpu-baseline "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; cut -c1-160 gpurun_out/$name.json; }
run 8 weak_8gpu
run 8 strong_1080p_spp8_8gpu --scaling strong --spp 8
run 8 strong_4k_spp8_8gpu --scaling strong --spp 8 --width 3840 --height 2160
run 2 weak_2gpu
