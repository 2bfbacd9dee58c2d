#!/bin/bash
# On the GPU box: the round's evidence set -> gpurun_out/ev_* (bench lines first, then the ncu passes of the same commands).
set -x
python bench.py > gpurun_out/ev_bench_neural.json 2> gpurun_out/ev_bench_neural.err
python bench.py --spp 8 --no-cpu-baseline > gpurun_out/ev_bench_neural_spp8.json 2> /dev/null
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/ev_bench_reference_arm.json 2> gpurun_out/ev_bench_reference_arm.err
# launch list of the same bench command (per-launch times are cold-cache and serialised: shares, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ev_launches_neural.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ev_ncu_launch.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ev_launches_gt.csv python bench.py --mode gt --steps 2 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ev_ncu_launch_gt.log 2>&1
# top kernels, full set, one warm frame each (frame_stages: 5 warm-up frames; skip their launches)
ncu --set full --clock-control none --import-source on -k regex:"k_bounce_begin|k_seg_exits|k_query_encode|k_query_mlp|k_entry_camera|k_path_tail|k_bounce_sample" --launch-skip 60 -c 10 -o gpurun_out/ev_top_neural -f python tools/frame_stages.py --steps 2 > gpurun_out/ev_ncu_full.log 2>&1
python tools/ncu_summary.py gpurun_out/ev_top_neural.ncu-rep > gpurun_out/ev_top_kernels_neural_ncu.txt
ncu --set full --clock-control none --import-source on -k regex:"k_gt_primary|k_gt_bounce|k_gt_tail" --launch-skip 12 -c 3 -o gpurun_out/ev_top_gt -f python tools/frame_stages.py --mode gt --steps 2 > gpurun_out/ev_ncu_full_gt.log 2>&1
python tools/ncu_summary.py gpurun_out/ev_top_gt.ncu-rep > gpurun_out/ev_top_kernels_gt_ncu.txt
