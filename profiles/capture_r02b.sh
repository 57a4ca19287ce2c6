#!/bin/bash
# ncu evidence of round 2, second half (5-operation on-chip kernel, variant 3 without a right-hand side and with
# the separator system in blocks).  Run on the GPU box from the repo root; outputs under gpurun_out/.
set -x
CMD="python bench.py --steps 2 --warmup 3 --no-long-legs --no-cpu-baseline"
$CMD > gpurun_out/r02b_plain.json 2> gpurun_out/r02b_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02b_launches_cfg3.csv $CMD > gpurun_out/r02b_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fused_tmem_pipe -s 2 -c 1 -o gpurun_out/r02b_fused_cfg3_full -f $CMD --no-e2e > gpurun_out/r02b_ncu_f.log 2>&1
python profiles/microbench/cfg5_steps.py > gpurun_out/r02b_cfg5_steps.txt 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 140 --csv --log-file gpurun_out/r02b_launches_cfg5.csv python profiles/microbench/cfg5_steps.py > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:tile_solve -s 5 -c 1 -o gpurun_out/r02b_tile_solve_cfg5 -f python profiles/microbench/cfg5_steps.py > gpurun_out/r02b_ncu_ts.log 2>&1
python profiles/microbench/theta_modes.py > gpurun_out/r02b_theta_modes.txt 2>&1
ls -la gpurun_out/*.ncu-rep
