#!/bin/bash
# one ncu --set full capture of the fused kernel (bench.py, 3 steps) after a plain run of the same command
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fused_mlp -s 3 -c 1 -o gpurun_out/prof_fused python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
tail -c 300 gpurun_out/plain.log; tail -3 gpurun_out/ncu_full.log
