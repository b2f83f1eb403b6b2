#!/bin/bash
# ncu --set full, first launch of each remaining stand-alone kernel (fp32 MLP forward/backward, confusion, devoxelize gather,
# multi-sweep accumulation, voxelizer 5-NN), one ncu run per kernel over tools/bench_rows.py --iters 1
mkdir -p gpurun_out
for k in ${KERNELS:-gemm_rows_kernel wgrad_kernel knn5_kernel}; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$k" -c 1 -o gpurun_out/prof_more_$k \
    python tools/bench_rows.py --iters 1 > gpurun_out/ncu_more_$k.log 2>&1
  tail -1 gpurun_out/ncu_more_$k.log
done
ls -la gpurun_out/prof_more_*.ncu-rep
