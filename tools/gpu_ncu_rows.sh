#!/bin/bash
# one ncu --set full pass over the stand-alone kernels of tools/bench_rows.py (first launch of each), after a plain run
mkdir -p gpurun_out
timeout 600 python tools/bench_rows.py --iters 3 > gpurun_out/rows_plain.jsonl 2> gpurun_out/rows_plain.err || exit 1
timeout 1500 ncu --set full --clock-control none --import-source on \
  -k regex:"project_kernel|gather_mean_kernel|mlp_fwd_f32|mlp_bwd|devox_grid_query|devox_nn_kernel|cels_|confusion_kernel|lovasz_|sort_|scan_ints" \
  -c 40 -o gpurun_out/prof_rows python tools/bench_rows.py --iters 1 > gpurun_out/ncu_rows.log 2>&1
tail -3 gpurun_out/ncu_rows.log
ls -la gpurun_out/prof_rows.ncu-rep
