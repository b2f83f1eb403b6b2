#!/bin/bash
# ncu --set full over the CE/Lovasz and bf16-backward kernels of one training step (second step of tools/train_probe.py)
mkdir -p gpurun_out
timeout 300 python tools/train_probe.py 3 > gpurun_out/train_plain.log 2>&1 || { tail -5 gpurun_out/train_plain.log; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on \
  -k regex:"sort_|cels_|lovasz_|scan_ints|mask_cvt|cvt3_bf16|confusion" -s 23 -c 25 \
  -o gpurun_out/prof_train python tools/train_probe.py 2 > gpurun_out/ncu_train.log 2>&1
tail -2 gpurun_out/ncu_train.log; ls -la gpurun_out/prof_train.ncu-rep
