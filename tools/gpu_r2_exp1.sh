#!/bin/bash
# round 2, experiment 1: how fast do the producers run alone (no epilogue math, no MMAs)?
mkdir -p gpurun_out
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
for pts in normal invalid hot; do
  timeout 300 python tools/exp_fused.py --fused-only --points $pts
  for tag in "$@"; do
    TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --points $pts --tag $tag 2> gpurun_out/var_$tag.err || tail -3 gpurun_out/var_$tag.err
  done
done
timeout 300 python tools/exp_fused.py --fused-only --iters 20000 --tag sustained20k
