#!/bin/bash
# A/B with repeats: default build against named variants, alternating, 3 rounds of 200 launches each
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
for round in 1 2 3; do
  timeout 300 python tools/exp_fused.py --fused-only --iters 200
  for tag in "$@"; do
    TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --iters 200 --tag $tag 2>/dev/null
  done
done
