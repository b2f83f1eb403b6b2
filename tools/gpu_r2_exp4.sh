#!/bin/bash
# timing of named fused-kernel variants against the default build (config 2, normal points)
mkdir -p gpurun_out; O=gpurun_out
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
timeout 300 python tools/exp_fused.py --fused-only
for tag in "$@"; do
  TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --tag $tag 2> $O/var_$tag.err || tail -3 $O/var_$tag.err
done
