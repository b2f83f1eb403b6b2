#!/bin/bash
# time the fused kernel (config 2) for the named build variants
mkdir -p gpurun_out
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
timeout 300 python tools/exp_fused.py --fused-only
for tag in "$@"; do
  TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --tag $tag 2> gpurun_out/var_$tag.err || tail -3 gpurun_out/var_$tag.err
done
