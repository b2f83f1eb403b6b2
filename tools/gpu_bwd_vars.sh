#!/bin/bash
# backward kernels: numerics of the default build, then kernel times of the timing-experiment variants
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
timeout 300 python tools/bwd_probe.py 2>&1 | grep -v -i warn | tail -8
for tag in "$@"; do
  echo "== $tag"
  TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/bwd_probe.py 2>&1 | grep " us x1" | head -4
done
