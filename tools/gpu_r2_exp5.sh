#!/bin/bash
# correctness of the default build (tensor-core tests), then timing of the default build against named variants
mkdir -p gpurun_out; O=gpurun_out
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
timeout 900 python -m pytest tests -m gpu -x -q -k "fus or tc or smoke or abi" > $O/pytest_var.log 2>&1; tail -5 $O/pytest_var.log
for pts in normal invalid; do
timeout 300 python tools/exp_fused.py --fused-only --points $pts
for tag in "$@"; do
  TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --points $pts --tag $tag 2> $O/var_$tag.err || tail -3 $O/var_$tag.err
done
done
timeout 300 python tools/exp_fused.py --fused-only --workload config4
timeout 300 python tools/exp_fused.py --fused-only --tok bfloat16
