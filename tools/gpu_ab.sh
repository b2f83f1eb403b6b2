#!/bin/bash
# A/B of fused-kernel build variants: parity tests of the default build first, then timings (normal / invalid / fp16 inputs)
mkdir -p gpurun_out
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
timeout 600 python -m pytest tests -m gpu -x -q -k "fus or tc" > gpurun_out/ab_pytest.log 2>&1; tail -2 gpurun_out/ab_pytest.log
for pts in normal invalid; do
  timeout 300 python tools/exp_fused.py --fused-only --points $pts
  for tag in "$@"; do
    TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --points $pts --tag $tag 2> gpurun_out/var_$tag.err || tail -3 gpurun_out/var_$tag.err
  done
done
timeout 300 python tools/exp_fused.py --fused-only --tok float16
timeout 300 python tools/exp_fused.py --fused-only --workload config4
