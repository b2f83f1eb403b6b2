#!/bin/bash
# fused kernel under input regimes that isolate the memory levels (profile build prints the cycle accounting)
mkdir -p gpurun_out
P=$PWD/3d-semantic-segmentation-fusion_b200/build/variants/TDSSF_PROFILE/libtdssf.so
for m in normal invalid hot sorted; do
  timeout 300 python tools/exp_fused.py --points $m --fused-only 2> gpurun_out/mode_$m.err
  TDSSF_LIB=$P timeout 300 python tools/exp_fused.py --points $m --fused-only --iters 5 2>&1 >/dev/null | sed -n 1,22p > gpurun_out/mode_${m}_profile.txt
done
timeout 300 python tools/exp_fused.py --points sorted --tok float16 --fused-only
