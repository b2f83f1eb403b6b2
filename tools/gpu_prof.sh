#!/bin/bash
mkdir -p gpurun_out
TDSSF_LIB=$PWD/3d-semantic-segmentation-fusion_b200/build/variants/TDSSF_PROFILE/libtdssf.so timeout 300 python tools/exp_fused.py --iters 5 > gpurun_out/exp_profile.json 2> gpurun_out/exp_profile.err
cat gpurun_out/exp_profile.err
