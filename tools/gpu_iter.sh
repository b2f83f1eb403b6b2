#!/bin/bash
# one GPU iteration: tensor-core tests, experiment harness (default, fp16 tokens, config4), named variants, profile variant
mkdir -p gpurun_out
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
timeout 900 python -m pytest tests/test_gpu_tc.py -m gpu -x -q > gpurun_out/pytest_tc.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_tc.log
tail -15 gpurun_out/pytest_tc.log
timeout 300 python tools/exp_fused.py > gpurun_out/exp_default.json 2> gpurun_out/exp_default.err; cat gpurun_out/exp_default.json; tail -3 gpurun_out/exp_default.err
timeout 300 python tools/exp_fused.py --tok float16 > gpurun_out/exp_f16.json 2> gpurun_out/exp_f16.err; cat gpurun_out/exp_f16.json
timeout 300 python tools/exp_fused.py --workload config4 > gpurun_out/exp_c4.json 2> gpurun_out/exp_c4.err; cat gpurun_out/exp_c4.json
for tag in "$@"; do
  echo "== variant $tag"
  TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/exp_fused.py > gpurun_out/exp_$tag.json 2> gpurun_out/exp_$tag.err; cat gpurun_out/exp_$tag.json; tail -2 gpurun_out/exp_$tag.err
done
TDSSF_LIB=$V/TDSSF_PROFILE/libtdssf.so timeout 300 python tools/exp_fused.py --iters 5 > gpurun_out/exp_profile.json 2> gpurun_out/exp_profile.err
tail -40 gpurun_out/exp_profile.err
