#!/bin/bash
# correctness of the default build under a hard timeout, then A/B with repeats, then the profile build's prologue counters
mkdir -p gpurun_out; O=gpurun_out
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q > $O/pytest_q.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_q.log
bash tools/gpu_r2_exp6.sh "$@"
for w in config4; do python tools/exp_fused.py --fused-only --iters 100 --workload $w; for tag in "$@"; do TDSSF_LIB=$V/$tag/libtdssf.so python tools/exp_fused.py --fused-only --iters 100 --workload $w --tag $tag 2>/dev/null; done; done
python tools/exp_fused.py --fused-only --iters 100 --tok bfloat16
if [ -f $V/TDSSF_PROFILE/libtdssf.so ]; then
TDSSF_LIB=$V/TDSSF_PROFILE/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --iters 5 --tag TDSSF_PROFILE > $O/prof.json 2> $O/prof.err
grep -v Warning $O/prof.err | sed -n 11,20p
fi
