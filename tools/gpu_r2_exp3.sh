#!/bin/bash
# round 2, experiment 3: camera means streamed behind the tile loop (helper warps) against the all-up-front prologue
mkdir -p gpurun_out; O=gpurun_out
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
timeout 900 python -m pytest tests -m gpu -x -q -k "fus or tc or smoke" > $O/pytest_stream.log 2>&1; tail -3 $O/pytest_stream.log
for pts in normal invalid; do
  timeout 300 python tools/exp_fused.py --fused-only --points $pts
  for tag in "$@"; do
    TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --points $pts --tag $tag 2> $O/var_$tag.err || tail -3 $O/var_$tag.err
  done
done
timeout 300 python tools/exp_fused.py --fused-only --tok bfloat16
TDSSF_LIB=$V/TDSSF_STREAM0/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --tok bfloat16 --tag TDSSF_STREAM0
timeout 300 python tools/exp_fused.py --fused-only --workload config4
TDSSF_LIB=$V/TDSSF_STREAM0/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --workload config4 --tag TDSSF_STREAM0
TDSSF_LIB=$V/TDSSF_PROFILE/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --iters 5 --tag TDSSF_PROFILE > $O/prof.json 2> $O/prof.err
grep -v Warning $O/prof.err | head -24
