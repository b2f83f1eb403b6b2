#!/bin/bash
# round 2, experiment 2: where do the producers spend a slab period?  (profile build with producer phase counters, slab trace),
# what does fence.proxy.async cost (timing-only variant), deferred camera means with an L2 prefetch in the prologue
mkdir -p gpurun_out; O=gpurun_out
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
tools/micro/l2_ingest.bin > $O/l2_ingest.jsonl 2>&1; cat $O/l2_ingest.jsonl
for pts in normal invalid; do
  timeout 300 python tools/exp_fused.py --fused-only --points $pts
  for tag in "$@"; do
    TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --points $pts --tag $tag 2> $O/var_$tag.err || tail -3 $O/var_$tag.err
  done
done
timeout 300 python tools/exp_fused.py --fused-only --points sorted
timeout 300 python tools/exp_fused.py --fused-only --tok bfloat16
for tag in TDSSF_PROFILE TDSSF_PRO_DEFER_TDSSF_PROFILE; do
  TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/exp_fused.py --fused-only --iters 5 --tag $tag > $O/prof_$tag.json 2> $O/prof_$tag.err
  grep -v Warning $O/prof_$tag.err | head -70
done
TDSSF_LIB=$V/TDSSF_PRO_DEFER/libtdssf.so timeout 600 python -m pytest tests -m gpu -x -q -k "fus or tc" > $O/pytest_defer.log 2>&1; tail -2 $O/pytest_defer.log
