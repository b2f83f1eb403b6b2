#!/bin/bash
# times the experiment variants of libtdssf.so under build/variants/
mkdir -p gpurun_out
: > gpurun_out/variants.jsonl
for lib in 3d-semantic-segmentation-fusion_b200/libtdssf.so 3d-semantic-segmentation-fusion_b200/build/variants/*/libtdssf.so; do
  case $lib in *PROFILE*) continue;; esac
  TDSSF_LIB=$PWD/$lib timeout 300 python tools/exp_fused.py --iters 20 >> gpurun_out/variants.jsonl 2>> gpurun_out/variants.err
done
python - <<'PY'
import json
for l in open('gpurun_out/variants.jsonl'):
    d=json.loads(l); print(d['tag'].split('/')[-2][:40].ljust(40), 'fused %.4f ms  dense %.4f ms' % (d['fused_ms'], d['mlp_tc_dense_ms']))
PY
