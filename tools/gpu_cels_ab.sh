#!/bin/bash
# CE/Lovasz loss: parity tests of the default build, then timing of the default build against the named variants
mkdir -p gpurun_out
V=$PWD/3d-semantic-segmentation-fusion_b200/build/variants
timeout 600 python -m pytest tests -m gpu -x -q -k "cels or lovasz or loss or sort" > gpurun_out/cels_pytest.log 2>&1; tail -2 gpurun_out/cels_pytest.log
for mode in randn peaked; do
  timeout 300 python tools/cels_probe.py $mode
  for tag in "$@"; do TDSSF_LIB=$V/$tag/libtdssf.so timeout 300 python tools/cels_probe.py $mode; done
done
