#!/bin/bash
# round 2, visit 3 (2 GPUs): the static-workspace entry point alone in a fresh process, the bench line at N ranks, then the whole GPU suite
N=${1:-2}
mkdir -p gpurun_out; O=gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_zz_alone.log 2>&1; echo "zz alone rc=$?"; tail -3 $O/pytest_zz_alone.log
bash tools/gpu_r2_multi.sh $N
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
