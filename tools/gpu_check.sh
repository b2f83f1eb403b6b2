#!/bin/bash
# One GPU-box visit: parity tests, bench, experiment harness, ncu launch list (the full capture is tools/gpu_ncu.sh: one ncu per box visit).
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 50 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench.json
python tools/exp_fused.py > gpurun_out/exp_default.json 2> gpurun_out/exp_default.err
TDSSF_LIB=$PWD/3d-semantic-segmentation-fusion_b200/build/variants/TDSSF_PROFILE/libtdssf.so python tools/exp_fused.py --iters 5 > gpurun_out/exp_profile.json 2> gpurun_out/exp_profile.err
cat gpurun_out/exp_default.json; tail -30 gpurun_out/exp_profile.err
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
ls -la gpurun_out
