#!/bin/bash
# One box visit for the round's records: parity tests, bench (own arm + reference arm), per-row timings, training-step kernel
# times, and the ncu launch list of the bench command (after the plain run of the same command exited 0).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 50 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 900 python tools/bench_rows.py > gpurun_out/rows.jsonl 2> gpurun_out/rows.err; echo "rows rc=$?"
python tools/kineto_probe.py > gpurun_out/kineto_train.txt 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
tail -c 400 gpurun_out/bench.json; echo; tail -c 300 gpurun_out/bench_ref.json
