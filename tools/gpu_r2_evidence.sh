#!/bin/bash
# round 2 evidence visit (1 GPU): parity suite, bench line + reference arm, ncu launch list of the bench command, ncu --set full
# captures (fused inference config2 / config4, backward kernels), compute-sanitizer runs, training-step kernel times
mkdir -p gpurun_out; O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -3 $O/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
timeout 300 python bench.py --workload config4 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_config4.json 2> $O/bench_config4.err; echo "config4 rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 300 python tools/kineto_probe.py > $O/kineto_train.txt 2>&1
timeout 300 python tools/bwd_probe.py > $O/bwd_probe.txt 2>&1
timeout 900 python tools/bench_rows.py > $O/rows.jsonl 2> $O/rows.err; echo "rows rc=$?"
# launch list of the bench command (after its plain run above exited 0)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/ncu_launches.log 2>&1; echo "ncu list rc=$?"
# --set full: the fused kernel of the bench command (config2), config4, bf16 tokens; then the backward kernels
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_mlp -s 6 -c 1 -o $O/prof_fused python bench.py --steps 3 --warmup 3 --no-cpu-baseline --headline-only > $O/ncu_fused.log 2>&1; echo "ncu fused rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_mlp -s 4 -c 1 -o $O/prof_fused_config4 python bench.py --workload config4 --steps 3 --warmup 3 --no-cpu-baseline --headline-only > $O/ncu_fused4.log 2>&1; echo "ncu config4 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"dgrad_chain|wgrad_kernel|wgrad_reduce|dw3_db3" -s 8 -c 4 -o $O/prof_bwd python tools/bwd_probe.py > $O/ncu_bwd.log 2>&1; echo "ncu bwd rc=$?"
# sanitizer
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python tools/san_probe.py > $O/san_memcheck.txt 2>&1; echo "memcheck rc=$?"; tail -3 $O/san_memcheck.txt
timeout 1200 compute-sanitizer --tool racecheck --print-limit 20 python tools/san_probe.py 600 > $O/san_racecheck.txt 2>&1; echo "racecheck rc=$?"; tail -3 $O/san_racecheck.txt
timeout 900 compute-sanitizer --tool synccheck --print-limit 20 python tools/san_probe.py 600 > $O/san_synccheck.txt 2>&1; echo "synccheck rc=$?"; tail -3 $O/san_synccheck.txt
ls -la $O/*.ncu-rep
