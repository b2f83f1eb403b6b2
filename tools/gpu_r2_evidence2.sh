#!/bin/bash
# round 2, second evidence visit (1 GPU) after the bias-MMA / 256-bit logits change: whole parity suite, bench line + reference arm,
# ncu launch list of the bench command, ncu --set full of the fused kernel (config 2 and config 4), training-step kernel times
mkdir -p gpurun_out; O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -3 $O/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"
timeout 300 python tools/kineto_probe.py > $O/kineto_train.txt 2>&1
timeout 900 python tools/bench_rows.py > $O/rows.jsonl 2> $O/rows.err; echo "rows rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_mlp -s 6 -c 1 -o $O/prof_fused python bench.py --steps 3 --warmup 3 --no-cpu-baseline --headline-only > $O/ncu_fused.log 2>&1; echo "ncu fused rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_mlp -s 4 -c 1 -o $O/prof_fused_config4 python bench.py --workload config4 --steps 3 --warmup 3 --no-cpu-baseline --headline-only > $O/ncu_fused4.log 2>&1; echo "ncu config4 rc=$?"
ls -la $O/*.ncu-rep
