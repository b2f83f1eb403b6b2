#!/bin/bash
# round 2 multi-GPU visit: bash tools/gpu_r2_multi.sh N   (N = 2, 4 or 8 GPUs of one box)
# collectives tests, oracle-per-shard check at world N, the bench line at N ranks (and N=1 for the ratio), L2 ingest micro-benchmark
N=${1:-2}
mkdir -p gpurun_out; O=gpurun_out
P=$((29500 + N))
timeout 600 python -m pytest tests/test_dist_gpu.py tests/test_multi_device.py -m gpu -x -q > $O/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_multi.log
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P \
  bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench n$N rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((P + 10)) \
  bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $O/bench_ref_n$N.json 2> $O/bench_ref_n$N.err; echo "bench ref n$N rc=$?"
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench n1 rc=$?"
python - <<'EOF'
import json, glob
for f in sorted(glob.glob("gpurun_out/bench_n*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "value %.3g" % d["value"], "e2e %.3g" % d["e2e"]["value"], "train %.3g" % d["train"]["value"],
              "eval %.3g" % d["eval"]["value"], "h2d/rank %.1f GB/s" % d["e2e"]["h2d_copy_only_gbs_per_rank"])
    except Exception as e:
        print(f, "unreadable:", repr(e)[:100])
EOF
tail -5 $O/bench_n$N.err
