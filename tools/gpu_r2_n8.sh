#!/bin/bash
# round 2, 8-GPU visit: the multi-rank collectives test at world 4 / 8, the bench line at 8 and 4 ranks, the reference arm under torchrun
N=${1:-8}
mkdir -p gpurun_out; O=gpurun_out
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q > $O/pytest_dist_n$N.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_dist_n$N.log
for W in $N 4; do
  [ $W -le $N ] || continue
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port $((29600 + W)) \
    bench.py --gpus $W --steps 20 --warmup 5 > $O/bench_n$W.json 2> $O/bench_n$W.err; echo "bench n$W rc=$?"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29650 \
  bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $O/bench_ref_n$N.json 2> $O/bench_ref_n$N.err; echo "bench ref n$N rc=$?"
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench n1 rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/bench_n*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "value %.3g" % d["value"], "e2e %.3g" % d["e2e"]["value"], "train %.3g (%.3f ms)" % (d["train"]["value"], d["train"]["ms_per_step"]),
              "eval %.3g" % d["eval"]["value"], "h2d/rank %.1f GB/s" % d["e2e"]["h2d_copy_only_gbs_per_rank"])
    except Exception as e:
        print(f, "unreadable:", repr(e)[:100])
PY
