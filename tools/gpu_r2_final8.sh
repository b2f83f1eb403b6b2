#!/bin/bash
# final 8-GPU line of the round with the single-GPU line of the same box
mkdir -p gpurun_out; O=gpurun_out
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29721 \
  bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_n8f.json 2> $O/bench_n8f.err; echo "n8 rc=$?"
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_n1f.json 2> $O/bench_n1f.err; echo "n1 rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/bench_n8f.json", "gpurun_out/bench_n1f.json"):
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, "%.4g" % d["value"], "e2e %.4g" % d["e2e"]["value"], "train %.4g %.3f ms" % (d["train"]["value"], d["train"]["ms_per_step"]),
          "eval %.4g" % d["eval"]["value"])
PY
