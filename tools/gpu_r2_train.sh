#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "cels or loops or dropin or trajectory" 2>&1 | tail -4
timeout 600 python tools/kineto_probe.py 2>&1 | grep -v -i warn | head -40
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2> gpurun_out/bench.err | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value', d['value'], 'frac', d['roofline']['frac'])
print('train', json.dumps(d['train'])[:900])
print('eval', d['eval']['value'], d['eval']['ms_per_batch'])"
