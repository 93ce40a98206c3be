#!/bin/bash
# 8-GPU evidence on one box: weak-scaling line, north-star target shape (10 M pairs) in both modes, config 3.
cd "$(dirname "$0")/.."
o=gpurun_out/scale8
mkdir -p $o
run() {  # name, extra flags
  name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu-baseline --no-mode-compare "$@" > $o/$name.json 2> $o/$name.err
  python -c "
import json
try:
    j=json.loads(open('$o/$name.json').read().strip().splitlines()[-1]); print('$name', 'ms=%.3f value=%.4g frac=%.3f' % (j['ms_per_step'], j['value'], j['roofline']['frac']), j.get('clocks'))
except Exception as e: print('$name FAILED', e)
"
}
run scale_n8
run target_bf16_n8 --mode bf16 --pairs 1250000
run target_refine_n8 --pairs 1250000
[ -n "$SKIP_CFG3" ] || run cfg3_n8 --config 3 --mode bf16
