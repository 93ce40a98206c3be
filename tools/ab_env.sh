#!/bin/bash
# Same-box A/B of a kernel variant selected by an environment variable (box-to-box variance is larger than most
# effects, so both arms must run inside ONE gpurun call):  tools/ab_env.sh ALINK_TC_STAGED [modes...]
# Runs the oracle sanity check and the role-delay stress test with the default (variant on), then alternates
# bench runs with VAR=0 / VAR=1.
cd "$(dirname "$0")/.."
var=${1:-ALINK_TC_STAGED}; shift
modes=${*:-bf16 refine}
o=gpurun_out/ab_$var
mkdir -p $o
for n in 1000 40000 300001; do
  timeout 300 python tools/tc_sanity.py $n >> $o/sanity.log 2>&1; echo "sanity $n rc=$?" | tee -a $o/rc.log
done
tail -3 $o/sanity.log
timeout 900 python tools/stress_roles.py > $o/stress.log 2>&1; echo "stress rc=$?" | tee -a $o/rc.log
tail -3 $o/stress.log
for rep in 1 2; do
  for d in 0 1; do
    for m in $modes; do
      f=$o/bench_${m}_${d}_$rep
      env $var=$d timeout 300 python bench.py --mode $m --no-mode-compare --no-cpu-baseline --no-clock-sampling \
        --steps 20 --warmup 5 > $f.json 2> $f.err
      python -c "
import json,sys
try:
    j=json.loads(open('$f.json').read().strip().splitlines()[-1])
    print('$var=$d mode=$m rep=$rep ms=%.3f score_ms=%.3f frac=%.3f' % (j['ms_per_step'], j['stage_ms']['ms_score'], j['roofline']['frac']))
except Exception as e:
    print('$var=$d mode=$m FAILED', e)
"
    done
  done
done 2>&1 | tee $o/ab.log
