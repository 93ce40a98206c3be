#!/bin/bash
# Same-box A/B of the current library against a baseline build (default build/lib_head.so, built from a git
# archive of HEAD): sanity + stress on the current one, then alternating bench runs.  tools/ab_lib.sh [modes...]
cd "$(dirname "$0")/.."
base=${ALINK_AB_BASE:-$PWD/build/lib_head.so}
modes=${*:-bf16 refine}
o=gpurun_out/ab_lib
mkdir -p $o
for n in 1000 40000 300001; do
  timeout 300 python tools/tc_sanity.py $n >> $o/sanity.log 2>&1; echo "sanity $n rc=$?" | tee -a $o/rc.log
done
tail -3 $o/sanity.log
timeout 900 python tools/stress_roles.py > $o/stress.log 2>&1; echo "stress rc=$?" | tee -a $o/rc.log
tail -2 $o/stress.log
for rep in 1 2; do
  for arm in base new; do
    for m in $modes; do
      f=$o/bench_${m}_${arm}_$rep
      if [ $arm = base ]; then export ALINK_LIB_PATH=$base; else unset ALINK_LIB_PATH; fi
      timeout 300 python bench.py --mode $m --no-mode-compare --no-cpu-baseline --no-clock-sampling \
        --steps 20 --warmup 5 > $f.json 2> $f.err
      python -c "
import json,sys
try:
    j=json.loads(open('$f.json').read().strip().splitlines()[-1])
    print('$arm mode=$m rep=$rep ms=%.3f score_ms=%.3f frac=%.3f' % (j['ms_per_step'], j['stage_ms']['ms_score'], j['roofline']['frac']))
except Exception as e:
    print('$arm mode=$m FAILED', e)
"
    done
  done
done 2>&1 | tee $o/ab.log
