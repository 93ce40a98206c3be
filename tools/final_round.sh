#!/bin/bash
# End-of-round evidence run on one GPU: GPU tests, smoke, the bench line (refine and bf16), the other configs,
# and the ncu launch list of the bench command. Outputs under gpurun_out/final/.
cd "$(dirname "$0")/.."
o=gpurun_out/final
mkdir -p $o
timeout 1200 python -m pytest tests -x -q -m gpu > $o/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $o/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $o/smoke.log 2>&1; tail -1 $o/smoke.log
timeout 600 python bench.py > $o/bench_refine.json 2> $o/bench_refine.err; tail -c 400 $o/bench_refine.json; echo
timeout 600 python bench.py --mode bf16 --no-cpu-baseline > $o/bench_bf16.json 2> $o/bench_bf16.err; tail -c 300 $o/bench_bf16.json; echo
for c in k1 5 4; do timeout 600 python tools/run_configs.py $c >> $o/configs.jsonl 2>> $o/configs.err; done
cut -c1-300 $o/configs.jsonl
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-mode-compare --no-clock-sampling > $o/ncu_launch.log 2>&1
echo "launch list rc=$?"
