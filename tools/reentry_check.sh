#!/bin/bash
# Short evidence run after a container re-entry: GPU tests, smoke, both bench arms, the launch list.
cd "$(dirname "$0")/.."
o=gpurun_out/reentry
mkdir -p $o
timeout 600 python -m pytest tests -x -q -m gpu > $o/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $o/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $o/smoke.log 2>&1; tail -1 $o/smoke.log
timeout 400 python bench.py --impl reference > $o/bench_reference.json 2> $o/bench_reference.err; tail -c 300 $o/bench_reference.json; echo
timeout 400 python bench.py > $o/bench_refine.json 2> $o/bench_refine.err; tail -c 600 $o/bench_refine.json; echo
timeout 300 python bench.py --mode bf16 --no-cpu-baseline > $o/bench_bf16.json 2> $o/bench_bf16.err; tail -c 300 $o/bench_bf16.json; echo
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-mode-compare --no-clock-sampling > $o/ncu_launch.log 2>&1
echo "launch list rc=$?"
