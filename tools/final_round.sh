#!/bin/bash
# End-of-round evidence on one B200 (run under gpurun): full GPU test suite, smoke, the default bench line, the reference
# arm, the BF16_TC line and the launch list of one FP32_REFINE step. Outputs under gpurun_out/final/.
cd "$(dirname "$0")/.."
o=gpurun_out/final
mkdir -p $o
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $o/gpu.txt
timeout 1800 python -m pytest tests -m gpu -q 2>&1 | tail -15 > $o/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $o/smoke.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > $o/bench_reference.json 2> $o/bench_reference.err
python bench.py > $o/bench_default.json 2> $o/bench_default.err
python bench.py --mode bf16 --no-cpu-baseline > $o/bench_bf16.json 2> $o/bench_bf16.err
python tools/refine_step.py > $o/refine_step_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $o/launches.csv python tools/refine_step.py > $o/refine_step_ncu.log 2>&1
python tools/launch_step.py $o/launches.csv > $o/launches_step.txt 2>&1
echo done
