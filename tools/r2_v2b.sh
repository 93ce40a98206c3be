set -x
cd /root/repo
timeout 600 python tools/probe_refine2.py 1000000 > gpurun_out/r2_probe_1m.jsonl 2> gpurun_out/r2_probe_1m.err
echo "rc=$?" >> gpurun_out/r2_probe_1m.err
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_gputest_v2b.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_v2b.json 2> gpurun_out/r2_bench_v2b.err
python tools/refine_step.py > gpurun_out/refine_step_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_v2b.csv python tools/refine_step.py > gpurun_out/refine_step_ncu.log 2>&1
echo done
