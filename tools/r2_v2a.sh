set -x
cd /root/repo
timeout 300 python tools/probe_refine2.py 200000 > gpurun_out/r2_probe_small.jsonl 2> gpurun_out/r2_probe_small.err
echo "rc=$?" >> gpurun_out/r2_probe_small.err
timeout 600 python tools/probe_refine2.py 1000000 > gpurun_out/r2_probe_1m.jsonl 2> gpurun_out/r2_probe_1m.err
echo "rc=$?" >> gpurun_out/r2_probe_1m.err
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_gputest_v2a.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_v2a.json 2> gpurun_out/r2_bench_v2a.err
echo done
