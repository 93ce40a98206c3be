cd /root/repo
timeout 1500 python -m pytest tests -m gpu -q -s 2>&1 | tail -60 > gpurun_out/r2_gputest_c.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_c.json 2> gpurun_out/r2_bench_c.err
echo done
