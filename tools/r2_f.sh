cd /root/repo
timeout 600 python -m pytest tests/test_gpu_a2link.py tests/test_gpu_refine2.py -q -x 2>&1 | tail -8 > gpurun_out/r2_gputest_f.log
python bench.py --config 4 --steps 3 --warmup 3 > gpurun_out/r2_bench_cfg4.json 2> gpurun_out/r2_bench_cfg4.err
python bench.py --config 5 --mode refine --steps 2 --warmup 3 > gpurun_out/r2_bench_cfg5_refine.json 2> gpurun_out/r2_bench_cfg5_refine.err
echo done
