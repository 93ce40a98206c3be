cd /root/repo
timeout 900 python -m pytest tests/test_gpu_bf16.py tests/test_gpu_refine2.py tests/test_gpu_parity.py tests/test_gpu_stress.py -q -x 2>&1 | tail -12 > gpurun_out/r2_gputest_h.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_atm1.json 2> gpurun_out/r2_bench_atm1.err
ALINK_TC_ATM=0 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_atm0.json 2> gpurun_out/r2_bench_atm0.err
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --mode bf16 > gpurun_out/r2_bench_atm1_bf16.json 2>> gpurun_out/r2_bench_atm1.err
ALINK_TC_ATM=0 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --mode bf16 > gpurun_out/r2_bench_atm0_bf16.json 2>> gpurun_out/r2_bench_atm0.err
echo done
