cd /root/repo
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -12 > gpurun_out/r2_gputest_j.log
python bench.py --mode bf16 --no-cpu-baseline > gpurun_out/r2_bench_j_bf16.json 2> gpurun_out/r2_bench_j_bf16.err
ALINK_TOPK_COOP=0 python bench.py --mode bf16 --no-cpu-baseline > gpurun_out/r2_bench_j_bf16_nocoop.json 2>> gpurun_out/r2_bench_j_bf16.err
echo done
