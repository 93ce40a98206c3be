set -x
cd /root/repo
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2_gputest_start.log
tools/ffma_peak > gpurun_out/r2_ffma_peak.json 2>&1
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_start.json 2> gpurun_out/r2_bench_start.err
python tools/hbm_stages.py > gpurun_out/hbm_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,l1tex__t_bytes.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,launch__grid_size,launch__registers_per_thread \
  --clock-control none -k regex:'utility_kernel|radix_pass|radix_select_small|compact_kernel|topk_small|bitonic|refine_|unpack_keys|select_init|head_kernel|ffma_layer' \
  --csv --page raw --log-file gpurun_out/r2_ncu_hbm_stages_raw.csv python tools/hbm_stages.py > gpurun_out/hbm_ncu.log 2>&1
echo done
