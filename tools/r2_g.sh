cd /root/repo
timeout 900 python -m pytest tests/test_gpu_refine2.py tests/test_gpu_bf16.py tests/test_gpu_fullsize.py tests/test_gpu_ref_pinned.py -q -x 2>&1 | tail -8 > gpurun_out/r2_gputest_g.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_g.json 2> gpurun_out/r2_bench_g.err
python tools/refine_step.py > gpurun_out/refine_step_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_g.csv python tools/refine_step.py > gpurun_out/refine_step_ncu.log 2>&1
echo done
