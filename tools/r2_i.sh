cd /root/repo
timeout 900 python -m pytest tests/test_gpu_refine2.py tests/test_gpu_bf16.py tests/test_gpu_fullsize.py tests/test_gpu_ref_pinned.py tests/test_gpu_allpairs.py -q -x 2>&1 | tail -8 > gpurun_out/r2_gputest_i.log
timeout 300 python tools/probe_refine2.py 300000 > gpurun_out/r2_probe_300k.jsonl 2> gpurun_out/r2_probe_300k.err
timeout 600 python tools/probe_refine2.py 1000000 > gpurun_out/r2_probe_1m.jsonl 2> gpurun_out/r2_probe_1m.err
echo done
